"""Dev helper: the Fast-Hessian scale space of a frame through the production kernels vs the CPU oracle, layer by layer
(positions of any differing responses), then one detect."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from boofcv_b200 import stages  # noqa: E402
from boofcv_b200.synth import blob_image  # noqa: E402
from oracle import bsoracle as orc  # noqa: E402  (dev tool, not the product path)

w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (400, 300)
img = blob_image(w, h, 234)
ii = stages.integral(img)
for L in stages.fh_intensity(ii):
    o = orc.hessian(ii, L["skip"], L["size"])
    g = L["intensity"]
    bad = np.argwhere(g.view(np.uint32) != o.view(np.uint32))
    print("size %3d skip %d  %4dx%-4d  differing %d" % (L["size"], L["skip"], g.shape[1], g.shape[0], len(bad)),
          "" if len(bad) == 0 else "first (y,x): %s  x range %d..%d  y range %d..%d" % (bad[:5].tolist(), bad[:, 1].min(), bad[:, 1].max(),
                                                                                     bad[:, 0].min(), bad[:, 0].max()))
    for y, x in bad[:3]:
        print("    gpu %r oracle %r" % (g[y, x], o[y, x]))
d = stages.fh_detect(ii)
print("points", len(d["x"]))
