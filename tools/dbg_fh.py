"""Dev helper: one Fast-Hessian detect of a 400x300 frame through the stage-level ABI (for compute-sanitizer / ncu runs)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from boofcv_b200 import stages  # noqa: E402
from boofcv_b200.synth import blob_image  # noqa: E402

img = blob_image(400, 300, 234)
ii = stages.integral(img)
d = stages.fh_detect(ii)
print("points", len(d["x"]))
