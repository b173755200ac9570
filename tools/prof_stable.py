import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, boofcv_b200 as bs
from boofcv_b200.synth import blob_images_torch
fr=blob_images_torch(1920,1080,16,seed=1,device=torch.device("cuda"))
det=bs.FactoryDetectDescribe.surfStable(None,None,None,np.float32)
for _ in range(3):
    det.detectBatch(fr); det.totalFeatures()
