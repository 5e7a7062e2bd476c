import sys, os
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo/splat-trainer_b200"); sys.path.insert(0, "/root/repo")
import numpy as np, torch
import test_gpu_raster as t
from gsb200 import ops
from gpu_util import cu, npy
for channels in (1,3,4):
  for (W,H),n in [((96,64),600),((101,75),3000),((320,240),20000),((640,480),100000)]:
    s = t._setup(W,H,n,seed=30,channels=channels)
    image,fT,nc,vis,med = ops.raster_fwd(s["lcfg"], s["g2d"], s["feat"], s["depth"], s["o2p"], s["ranges"], W, H, visibility=True, median=True)
    r = s["orc"].raster_fwd(s["cfg"], npy(s["g2d"]), npy(s["feat"]), npy(s["depth"]), npy(s["o2p"])[:s["total"]], npy(s["ranges"]), W, H, visibility=True, median=True)
    image,fT,nc,vis,med = map(npy,(image,fT,nc,vis,med))
    print(channels,W,H,n,"img",np.abs(image-r[0]).max(),"fT",np.abs(fT-r[1]).max(),"nc!=",(nc!=r[2]).sum(),"vis",np.abs(vis-r[3]).max()/max(r[3].max(),1),"med!=",(med!=r[4]).sum())
