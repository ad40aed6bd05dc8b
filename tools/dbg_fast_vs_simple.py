import os, sys, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import cases
from core_b200 import scenes
from core_b200.interface import YafrayInterface
name = "rays_cornell"
fn, kw = cases.RAY_CASES[name]
yi = YafrayInterface(); fn(yi, **kw); yi.updateScene()
rays = cases.ray_batch(name, n=4096)
os.environ["YB200_SIMPLE_TRACE"]="1"; a = yi.traceBatch(rays[:32],0)
os.environ["YB200_SIMPLE_TRACE"]="0"; b = yi.traceBatch(rays[:32],0)
single = np.concatenate([yi.traceBatch(rays[i:i+1],0) for i in range(32)])
print("simple", a["prim"]); print("fast32", b["prim"]); print("fast1 ", single["prim"])
print("dirs sign of failing:", [(int(i), (rays[i,4:7]>=0).astype(int).tolist()) for i in np.nonzero(a["prim"]!=b["prim"])[0]])
print("dirs sign of ok hits:", [(int(i), (rays[i,4:7]>=0).astype(int).tolist()) for i in np.nonzero((a["prim"]==b["prim"])&(a["prim"]>=0))[0]])
