import os, sys, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import cases
from core_b200.interface import YafrayInterface
emu = len(sys.argv) > 1
if emu:
    import refapi; mk = refapi.EmuInterface
else:
    mk = YafrayInterface
fn, kw = cases.RAY_CASES["rays_cornell"]
yi = mk(); fn(yi, **kw); yi.updateScene()
rays = cases.ray_batch("rays_cornell", n=4096)
print(rays[0])
print(yi.traceBatch(rays[:1],0))
