"""Several GPUs (-m gpu, skipped on a one-GPU box; run with `gpurun --gpus 2`):
  (1) yb200_set_devices: ONE context / one process drives all devices - the reference-facing render() call itself scales; the
      BVH is built on the leader and reaches the other devices by peer copy, films are summed on the leader;
  (2) one process per GPU under torchrun, films reduced with NCCL (what bench.py does): the reduced frame must equal the golden
      frame, with the library stream ordered after the reduce; the adaptive path exchanges the film after every pass."""
import os
import subprocess
import sys

import numpy as np
import pytest

import cases
from core_b200 import scenes
from core_b200.interface import YafrayInterface

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "tests", "golden")
pytestmark = pytest.mark.gpu


def _ndev():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


needs2 = pytest.mark.skipif(_ndev() < 2, reason="needs at least two GPUs")


@needs2
@pytest.mark.parametrize("name", ["c1_cornell_path", "c4_materials_path", "c2_field_adaptive", "c4_imagetex_direct", "c4_meshlight_direct"])
def test_set_devices_frame_equals_golden(name):
    """All devices of the box behind one context: same golden bars as the one-device frame (tiles of 16 px so every device owns some)."""
    fn, kw = cases.IMAGE_CASES[name]
    n = min(_ndev(), 8)
    yi = YafrayInterface()
    yi.setDevices(list(range(n)))
    fn(yi, tile_size=16, **kw)
    img = yi.render(kw["width"], kw["height"])
    ref = np.load(os.path.join(G, "img_" + name + ".npz"))["img"]
    rel, psnr, da = cases.compare_images(ref, img)
    bar = cases.GPU_IMAGE_BARS.get(name, cases.GPU_IMAGE_BAR)
    assert rel <= bar[0] and psnr >= bar[1] and da <= bar[2], (rel, psnr, da)
    st = yi.getStats()
    one = YafrayInterface(); fn(one, tile_size=16, **kw); one.render(kw["width"], kw["height"])
    s1 = one.getStats()
    assert st["rays_closest"] == s1["rays_closest"] and st["camera_samples"] == s1["camera_samples"]     # same work, split over devices
    # a second frame through the same group (buffers and helper scenes are reused)
    img2 = yi.render(kw["width"], kw["height"])
    assert np.allclose(img, img2, rtol=2e-5, atol=1e-6)


@needs2
def test_set_devices_depth_channel():
    name = "z_field_raw"
    fn, kw = cases.DEPTH_CASES[name]
    yi = YafrayInterface()
    yi.setDevices([0, 1])
    fn(yi, tile_size=16, **kw)
    img = yi.render(kw["width"], kw["height"])
    z = yi.getDepth(kw["width"], kw["height"])
    g = np.load(os.path.join(G, "depth_" + name + ".npz"))
    rel, psnr, da = cases.compare_images(g["img"], img)
    assert rel <= 1e-5 and psnr >= 100.0, (rel, psnr, da)
    assert np.allclose(z, g["z"], rtol=2e-5, atol=1e-6)


@needs2
def test_set_devices_rejects_bad_lists():
    yi = YafrayInterface()
    with pytest.raises(RuntimeError, match="twice"):
        yi.setDevices([0, 0])
    with pytest.raises(RuntimeError, match="no such"):
        yi.setDevices([0, 99])
    yi.setDevices([1, 0])                       # leader = device 1
    scenes.cornell_box(yi, 48, 32, spp=2, light_samples=2, tile_size=16)
    a = yi.render(48, 32)
    yi.setDevices([1])                          # back to one device
    b = yi.render(48, 32)
    assert np.allclose(a, b, rtol=2e-5, atol=1e-6)


WORKER = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
ROOT = %(root)r
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases
from core_b200.interface import YafrayInterface
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
out = sys.argv[1]
res = {}
for name in ("c1_cornell_path", "c2_field_adaptive"):
    fn, kw = cases.IMAGE_CASES[name]
    W, H = kw["width"], kw["height"]
    yi = YafrayInterface(device=local)
    fn(yi, tile_size=16, **kw)
    yi.updateScene()
    ext = torch.cuda.ExternalStream(yi.cudaStream())
    film = torch.zeros(H, W, 5, device="cuda")
    torch.cuda.synchronize()
    passes = kw.get("passes", 1)
    if passes == 1:
        yi.renderFilm(film, rank, world)
        dist.reduce(film, 0)
        ext.wait_stream(torch.cuda.current_stream())       # the library's stream is non-blocking: order it after the NCCL reduce
        total = film
    else:
        total = None
        for p in range(passes):
            yi.renderFilmPass(film, total, rank, world, p)
            total = film.clone()
            dist.all_reduce(total)
            ext.wait_stream(torch.cuda.current_stream())
    if rank == 0:
        res[name] = yi.filmResolve(total, W, H)
    dist.barrier()
if rank == 0:
    np.savez(out, **res)
dist.destroy_process_group()
'''


@needs2
def test_nccl_reduced_frame_equals_golden(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(WORKER % dict(root=ROOT))
    out = str(tmp_path / "o.npz")
    n = 2
    port = 29500 + os.getpid() % 400
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
                        "--master-port", str(port), str(script), out], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-3000:]
    got = np.load(out)
    for name in ("c1_cornell_path", "c2_field_adaptive"):
        ref = np.load(os.path.join(G, "img_" + name + ".npz"))["img"]
        rel, psnr, da = cases.compare_images(ref, got[name])
        assert rel <= 1e-5 and psnr >= 100.0, (name, rel, psnr, da)
