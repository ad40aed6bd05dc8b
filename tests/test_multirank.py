"""World-size-2 tile partition on CPU (gloo): each rank accumulates its round-robin tiles into a full-frame film,
the films are summed with a reduce to rank 0 (the NCCL reduce of the GPU path) and resolved; the result must equal
the single-rank frame up to float summation order."""
import os
import subprocess
import sys

import numpy as np
import pytest

import refapi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import refapi
from core_b200 import scenes
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%(port)d", rank=int(sys.argv[1]), world_size=2)
rank = dist.get_rank()
W, H = 80, 48
yi = refapi.EmuInterface()
scenes.cornell_box(yi, W, H, spp=2, light_samples=2, integrator="pathtracing", integ_params=dict(bounces=3, path_samples=1), tile_size=16)
film = torch.zeros(H, W, 5, dtype=torch.float32)
yi.renderFilm(film, rank, 2)
dist.reduce(film, 0)
if rank == 0:
    np.save(sys.argv[2], yi.filmResolve(film, W, H))
dist.barrier()
'''


@pytest.mark.skipif(not refapi.have_emu(), reason="host emulator could not be built")
def test_two_ranks_equal_one(tmp_path):
    from core_b200 import scenes
    port = 29000 + os.getpid() % 2000
    script = tmp_path / "w.py"
    script.write_text(WORKER % dict(root=ROOT, port=port))
    out = str(tmp_path / "o.npy")
    procs = [subprocess.Popen([sys.executable, str(script), str(r), out]) for r in range(2)]
    for p in procs:
        assert p.wait(timeout=300) == 0
    two = np.load(out)
    yi = refapi.EmuInterface()
    scenes.cornell_box(yi, 80, 48, spp=2, light_samples=2, integrator="pathtracing",
                       integ_params=dict(bounces=3, path_samples=1), tile_size=16)
    one = yi.render(80, 48)
    assert np.allclose(one, two, rtol=2e-5, atol=1e-6)


WORKER_ADAPTIVE = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import refapi, cases
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%(port)d", rank=int(sys.argv[1]), world_size=2)
rank = dist.get_rank()
fn, kw = cases.IMAGE_CASES["c2_field_adaptive"]
W, H = kw["width"], kw["height"]
yi = refapi.EmuInterface()
fn(yi, tile_size=16, **kw)
film = torch.zeros(H, W, 5, dtype=torch.float32)
film_all = None
for p in range(kw["passes"]):
    yi.renderFilmPass(film, film_all, rank, 2, p)     # pass p on this rank's tiles
    film_all = film.clone()
    dist.all_reduce(film_all)                          # nextPass needs the film of ALL ranks (NCCL all-reduce on GPUs)
if rank == 0:
    np.save(sys.argv[2], yi.filmResolve(film_all, W, H))
dist.barrier()
'''


@pytest.mark.skipif(not refapi.have_emu(), reason="host emulator could not be built")
def test_two_ranks_adaptive_passes_equal_golden(tmp_path):
    """Adaptive AA on two ranks: per-pass film exchange, result = the reference's single-process frame."""
    import cases
    port = 31000 + os.getpid() % 2000
    script = tmp_path / "w.py"
    script.write_text(WORKER_ADAPTIVE % dict(root=ROOT, port=port))
    out = str(tmp_path / "o.npy")
    procs = [subprocess.Popen([sys.executable, str(script), str(r), out]) for r in range(2)]
    for p in procs:
        assert p.wait(timeout=300) == 0
    two = np.load(out)
    ref = np.load(os.path.join(ROOT, "tests", "golden", "img_c2_field_adaptive.npz"))["img"]
    rel, psnr, da = cases.compare_images(ref, two)
    assert rel <= 1e-4 and psnr >= 80.0, (rel, psnr, da)
    # and yb200_render_film refuses to run adaptive passes blind on several ranks
    import torch
    fn, kw = cases.IMAGE_CASES["c2_field_adaptive"]
    yi = refapi.EmuInterface(); fn(yi, **kw)
    with pytest.raises(RuntimeError, match="render_film_pass"):
        yi.renderFilm(torch.zeros(kw["height"], kw["width"], 5), 0, 2)
