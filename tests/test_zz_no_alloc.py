"""yb200_trace_batch_static (include/yb200.h): the allocation-free form of the ray-batch hook, whose scratch (ray-fetch
counter of the persistent traversal kernel + staging of host batches) is a file-scope `static __device__` array.  It has no
counterpart in the reference, so the oracle here is the existing entry point
yb200_trace_batch on the same rays: integer ids / flags and float t / barycentrics must be bit-identical (same kernel, same
per-ray traversal order).  Two calls in a row: the array keeps the first call's contents, the second must not see them.

Collected last on purpose (pytest -x would hide every test after a failure here)."""
import ctypes as C

import numpy as np
import pytest

import cases
from core_b200 import capi, scenes
from core_b200.interface import YafrayInterface

pytestmark = pytest.mark.gpu

MAX_HOST = 65536            # YB200_STATIC_MAX_RAYS_HOST


def _static(yi, rays_ptr, n, mode, out_ptr, on_device):
    return yi._l.fn("trace_batch_static")(yi._c, rays_ptr, n, mode, out_ptr, on_device)


def _out(n, mode):
    return np.zeros(n, dtype=[("prim", "<i4"), ("t", "<f4"), ("b1", "<f4"), ("b2", "<f4")]) if mode == 0 else np.zeros(n, np.int32)


def _scene():
    yi = YafrayInterface()
    scenes.icosphere_field(yi, n_tris=20000, subdiv=2, width=64, height=36, spp=1)
    yi.updateScene()
    return yi


def test_static_entry_matches_trace_batch_on_two_calls_in_a_row():
    yi = _scene()
    # two different batches of different sizes, the second smaller: stale rays / results / counter of the first call stay in
    # the scratch array behind it
    batches = [cases.ray_batch("rays_field", n=MAX_HOST, seed=11), cases.ray_batch("rays_field", n=5000, seed=12)]
    for mode in (0, 1):
        for rays in batches:
            want = yi.traceBatch(rays, mode)
            got = _out(len(rays), mode)
            assert _static(yi, rays.ctypes.data, len(rays), mode, got.ctypes.data, 0) == 1, yi.lastError()
            assert np.array_equal(got.view(np.uint8), want.view(np.uint8)), (mode, len(rays))
    assert (yi.traceBatch(batches[0], 0)["prim"] >= 0).sum() > 1000          # the comparison is not over misses only


def test_static_entry_device_pointers_and_limits():
    import torch
    yi = _scene()
    rays = cases.ray_batch("rays_field", n=200000, seed=13)
    d_r = torch.from_numpy(rays).cuda()
    for mode, width in ((0, 4), (1, 1)):
        want = yi.traceBatch(rays, mode)
        for _ in range(2):
            d_o = torch.full((len(rays), width), -7, dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            assert _static(yi, d_r.data_ptr(), len(rays), mode, d_o.data_ptr(), 1) == 1, yi.lastError()
            torch.cuda.synchronize()          # device-wide: covers the context's own (non-blocking) stream
            assert np.array_equal(d_o.cpu().numpy().view(np.uint8).reshape(-1), want.view(np.uint8).reshape(-1)), mode
    # beyond the limits: a clear error, nothing written
    big = np.zeros((MAX_HOST + 1, 8), np.float32)
    out = np.full(MAX_HOST + 1, 123, np.int32)
    assert _static(yi, big.ctypes.data, len(big), 1, out.ctypes.data, 0) == 0
    assert "YB200_STATIC_MAX_RAYS_HOST" in yi.lastError() and np.all(out == 123)
    assert _static(yi, d_r.data_ptr(), (1 << 30) + 1, 0, d_r.data_ptr(), 1) == 0 and "YB200_STATIC_MAX_RAYS_DEVICE" in yi.lastError()
    assert _static(yi, d_r.data_ptr(), 16, 2, d_r.data_ptr(), 1) == 0 and "mode" in yi.lastError()
    fresh = YafrayInterface()
    assert _static(fresh, rays.ctypes.data, 16, 0, out.ctypes.data, 0) == 0 and "update_scene" in fresh.lastError()
