"""Generate the golden fixtures of tests/golden/ from the UNMODIFIED reference compiled into oracle/_ref
(run here, where /root/reference exists:  python tests/golden/make_golden.py).  The reference ships no golden
vectors of its own, so these pin the oracle port, the emulator and the CUDA path:
  kat.npz    : QMC / hash / fast-math known answers from the reference's inline headers (yref_* hooks)
  rays_*.npz : triKdTree_t::Intersect / scene_t::isShadowed results for seeded ray batches
  img_*.npz  : float RGBA frames rendered with threads=1 (bit-stable)
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cases  # noqa: E402
import refapi  # noqa: E402


def kat(lib):
    L = lib.lib
    L.yref_scr_halton.restype = C.c_double; L.yref_scr_halton.argtypes = [C.c_int, C.c_uint]
    for f in ("yref_ri_vdc", "yref_ri_s", "yref_ri_lp"):
        getattr(L, f).restype = C.c_float; getattr(L, f).argtypes = [C.c_uint, C.c_uint]
    L.yref_fnv.restype = C.c_uint; L.yref_fnv.argtypes = [C.c_uint]
    L.yref_halton_seq.argtypes = [C.c_int, C.c_uint, C.c_int, C.c_void_p]
    for f in ("yref_fsin", "yref_fcos", "yref_fexp", "yref_flog", "yref_fsqrt"):
        getattr(L, f).restype = C.c_float; getattr(L, f).argtypes = [C.c_float]
    L.yref_fpow.restype = C.c_float; L.yref_fpow.argtypes = [C.c_float, C.c_float]
    rng = np.random.Generator(np.random.PCG64(11))
    n_idx = np.concatenate([np.arange(0, 64), rng.integers(0, 2 ** 32, 192, dtype=np.uint64)]).astype(np.uint32)
    dims = np.arange(1, 50)   # dim 0 has base 1: the reference loop never terminates for n > 0
    scr = np.array([[L.yref_scr_halton(int(d), int(n)) for n in n_idx] for d in dims])
    r_idx = rng.integers(0, 2 ** 32, len(n_idx), dtype=np.uint64).astype(np.uint32)
    vdc = np.array([L.yref_ri_vdc(int(n), int(r)) for n, r in zip(n_idx, r_idx)], np.float32)
    ris = np.array([L.yref_ri_s(int(n), int(r)) for n, r in zip(n_idx, r_idx)], np.float32)
    rlp = np.array([L.yref_ri_lp(int(n), int(r)) for n, r in zip(n_idx, r_idx)], np.float32)
    fnv = np.array([L.yref_fnv(int(n)) for n in n_idx], np.uint32)
    starts = np.array([0, 1, 5, 63, 1000, 123456789, 4294967295], np.uint32)
    hal = np.zeros((3, len(starts), 16), np.float32)
    for bi, base in enumerate((2, 3, 5)):
        for si, st in enumerate(starts):
            buf = (C.c_float * 16)()
            L.yref_halton_seq(base, int(st), 16, buf)
            hal[bi, si] = np.frombuffer(buf, np.float32)
    x = np.concatenate([np.linspace(-10, 10, 201), rng.uniform(-100, 100, 55)]).astype(np.float32)
    fsin = np.array([L.yref_fsin(float(v)) for v in x], np.float32)
    fcos = np.array([L.yref_fcos(float(v)) for v in x], np.float32)
    pa = rng.uniform(1e-3, 10, 256).astype(np.float32); pb = rng.uniform(-4, 60, 256).astype(np.float32)
    fpow = np.array([L.yref_fpow(float(a), float(b)) for a, b in zip(pa, pb)], np.float32)
    np.savez_compressed(os.path.join(HERE, "kat.npz"), n_idx=n_idx, r_idx=r_idx, scr=scr, vdc=vdc, ris=ris, rlp=rlp, fnv=fnv,
                        hal_starts=starts, hal=hal, trig_x=x, fsin=fsin, fcos=fcos, pow_a=pa, pow_b=pb, fpow=fpow)


def rays(only=None):
    for name, (fn, kw) in cases.RAY_CASES.items():
        if only and name not in only:
            continue
        yi = refapi.RefInterface(); yi.setVerbosityMute()
        fn(yi, **kw)
        yi.updateScene()
        r = cases.ray_batch(name)
        closest = yi.traceBatch(r, 0)
        shadow = yi.traceBatch(r, 1)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), rays=r, closest=closest, shadow=shadow)
        print(name, "hits", int((closest["prim"] >= 0).sum()), "shadowed", int(shadow.sum()))


def depth(only=None):
    for name, (fn, kw) in cases.DEPTH_CASES.items():
        if only and name not in only:
            continue
        yi = refapi.RefInterface(); yi.setVerbosityMute()
        fn(yi, threads=1, **kw)
        img = yi.render(kw["width"], kw["height"])
        z = yi.getDepth(kw["width"], kw["height"])
        np.savez_compressed(os.path.join(HERE, "depth_" + name + ".npz"), img=img, z=z)
        print(name, img.mean(axis=(0, 1)), z.min(), z.max())


def images(only=None):
    for name, (fn, kw) in cases.IMAGE_CASES.items():
        if only and name not in only:
            continue
        yi = refapi.RefInterface(); yi.setVerbosityMute()
        fn(yi, threads=1, **kw)
        img = yi.render(kw["width"], kw["height"])
        np.savez_compressed(os.path.join(HERE, "img_" + name + ".npz"), img=img)
        print(name, img.mean(axis=(0, 1)))


if __name__ == "__main__":
    assert refapi.have_ref(), "build oracle/_ref first (make -C oracle ref)"
    only = set(sys.argv[1:])      # optional: regenerate just the named cases (fixtures are deterministic)
    if not only:
        kat(refapi.ref_lib())
    rays(only)
    images(only)
    depth(only)
