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


BSDF_MATERIALS = ["ground", "glossy", "aniso", "gdiff", "coated", "glass", "glass2", "mirror", "shiny"]    # creation order of scenes.materials_scene


def bsdf_queries(n_per=96, seed=21):
    """Seeded surface frames / directions / sample values for the BSDF known answers: record layout of yb200_selftest
    (include/yb200.h): {material index, N, Ng, wo, wi, s1, s2, flags}.  Directions cover both hemispheres; Ng = N for half of
    the records and a tilted geometric normal otherwise; flags cycle through BSDF_ALL and single-component masks."""
    rng = np.random.Generator(np.random.PCG64(seed))
    BSDF_SPECULAR, BSDF_GLOSSY, BSDF_DIFFUSE, BSDF_REFLECT, BSDF_TRANSMIT = 0x1, 0x2, 0x4, 0x10, 0x20
    ALL = 0x1 | 0x2 | 0x4 | 0x8 | 0x10 | 0x20 | 0x40            # BSDF_ALL (material.h:62)
    masks = [ALL, ALL, BSDF_DIFFUSE | BSDF_REFLECT, BSDF_GLOSSY | BSDF_REFLECT, BSDF_SPECULAR | BSDF_REFLECT | BSDF_TRANSMIT,
             BSDF_GLOSSY | BSDF_DIFFUSE | BSDF_REFLECT | BSDF_TRANSMIT]

    def unit(v):
        return v / np.linalg.norm(v, axis=-1, keepdims=True)
    recs = []
    for mi in range(len(BSDF_MATERIALS)):
        N = unit(rng.normal(size=(n_per, 3)))
        Ng = np.where(rng.uniform(size=(n_per, 1)) < 0.5, N, unit(N + 0.3 * rng.normal(size=(n_per, 3))))
        wo = unit(N * rng.uniform(0.05, 1.0, (n_per, 1)) + 0.7 * rng.normal(size=(n_per, 3)))
        wi = unit(N * rng.uniform(-0.3, 1.0, (n_per, 1)) + 0.7 * rng.normal(size=(n_per, 3)))
        r = np.zeros((n_per, 16), np.float32)
        r[:, 1:4] = N; r[:, 4:7] = Ng; r[:, 7:10] = wo; r[:, 10:13] = wi
        r[:, 13] = rng.uniform(0, 1, n_per); r[:, 14] = rng.uniform(0, 1, n_per)
        w = r.view(np.uint32)
        w[:, 0] = mi
        w[:, 15] = np.array([masks[k % len(masks)] for k in range(n_per)], np.uint32)
        recs.append(r)
    return np.concatenate(recs)


def bsdf_kat(lib):
    """material_t::initBSDF / eval / pdf / sample of the nine materials of scenes.materials_scene, called on the compiled reference."""
    from core_b200 import scenes
    yi = refapi.RefInterface(); yi.setVerbosityMute()
    scenes.materials_scene(yi, width=8, height=8, spp=1)
    q = bsdf_queries()
    out = np.zeros_like(q)
    L = lib.lib
    L.yref_bsdf_kat.restype = C.c_int
    L.yref_bsdf_kat.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_size_t, C.c_void_p]
    mi = q.view(np.uint32)[:, 0]
    for k, name in enumerate(BSDF_MATERIALS):
        sel = np.nonzero(mi == k)[0]
        qi = np.ascontiguousarray(q[sel]); oi = np.zeros_like(qi)
        assert L.yref_bsdf_kat(yi._c, name.encode(), qi.ctypes.data, len(qi), oi.ctypes.data) == 1
        out[sel] = oi
    np.savez_compressed(os.path.join(HERE, "bsdf_kat.npz"), q=q, out=out, materials=np.array(BSDF_MATERIALS))
    print("bsdf_kat", len(q), "records; non-black eval", int((out[:, :3].sum(1) > 0).sum()), "non-zero sample pdf", int((out[:, 11] > 0).sum()))


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
    if not only or "bsdf_kat" in only:
        bsdf_kat(refapi.ref_lib())
    rays(only)
    images(only)
    depth(only)
