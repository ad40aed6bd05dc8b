"""GPU suite, part 2 (-m gpu): the CUDA path against the LIVE reference (oracle/_ref = the unmodified reference compiled by
oracle/Makefile; it travels to the GPU box) at the geometry sizes and sample counts BASELINE.json names, with north_star's bars:
primitive ids exact except t-ties (< 1e-5 of the rays), t and barycentrics to 1e-5 relative, shadow flags equal except on
ties, frames to 1e-5 mean relative error / 100 dB PSNR (north_star asks for 1 % / 40 dB).

Two builds of the reference are used (oracle/Makefile).  `release` carries the reference's own flags, -O3 -ffast-math: GCC
reassociates and approximates, and the release build differs FROM ITS OWN strict-IEEE build on about 1e-4 of the camera
samples of a large frame - a light sample whose shadow ray passes within an ulp of a silhouette edge flips (measured in
this repo on the CPU: 18 of 230 400 pixels between the two builds of the reference, 0 between the strict build and these
kernels).  Identical rays give identical hits (first test), so frames are held to 1e-5 / 85 dB against
`strict` (measured on B200: 1.5e-6 / 92.6 dB on the C2 crop, 9e-8 / 111 dB on the terrain; what is left are libm vs CUDA
transcendental functions in the samplers) and to the release build's own noise floor (2e-4 rel, 50 dB, < 5 % of the pixels
off by > 1e-4; measured 6.1e-5 / 71.6 dB - exactly what the two builds of the reference show between themselves).

Sizes: C2 at its full 1 M triangles / 64 spp and C4 at its full 256 spp / 8 bounces always run.  The reference's
single-threaded kd-tree build (kdtree.cc:75-156) takes minutes on the 10 M-triangle soup and the 50 M-triangle terrain, so
C3 / C5 run at 1 M / 3.1 M triangles by default and at full size with YB200_FULL_PARITY=1 (tools/parity_full.py runs that
once per round and commits the result under profiles/)."""
import os

import numpy as np
import pytest

import cases
import refapi
from core_b200 import scenes
from core_b200.interface import YafrayInterface

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not refapi.have_ref(), reason="oracle/_ref not present")]
FULL = os.environ.get("YB200_FULL_PARITY", "0") == "1"
THREADS = os.cpu_count() or 8


def _frames(builder, **kw):
    """(release-build frame, strict-build frame or None, CUDA frame) of the same scene"""
    a = refapi.render_in_subprocess("release", builder.__name__, threads=THREADS, **kw)
    s = refapi.render_in_subprocess("strict", builder.__name__, threads=THREADS, **kw) if refapi.have_ref_strict() else None
    yi = YafrayInterface()
    builder(yi, **kw)
    return a, s, yi.render(kw["width"], kw["height"])


def _check_frames(a, s, b, tag, strict_bar=(1e-5, 85.0), release_bar=(2e-4, 50.0, 0.05)):
    def stats(r):
        rel, psnr, da = cases.compare_images(r, b)
        off = (np.abs(r[..., :3].astype(np.float64) - b[..., :3]) / np.maximum(np.abs(r[..., :3]), 1e-2)).max(axis=2) > 1e-4
        return rel, psnr, da, off.mean()
    ra = stats(a)
    msg = "%s: vs release build rel %.2e psnr %.1f dB, %.2f %% of the pixels off by > 1e-4" % (tag, ra[0], ra[1], 100 * ra[3])
    rs = None
    if s is not None:
        rs = stats(s)
        rel3, psnr3, _ = cases.compare_images(s, a)
        msg += "; vs strict build rel %.2e psnr %.1f dB, %.2f %% off (the two builds of the reference between themselves: %.2e, %.1f dB)" % (
            rs[0], rs[1], 100 * rs[3], rel3, psnr3)
    # alpha: exact against the strict build; against the release build as far as the reference's two builds agree with each other
    # (50 M-triangle terrain: one camera sample of the -ffast-math build slips between two sub-pixel triangles, alpha 6/7 in that pixel)
    alpha_bar = 1e-5
    if s is not None:
        alpha_bar += float(np.abs(s[..., 3].astype(np.float64) - a[..., 3]).max())
        msg += "; alpha: vs release %.3g, vs strict %.3g, release vs strict %.3g" % (ra[2], rs[2], alpha_bar - 1e-5)
    print(msg)
    assert np.isfinite(b).all()
    assert ra[0] <= release_bar[0] and ra[1] >= release_bar[1] and ra[2] <= alpha_bar and ra[3] <= release_bar[2], msg
    if rs is not None:
        assert rs[0] <= strict_bar[0] and rs[1] >= strict_bar[1] and rs[2] <= 1e-5, msg


def _pair(builder, ref_kw=None, **kw):
    yr = refapi.RefInterface(); yr.setVerbosityMute()
    builder(yr, threads=THREADS, **dict(kw, **(ref_kw or {})))
    yi = YafrayInterface()
    builder(yi, **kw)
    return yr, yi


def _surface_rays(rays, hits, rng, tmin=0.0005, far=0.5):
    """Rays leaving the surfaces the first batch hit: cosine-ish directions around the geometric side the ray came from -
    what the integrators actually trace (bounce and shadow rays), including origins that lie ON triangles."""
    h = hits["prim"] >= 0
    P = rays[h, 0:3] + rays[h, 4:7] * hits["t"][h, None]
    d = rng.normal(size=P.shape)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    r = np.zeros((len(P), 8), np.float32)
    r[:, 0:3] = P; r[:, 3] = tmin; r[:, 4:7] = d
    r[:, 7] = np.where(rng.uniform(size=len(P)) < far, -1.0, rng.uniform(0.05, 3.0, len(P)))
    return r


def _check_rays(yr, yi, rays, tie_frac=1e-5):
    want = yr.traceBatch(rays, 0)
    got = yi.traceBatch(rays, 0)
    bad, terr, berr, nties = cases.compare_hits(want, got)
    assert bad == 0, (bad, terr, berr, nties)
    assert nties <= max(1, int(tie_frac * len(rays))), nties
    ok = (want["prim"] >= 0) & (want["prim"] == got["prim"])
    for k in ("b1", "b2"):                                         # barycentrics: 1e-5 relative (to the unit range they live in)
        assert np.abs(want[k][ok] - got[k][ok]).max() <= 1e-5, k
    assert terr <= 1e-5
    s_want = yr.traceBatch(rays, 1)
    s_got = yi.traceBatch(rays, 1)
    diff = np.nonzero(s_want != s_got)[0]
    # a shadow flag may only differ where the any-hit interval ends exactly on a surface (t-tie with tmax)
    assert len(diff) <= max(1, int(tie_frac * len(rays))), len(diff)
    return int((want["prim"] >= 0).sum()), nties, len(diff)


def test_c2_rays_at_1m_triangles():
    """BASELINE configs[1] geometry: 1 M seeded rays (half from outside, half leaving surfaces) through
    triKdTree_t::Intersect / scene_t::isShadowed of the reference vs yb200_trace_batch."""
    kw = dict(n_tris=1_000_000, width=64, height=36, spp=1)
    yr, yi = _pair(scenes.icosphere_field, **kw)
    yr.updateScene(); yi.updateScene()
    rng = np.random.Generator(np.random.PCG64(41))
    first = cases.ray_batch("rays_field", n=500_000, seed=40)
    first[:, 0:2] *= 3.5                                            # spread over the whole field
    nh, nt, ns = _check_rays(yr, yi, first)
    assert nh > 100_000
    second = _surface_rays(first, yr.traceBatch(first, 0), rng)
    second = np.concatenate([second, _surface_rays(second, yr.traceBatch(second, 0), rng)])[:500_000]
    nh2, nt2, ns2 = _check_rays(yr, yi, second)
    assert nh2 > 50_000
    print("c2 rays: %d + %d hits, ties %d + %d, shadow flag diffs %d + %d" % (nh, nh2, nt, nt2, ns, ns2))


def test_c2_frame_crop_at_1m_triangles_64spp():
    """BASELINE configs[1] itself - 1 M triangles, pathtracing 4 bounces, 1920x1080 camera, 64 spp - on a cropped window
    (xstart / ystart, environment.cc:498-499): the benchmark's own pixels against the reference."""
    kw = dict(n_tris=1_000_000, width=160, height=96, spp=64, bounces=4, cam_res=(1920, 1080), xstart=880, ystart=560)
    a, s, b = _frames(scenes.icosphere_field, **kw)
    assert a[..., :3].std() > 0.05
    _check_frames(a, s, b, "c2 1M triangles 64 spp crop")


def test_c4_materials_at_256spp_8_bounces():
    """BASELINE configs[3] at its full sample count and depth (240x135): glossy / coated / glass, sun + point + area,
    pathtracing 8 bounces.  The reference's estimator has no clamping (W = |wi.N| / (0.99 pdf + 0.01) up to 100 per bounce,
    sun colPdf ~ 1e3), so single samples reach 1e5..1e7 and the frame carries fireflies on both sides (reference at 64 spp,
    240x135: max pixel 1.1e5, mean 2.7; at 256 spp: max 8.4e2) - that is where the 1080p mean of 1.4e4 comes from
    (DESIGN.md section 5).  The worst of them are chains of grazing mirror reflections: mirrorMat_t::sample returns
    refCol / |N.wi| with W = 1, a ray that leaves a smooth-shaded sphere almost tangentially re-hits the next facet after 1e-3
    and multiplies the throughput by ~16 per bounce (traced in the emulator: sample 111 of pixel (116, 78), x 16.7 five times).
    Whether such a chain starts is decided by the last bit of the hit point, so it differs between ANY two builds; identical
    rays give identical hits (60 000 grazing rays: reference kd-tree = this BVH = brute force).  A flipped firefly moves the
    global error measures by its own size, so the frame is compared pixel by pixel and with the fireflies clipped."""
    kw = dict(width=240, height=135, spp=256, bounces=8)
    a, s, b = _frames(scenes.materials_scene, **kw)
    assert np.isfinite(a).all() and np.isfinite(b).all()
    rows = []
    for name, r in (("release", a), ("strict", s)):
        if r is None:
            continue
        pr = (np.abs(r[..., :3].astype(np.float64) - b[..., :3]) / np.maximum(np.abs(r[..., :3]), 1e-2)).max(axis=2)
        rel, psnr, da = cases.compare_images(r, b)
        relc, psnrc, _ = cases.compare_images(np.minimum(r, 4.0), np.minimum(b, 4.0))      # fireflies clipped: robust global measure
        rows.append((name, rel, relc, psnrc, float(np.median(pr)), float((pr > 1e-3).mean()), da))
        print("c4 256 spp vs %s build: mean %.4f / %.4f, max %.1f / %.1f, rel %.2e (clipped at 4: %.2e, %.1f dB), median pixel rel %.1e, "
              "pixels off by > 1e-3: %.2f %%" % (name, r[..., :3].mean(), b[..., :3].mean(), r[..., :3].max(), b[..., :3].max(), rel, relc, psnrc,
                                                  float(np.median(pr)), 100 * float((pr > 1e-3).mean())))
    if s is not None:
        print("c4 256 spp, the two builds of the reference between themselves: rel %.2e, clipped %.2e" % (
            cases.compare_images(s, a)[0], cases.compare_images(np.minimum(s, 4.0), np.minimum(a, 4.0))[0]))
    for name, rel, relc, psnrc, med, off, da in rows:
        assert da <= 1e-5
        assert med <= 1e-5                                         # the bulk of the frame agrees to float noise
        assert off <= 0.01                                         # pixels holding a sample that took another path (measured 0.4 - 0.5 %)
        assert relc <= 2e-3 and psnrc >= 40.0                      # measured 8e-4 / 42 dB: north_star's 1 % / 40 dB with the fireflies clipped
        assert rel <= 5e-2                                         # unclipped: one flipped firefly sample moves the whole-frame mean by a per cent


def test_c3_soup_any_hit_and_ao_frame():
    """BASELINE configs[2]: random triangle soup - closest / any-hit rays incl. rays leaving surfaces with the AO interval,
    and the directlighting + AO frame on a 128x128 crop."""
    n = 10_000_000 if FULL else 1_000_000
    kw = dict(n_tris=n, width=128, height=128, spp=2, ao_samples=32, ao_distance=0.1, cam_res=(512, 512), xstart=192, ystart=192)
    yr, yi = _pair(scenes.triangle_soup, **kw)
    yr.updateScene(); yi.updateScene()
    rng = np.random.Generator(np.random.PCG64(43))
    first = cases.ray_batch("rays_soup", n=200_000, seed=42)
    nh, nt, ns = _check_rays(yr, yi, first, tie_frac=2e-5)
    ao = _surface_rays(first, yr.traceBatch(first, 0), rng, tmin=0.0005, far=0.5)
    ao[:, 7] = np.where(ao[:, 7] < 0, -1.0, 0.1)                    # AO_distance 0.1 and infinity
    _check_rays(yr, yi, ao, tie_frac=2e-5)
    del yr, yi
    a, s, b = _frames(scenes.triangle_soup, **kw)
    _check_frames(a, s, b, "c3 soup %d triangles AO crop" % n)


def test_c5_terrain_frame_crop():
    """BASELINE configs[4]: tessellated terrain, pathtracing 4 bounces, 3840x2160 camera, 128x128 crop."""
    grid = 5000 if FULL else 1250
    kw = dict(grid=grid, width=128, height=128, spp=4, bounces=4, cam_res=(3840, 2160), xstart=1856, ystart=1400)
    a, s, b = _frames(scenes.terrain, **kw)
    assert a[..., :3].std() > 0.01
    _check_frames(a, s, b, "c5 terrain %d triangles crop" % (2 * grid * grid))
