"""CPU suite: the kernel bodies (compiled for the host by tests/emu, -DYB_HOSTEMU) against the golden frames and
kd-tree ray results of the compiled reference.  Checks the host logic + every device function without a GPU.
The same assertions run against the real CUDA library in test_gpu_parity.py."""
import os

import numpy as np
import pytest

import cases
import refapi

G = os.path.join(os.path.dirname(__file__), "golden")
pytestmark = pytest.mark.skipif(not refapi.have_emu(), reason="host emulator could not be built")


@pytest.mark.parametrize("name", list(cases.IMAGE_CASES))
def test_image_parity(name):
    fn, kw = cases.IMAGE_CASES[name]
    yi = refapi.EmuInterface()
    fn(yi, **kw)
    img = yi.render(kw["width"], kw["height"])
    ref = np.load(os.path.join(G, "img_" + name + ".npz"))["img"]
    rel, psnr, da = cases.compare_images(ref, img)
    # north_star bar: <= 1 % mean relative error, PSNR >= 40 dB; measured here: ~1e-7 / > 120 dB
    bar = cases.EMU_IMAGE_BARS.get(name, cases.EMU_IMAGE_BAR)
    assert rel <= bar[0] and psnr >= bar[1] and da <= bar[2], (rel, psnr, da)


@pytest.mark.parametrize("name", list(cases.DEPTH_CASES))
def test_depth_channel_parity(name):
    fn, kw = cases.DEPTH_CASES[name]
    yi = refapi.EmuInterface()
    fn(yi, **kw)
    img = yi.render(kw["width"], kw["height"])
    z = yi.getDepth(kw["width"], kw["height"])
    g = np.load(os.path.join(G, "depth_" + name + ".npz"))
    rel, psnr, da = cases.compare_images(g["img"], img)
    assert rel <= 1e-4 and psnr >= 80.0, (rel, psnr, da)
    assert np.allclose(z, g["z"], rtol=2e-5, atol=1e-6), float(np.abs(z - g["z"]).max())


@pytest.mark.parametrize("name", list(cases.RAY_CASES))
def test_ray_parity(name):
    fn, kw = cases.RAY_CASES[name]
    yi = refapi.EmuInterface()
    fn(yi, **kw)
    yi.updateScene()
    g = np.load(os.path.join(G, name + ".npz"))
    got = yi.traceBatch(g["rays"], 0)
    bad, terr, berr, ntie = cases.compare_hits(g["closest"], got)
    # north_star: ids exact except t-ties, t and barycentrics to 1e-5; shadow flags equal
    assert bad == 0 and terr <= 1e-5 and berr <= 1e-5, (bad, terr, berr, ntie)
    sh = yi.traceBatch(g["rays"], 1)
    assert np.count_nonzero(sh != g["shadow"]) == 0


def test_empty_scene_and_miss():
    yi = refapi.EmuInterface()
    from core_b200 import scenes
    yi.startScene(0)
    scenes.create(yi, "camera", "cam", type="perspective", resx=8, resy=8)
    scenes.create(yi, "background", "bg", type="constant", color=(0.25, 0.5, 0.75, 1.0))
    scenes.create(yi, "integrator", "default", type="directlighting")
    scenes.std_volintegr(yi)
    scenes.set_render(yi, 8, 8, spp=2)
    img = yi.render(8, 8)
    assert np.allclose(img[..., :3], [0.25, 0.5, 0.75]) and np.all(img[..., 3] == 0.0)   # bg_transp default (directlight.cc:196)
    assert len(yi.traceBatch(np.zeros((0, 8), np.float32), 0)) == 0


def test_russian_roulette_extension_is_unbiased():
    """Opt-in extension (OFF in every parity case): image mean within noise of the RR-off render."""
    from core_b200 import scenes
    out = []
    for rr in (False, True):
        yi = refapi.EmuInterface()
        scenes.cornell_box(yi, 32, 32, spp=16, light_samples=1, mirror_box=False, integrator="pathtracing",
                           integ_params=dict(bounces=6, path_samples=1, russian_roulette=rr))
        out.append(yi.render(32, 32)[..., :3].mean())
    assert abs(out[0] - out[1]) / out[0] < 0.03, out


@pytest.mark.parametrize("name,var", [("c4_materials_path", "YB200_STAGE_DIRECT"), ("c4_materials_direct", "YB200_STAGE_DIRECT"),
                                      ("c2_field_path", "YB200_KIND_SHADE"), ("c1_cornell_direct", "YB200_KIND_SHADE")])
def test_shade_variants_render_the_same_frame(name, var, monkeypatch):
    """Staged direct lighting (job queue + one evaluation per light type x material class) and the shade builds per kind of task /
    material set against the single inline kernel: the emulator runs tasks in order, so the frames agree to float rounding of the
    accumulation order and the ray counts are equal (same test on the GPU: test_gpu_parity.py)."""
    fn, kw = cases.IMAGE_CASES[name]
    frames, rays = [], []
    for val in ("1", "0"):
        monkeypatch.setenv(var, val)
        yi = refapi.EmuInterface()
        fn(yi, **kw)
        frames.append(yi.render(kw["width"], kw["height"]))
        st = yi.getStats()
        rays.append((st["rays_closest"], st["rays_shadow"]))
        monkeypatch.delenv(var)
    rel, psnr, da = cases.compare_images(frames[0], frames[1])
    assert rays[0] == rays[1], (name, var, rays)
    assert rel <= 1e-6 and psnr >= 110.0 and da <= 1e-6, (name, var, rel, psnr, da)
