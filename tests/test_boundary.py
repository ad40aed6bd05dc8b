"""The drop-in boundary: libyb200.so exports exactly what include/yb200.h declares, mirrors the reference's error
behaviour, and fails LOUDLY (no CPU fallback) for unsupported features or when no CUDA device exists."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import refapi
from core_b200 import capi, scenes
from core_b200.interface import YafrayInterface

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    h = open(os.path.join(ROOT, "include", "yb200.h")).read()
    return sorted(set(re.findall(r"\b(yb200_[a-z_0-9]+)\s*\(", h)))


@pytest.mark.skipif(not os.path.exists(capi.LIB_PATH), reason="libyb200.so not built")
def test_product_library_exports_every_declared_symbol():
    lib = C.CDLL(capi.LIB_PATH)
    for s in header_symbols():
        assert hasattr(lib, s), s
    assert len(header_symbols()) >= 45


@pytest.mark.skipif(not os.path.exists(capi.LIB_PATH), reason="libyb200.so not built")
def test_product_has_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    yi = YafrayInterface()
    assert "EMULATOR" not in yi.getVersion()
    scenes.cornell_box(yi, 8, 8, spp=1)
    with pytest.raises(RuntimeError, match="no CUDA device|CPU fallback"):
        yi.render(8, 8)


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(RuntimeError, match="missing"):
        capi.CLib(str(tmp_path / "nope.so"))


@pytest.mark.skipif(not refapi.have_emu(), reason="host emulator could not be built")
class TestHostLogic:
    def test_state_machine_mirrors_scene_t(self):
        yi = refapi.EmuInterface()
        yi.setVerbosityMute()
        yi.startScene(0)
        assert not yi.startTriMesh(1, 3, 1, False)          # not inside startGeometry (scene.cc:268)
        assert yi.startGeometry() and not yi.startGeometry()
        assert yi.addVertex(0, 0, 0) == -1                   # no object open (scene.cc:524)
        assert yi.getNextFreeID() == 2 ** 31 - 1 and yi.getNextFreeID() == 2 ** 31 - 2
        m = scenes.create(yi, "material", "m", type="shinydiffusemat")
        assert yi.startTriMesh(5, 3, 1, False)
        assert [yi.addVertex(*p) for p in ((0, 0, 0), (1, 0, 0), (0, 1, 0))] == [0, 1, 2]
        assert yi.addTriangle(0, 1, 2, m) and not yi.addTriangle(0, 1, 2, 12345)
        assert yi.endTriMesh() and not yi.endTriMesh() and yi.endGeometry()
        assert not yi.startScene(1)                          # universal mode unsupported

    def test_unsupported_features_fail(self):
        yi = refapi.EmuInterface()
        yi.setVerbosityMute()
        yi.startScene(0)
        for kind, ty in (("material", "translucent"), ("material", "glossy2"), ("light", "ieslight"),
                         ("camera", "equirectangular"), ("background", "sunsky"), ("integrator", "photonmapping")):
            yi.paramsClearAll(); yi.paramsSetString("type", ty)
            assert getattr(yi, "create" + kind.capitalize())("x") == 0
            assert "not supported" in yi.lastError()
        yi.paramsClearAll(); yi.paramsSetString("type", "blend_mat"); yi.paramsSetString("material1", "nope"); yi.paramsSetString("material2", "nope")
        assert yi.createMaterial("bl") == 0 and "does not exist" in yi.lastError()
        yi.paramsClearAll(); yi.paramsSetString("type", "rough_glass"); yi.paramsSetFloat("dispersion_power", 0.2)
        assert yi.createMaterial("rgd") == 0 and "dispersion" in yi.lastError()
        yi.paramsClearAll(); yi.paramsSetString("type", "blend_mat"); yi.paramsSetString("material1", "m1"); yi.paramsSetString("material2", "m1")
        yi.paramsSetString("mask", "node")
        assert yi.createMaterial("masked") == 0
        yi.paramsClearAll(); yi.paramsSetString("type", "shinydiffusemat"); yi.paramsSetString("diffuse_shader", "n")
        yi.paramsPushList(); yi.paramsSetString("element", "shader_node"); yi.paramsSetString("name", "n")
        yi.paramsSetString("type", "texture_mapper"); yi.paramsSetString("texture", "missing"); yi.paramsEndList()
        assert yi.createMaterial("withnodes") == 0 and "does not exist" in yi.lastError()
        yi.paramsClearAll(); yi.paramsSetString("type", "image"); yi.paramsSetString("filename", "/nonexistent/picture.png")
        assert yi.createTexture("png") == 0 and "register" in yi.lastError()
        yi.paramsClearAll(); yi.paramsSetString("type", "image"); yi.paramsSetString("filename", "/nonexistent/picture.tga")
        assert yi.createTexture("tga") == 0 and "Cannot open" in yi.lastError()
        yi.paramsClearAll(); yi.paramsSetString("type", "fractal_flames")
        assert yi.createTexture("ff") == 0 and "not supported" in yi.lastError()
        yi.paramsClearAll(); yi.paramsSetString("type", "glass"); yi.paramsSetFloat("dispersion_power", 0.1)
        assert yi.createMaterial("disp") == 0
        yi.paramsClearAll()
        assert yi.createTexture("t") == 0 and yi.createVolumeRegion("v") == 0
        yi.paramsClearAll(); yi.paramsSetString("type", "shinydiffusemat")
        assert yi.createMaterial("dup") != 0 and yi.createMaterial("dup") == 0     # name clash (environment.cc:226)

    def test_registered_image_equals_tga_file(self, tmp_path):
        """yb200_register_image (decoded pixels handed over by the caller) and the built-in TGA reader feed the same texture."""
        import numpy as np
        from core_b200 import scenes as sc
        img8 = sc.test_image(24, 12, seed=9)
        path = sc.write_tga(str(tmp_path / "t.tga"), img8, top=False, rle=True)
        frames = []
        for registered in (False, True):
            yi = refapi.EmuInterface(); yi.setVerbosityMute()
            name = path if not registered else "memory://picture.png"
            if registered:
                px = (img8.astype(np.float64) * 0.00392156862745098039).astype(np.float32)
                assert yi.registerImage(name, px)
            yi.startScene(0)
            sc.create(yi, "texture", "pic", type="image", filename=name, interpolate="bicubic")
            m = sc.create(yi, "material", "m", type="shinydiffusemat", color=(1.0, 1.0, 1.0, 1.0), diffuse_shader="map",
                          nodes=[dict(name="map", type="texture_mapper", texture="pic", texco="global", scale=(0.5, 0.5, 0.5))])
            yi.startGeometry()
            sc.add_mesh(yi, *sc.quad((-2, -2, 0), (2, -2, 0), (2, 2, 0), (-2, 2, 0)), m, bulk=False)
            yi.endGeometry()
            sc.create(yi, "light", "pt", type="pointlight", color=(1.0, 1.0, 1.0, 1.0), power=30.0, **{"from": (0.5, 0.3, 3.0)})
            sc.create(yi, "camera", "cam", type="perspective", to=(0.0, 0.0, 0.0), up=(0.0, 1.0, 5.0), resx=24, resy=24, focal=1.0,
                      **{"from": (0.0, 0.0, 5.0)})
            sc.create(yi, "background", "bg", type="constant", color=(0.0, 0.0, 0.0, 1.0))
            sc.create(yi, "integrator", "default", type="directlighting")
            sc.std_volintegr(yi)
            sc.set_render(yi, 24, 24, spp=1)
            frames.append(yi.render(24, 24))
        assert frames[0][..., :3].std() > 0.01 and np.array_equal(frames[0], frames[1])

    def test_render_requires_volume_integrator_none(self):
        yi = refapi.EmuInterface()
        yi.setVerbosityMute()
        scenes.cornell_box(yi, 8, 8, spp=1)
        yi.paramsSetString("volintegrator_name", "missing")
        with pytest.raises(RuntimeError):
            yi.render(8, 8)


def test_getyafray_library_exports_the_reference_entry_point():
    """libyb200_yafray.so exports `getYafray` (yafrayinterface.cc:386-392) and the object it returns answers
    getVersion through the C++ vtable of include/yb200_interface.hpp (no device work)."""
    import ctypes
    import os
    so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "core_b200", "libyb200_yafray.so")
    assert os.path.exists(so), "run __graft_entry__.build()"
    lib = ctypes.CDLL(so)
    lib.getYafray.restype = ctypes.c_void_p
    obj = lib.getYafray()
    assert obj
