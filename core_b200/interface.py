"""Host-side mirror of ``yafrayInterface_t`` (include/interface/yafrayinterface.h:45-146 of the reference).

Same method names, argument meaning and error behaviour (bool / 0-handle returns, ``render`` silently
returning when the scene is not set up, yafrayinterface.cc:336-342) — over the C ABI of
include/yb200.h, so a test reads like a reference test.  Adds ``traceBatch`` (parity hook) and
``renderFilm``/``filmResolve`` (multi-GPU).  numpy arrays are host buffers; torch tensors on CUDA are
passed by device pointer.
"""
import ctypes as C

import numpy as np

from . import capi


def _b(s):
    return s.encode() if isinstance(s, str) else s


class YafrayInterface:
    def __init__(self, clib=None, device=None):
        self._l = clib or capi.product_lib()
        self._c = C.c_void_p(self._l.fn("create")())
        if not self._c:
            raise RuntimeError("yb200_create failed")
        if device is not None and not self._call("set_device", int(device)):
            raise RuntimeError(self.lastError())
        self._keep = []

    def _call(self, name, *a):
        return self._l.fn(name)(self._c, *a)

    def close(self):
        if self._c:
            self._call("destroy")
            self._c = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --- misc
    def lastError(self):
        return (self._call("last_error") or b"").decode()

    def getVersion(self):
        return self._l.fn("version")().decode()

    def loadPlugins(self, path=None):
        """No-op: the supported plugin set is compiled in (environment.cc:119-135 replaced)."""

    def setVerbosityLevel(self, lvl): self._call("set_verbosity", int(lvl))
    def setVerbosityMute(self): self._call("set_verbosity", 0)
    def setVerbosityError(self): self._call("set_verbosity", 1)
    def setVerbosityWarning(self): self._call("set_verbosity", 2)
    def setVerbosityInfo(self): self._call("set_verbosity", 3)
    def clearAll(self): self._call("clear_all")

    # --- scene
    def startScene(self, type=0): return bool(self._call("start_scene", type))
    def startGeometry(self): return bool(self._call("start_geometry"))
    def endGeometry(self): return bool(self._call("end_geometry"))
    def getNextFreeID(self): return self._call("get_next_free_id")

    def startTriMesh(self, id, vertices, triangles, hasOrco, hasUV=False, type=0):
        return bool(self._call("start_trimesh", id, vertices, triangles, int(hasOrco), int(hasUV), type))

    def endTriMesh(self): return bool(self._call("end_trimesh"))

    def addVertex(self, x, y, z, ox=None, oy=None, oz=None):
        if ox is None:
            return self._call("add_vertex", x, y, z)
        return self._call("add_vertex_orco", x, y, z, ox, oy, oz)

    def addNormal(self, nx, ny, nz): self._call("add_normal", nx, ny, nz)

    def addTriangle(self, a, b, c, *rest):
        if len(rest) == 1:
            return bool(self._call("add_triangle", a, b, c, rest[0]))
        ua, ub, uc, mat = rest
        return bool(self._call("add_triangle_uv", a, b, c, ua, ub, uc, mat))

    def addUV(self, u, v): return self._call("add_uv", u, v)
    def smoothMesh(self, id, angle): return bool(self._call("smooth_mesh", id, angle))

    def addInstance(self, baseObjectId, objToWorld):
        m = (C.c_float * 16)(*np.asarray(objToWorld, dtype=np.float32).reshape(16))
        return bool(self._call("add_instance", baseObjectId, m))

    # bulk forms (numpy): same semantics as n single calls
    def addVertices(self, xyz, normals=None):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        nrm = None if normals is None else np.ascontiguousarray(normals, dtype=np.float32)
        return bool(self._call("add_vertices", xyz.ctypes.data, None if nrm is None else nrm.ctypes.data,
                               xyz.shape[0]))

    def addTriangles(self, abc, mat):
        abc = np.ascontiguousarray(abc, dtype=np.int32)
        return bool(self._call("add_triangles", abc.ctypes.data, abc.shape[0], mat))

    # --- params
    def paramsSetPoint(self, n, x, y, z): self._call("params_set_point", _b(n), x, y, z)
    def paramsSetString(self, n, s): self._call("params_set_string", _b(n), _b(s))
    def paramsSetBool(self, n, b): self._call("params_set_bool", _b(n), int(bool(b)))
    def paramsSetInt(self, n, i): self._call("params_set_int", _b(n), int(i))
    def paramsSetFloat(self, n, f): self._call("params_set_float", _b(n), float(f))
    def paramsSetColor(self, n, r, g, b, a=1.0): self._call("params_set_color", _b(n), r, g, b, a)

    def paramsSetMatrix(self, n, m, transpose=False):
        mm = (C.c_float * 16)(*np.asarray(m, dtype=np.float32).reshape(16))
        self._call("params_set_matrix", _b(n), mm, int(transpose))

    def paramsClearAll(self): self._call("params_clear_all")
    def paramsStartList(self): self._call("params_start_list")
    def paramsPushList(self): self._call("params_push_list")
    def paramsEndList(self): self._call("params_end_list")

    # --- factories (0 == NULL on failure, like the reference)
    def createLight(self, name): return self._call("create_light", _b(name))
    def createMaterial(self, name): return self._call("create_material", _b(name))
    def createCamera(self, name): return self._call("create_camera", _b(name))
    def createBackground(self, name): return self._call("create_background", _b(name))
    def createIntegrator(self, name): return self._call("create_integrator", _b(name))
    def registerImage(self, file_name, rgba, is_hdr=False):
        """Decoded pixels (H, W, 4) float32, row 0 on top, for image textures whose `filename` parameter is file_name."""
        rgba = np.ascontiguousarray(rgba, dtype=np.float32)
        return bool(self._call("register_image", _b(file_name), rgba.ctypes.data, rgba.shape[1], rgba.shape[0], int(is_hdr)))

    def createTexture(self, name): return self._call("create_texture", _b(name))
    def createVolumeRegion(self, name): return self._call("create_volume_region", _b(name))

    # --- render
    def render(self, width, height):
        """yafrayInterface_t::render into a memoryIO_t-layout buffer; returns (H, W, 4) float32."""
        out = np.zeros((height, width, 4), dtype=np.float32)
        ok = self._call("render", out.ctypes.data, width, height, None)
        if not ok:
            raise RuntimeError("render failed: " + self.lastError())
        return out

    def getDepth(self, width, height):
        """Depth channel of the last render (render parameter z_channel = true): what imageFilm_t::flush passes to
        colorOutput_t::putPixel(..., depth=true, z); returns (H, W) float32."""
        z = np.zeros((height, width), dtype=np.float32)
        if not self._call("get_depth", z.ctypes.data):
            raise RuntimeError("get_depth failed: " + self.lastError())
        return z

    def depthFilm(self):
        """Device pointer of this rank's raw depth film (H, W, 2) or None."""
        return self._call("depth_film")

    def depthResolve(self, depth_film_ptr, width, height):
        z = np.zeros((height, width), dtype=np.float32)
        if not self._call("depth_resolve", depth_film_ptr, z.ctypes.data):
            raise RuntimeError("depth_resolve failed: " + self.lastError())
        return z

    def abort(self): self._call("abort")
    def updateScene(self):
        if not self._call("update_scene"):
            raise RuntimeError("update_scene failed: " + self.lastError())

    def renderFilm(self, film, rank=0, world=1):
        """film: zeroed CUDA float32 torch tensor (H, W, 5); accumulates this rank's tiles."""
        if not self._call("render_film", film.data_ptr(), rank, world, None):
            raise RuntimeError("render_film failed: " + self.lastError())

    def renderFilmPass(self, film, film_all, rank, world, aa_pass):
        """One AA pass (adaptive sampling on several ranks): film_all = sum of the films of all ranks after the
        previous pass (None for pass 0 / a single rank)."""
        if not self._call("render_film_pass", film.data_ptr(), film_all.data_ptr() if film_all is not None else None,
                          rank, world, int(aa_pass), None):
            raise RuntimeError("render_film_pass failed: " + self.lastError())

    def filmResolve(self, film, width, height):
        out = np.zeros((height, width, 4), dtype=np.float32)
        if not self._call("film_resolve", film.data_ptr(), out.ctypes.data, 0):
            raise RuntimeError("film_resolve failed: " + self.lastError())
        return out

    # --- parity hook
    def traceBatch(self, rays, mode=0, shadow_depth=5):
        """rays (n, 8) float32 {from, tmin, dir, tmax}. Host numpy in, numpy structured array out."""
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        n = rays.shape[0]
        if mode == 0:
            out = np.zeros(n, dtype=[("prim", "<i4"), ("t", "<f4"), ("b1", "<f4"), ("b2", "<f4")])
        elif mode == 1:
            out = np.zeros(n, dtype=np.int32)
        else:
            out = np.zeros(n, dtype=[("shadowed", "<i4"), ("r", "<f4"), ("g", "<f4"), ("b", "<f4")])
        if n and not self._call("trace_batch", rays.ctypes.data, n, mode, shadow_depth, out.ctypes.data, 0):
            raise RuntimeError("trace_batch failed: " + self.lastError())
        return out

    def traceBatchDevice(self, rays_ptr, n, mode, out_ptr, shadow_depth=5):
        if not self._call("trace_batch", rays_ptr, n, mode, shadow_depth, out_ptr, 1):
            raise RuntimeError("trace_batch failed: " + self.lastError())

    def getStats(self):
        s = capi.Stats()
        self._call("get_stats", C.byref(s))
        return {k: getattr(s, k) for k, _ in capi.Stats._fields_}

    def setDevices(self, devices):
        """Several GPUs behind this one context (yb200_set_devices): render() then spreads the tiles over them."""
        arr = (C.c_int * len(devices))(*[int(d) for d in devices])
        if not self._call("set_devices", arr, len(devices)):
            raise RuntimeError("set_devices failed: " + self.lastError())

    def probeLatency(self, nbytes, steps=20000):
        """ns per dependent load through a random cycle over nbytes (yb200_probe_latency)."""
        ns = C.c_double(0.0)
        if not self._call("probe_latency", int(nbytes), int(steps), C.byref(ns)):
            raise RuntimeError("probe_latency failed: " + self.lastError())
        return ns.value

    def touchGeometry(self):
        self._call("touch_geometry")

    def cudaStream(self):
        return self._call("cuda_stream")
