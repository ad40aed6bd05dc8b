"""ctypes access to oracle/libyb_oracle.so (the plain-C restatement) + a recording yafrayInterface_t stand-in that
captures geometry in the reference's global primitive order.  TEST INFRASTRUCTURE."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "oracle", "libyb_oracle.so")
HIT = np.dtype([("prim", "<i4"), ("t", "<f4"), ("b1", "<f4"), ("b2", "<f4")])
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO):
            subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "port"], check=True, stdout=subprocess.DEVNULL)
        L = C.CDLL(SO)
        L.yo_halton.restype = C.c_double; L.yo_halton.argtypes = [C.c_uint, C.c_uint, C.c_int]
        for f in ("yo_ri_vdc", "yo_ri_s", "yo_ri_lp"):
            getattr(L, f).restype = C.c_float; getattr(L, f).argtypes = [C.c_uint, C.c_uint]
        L.yo_fnv.restype = C.c_uint; L.yo_fnv.argtypes = [C.c_uint]
        L.yo_scr_halton.restype = C.c_double; L.yo_scr_halton.argtypes = [C.c_int, C.c_uint]
        for f in ("yo_fsin", "yo_fcos"):
            getattr(L, f).restype = C.c_float; getattr(L, f).argtypes = [C.c_float]
        L.yo_fpow.restype = C.c_float; L.yo_fpow.argtypes = [C.c_float, C.c_float]
        L.yo_closest.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.yo_shadow.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        _lib = L
    return _lib


def closest(tris, rays):
    tris = np.ascontiguousarray(tris, np.float32); rays = np.ascontiguousarray(rays, np.float32)
    out = np.zeros(len(rays), HIT); ties = np.zeros(len(rays), np.int32)
    lib().yo_closest(tris.ctypes.data, len(tris), rays.ctypes.data, len(rays), out.ctypes.data, ties.ctypes.data)
    return out, ties


def shadow(tris, rays):
    tris = np.ascontiguousarray(tris, np.float32); rays = np.ascontiguousarray(rays, np.float32)
    out = np.zeros(len(rays), np.int32)
    lib().yo_shadow(tris.ctypes.data, len(tris), rays.ctypes.data, len(rays), out.ctypes.data)
    return out


class Recorder:
    """Accepts the scene-building calls of yafrayInterface_t and keeps the triangles in the global primitive
    order of scene_t::update (ascending object id; ids from getNextFreeID count down from INT_MAX,
    scene.cc:53,760-781,950-966)."""

    def __init__(self):
        self.meshes = {}
        self.next_id = 2 ** 31 - 1
        self.cur = None

    def __getattr__(self, name):          # params*, create*, smoothMesh ...: accepted and ignored
        def f(*a, **k):
            return 1
        return f

    def getNextFreeID(self):
        i = self.next_id
        self.next_id -= 1
        return i

    def startTriMesh(self, id, nv, nt, orco, uv=False, type=0):
        self.cur = {"v": [], "t": []}
        self.meshes[id] = self.cur
        return True

    def addVertex(self, x, y, z, *o):
        self.cur["v"].append((x, y, z))
        return len(self.cur["v"]) - 1

    def addVertices(self, xyz, normals=None):
        self.cur["v"] += [tuple(p) for p in np.asarray(xyz, np.float32)]
        return True

    def addTriangle(self, a, b, c, *rest):
        self.cur["t"].append((a, b, c))
        return True

    def addTriangles(self, abc, mat):
        self.cur["t"] += [tuple(t) for t in np.asarray(abc)]
        return True

    def triangles(self):
        out = []
        for i in sorted(self.meshes):
            m = self.meshes[i]
            v = np.asarray(m["v"], np.float32).reshape(-1, 3)
            t = np.asarray(m["t"], np.int64).reshape(-1, 3)
            out.append(v[t].reshape(-1, 9))
        return np.concatenate(out).astype(np.float32)
