"""ctypes binding of the C ABI declared in include/yb200.h.

The same binding class also fits any library exporting that entry-point set under another prefix
(the tests bind oracle/_ref/libyref.so, prefix ``yref_``, through it); the product only ever loads
``core_b200/libyb200.so`` and raises if it is missing — there is no CPU fallback.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libyb200.so")

HANDLE = C.c_size_t


class Callbacks(C.Structure):
    _fields_ = [("user", C.c_void_p),
                ("should_abort", C.CFUNCTYPE(C.c_int, C.c_void_p)),
                ("progress", C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int))]


class Stats(C.Structure):
    _fields_ = [("n_triangles", C.c_uint64), ("n_bvh_nodes", C.c_uint64),
                ("rays_closest", C.c_uint64), ("rays_shadow", C.c_uint64),
                ("camera_samples", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("build_ms", C.c_double), ("render_ms", C.c_double),
                ("trace_closest_ms", C.c_double), ("trace_shadow_ms", C.c_double),
                ("shade_ms", C.c_double), ("film_ms", C.c_double),
                ("bvh_bytes", C.c_uint64), ("tri_bytes", C.c_uint64)]


# name -> (restype, argtypes after the ctx pointer); mirrors include/yb200.h one to one
_CTX_FUNCS = {
    "destroy": (None, []),
    "last_error": (C.c_char_p, []),
    "set_device": (C.c_int, [C.c_int]),
    "set_verbosity": (None, [C.c_int]),
    "clear_all": (None, []),
    "start_scene": (C.c_int, [C.c_int]),
    "start_geometry": (C.c_int, []),
    "end_geometry": (C.c_int, []),
    "get_next_free_id": (C.c_uint, []),
    "start_trimesh": (C.c_int, [C.c_uint, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "end_trimesh": (C.c_int, []),
    "add_vertex": (C.c_int, [C.c_double] * 3),
    "add_vertex_orco": (C.c_int, [C.c_double] * 6),
    "add_normal": (None, [C.c_double] * 3),
    "add_triangle": (C.c_int, [C.c_int, C.c_int, C.c_int, HANDLE]),
    "add_triangle_uv": (C.c_int, [C.c_int] * 6 + [HANDLE]),
    "add_uv": (C.c_int, [C.c_float, C.c_float]),
    "smooth_mesh": (C.c_int, [C.c_uint, C.c_double]),
    "add_instance": (C.c_int, [C.c_uint, C.POINTER(C.c_float)]),
    "add_vertices": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "add_triangles": (C.c_int, [C.c_void_p, C.c_int, HANDLE]),
    "params_set_point": (None, [C.c_char_p, C.c_double, C.c_double, C.c_double]),
    "params_set_string": (None, [C.c_char_p, C.c_char_p]),
    "params_set_bool": (None, [C.c_char_p, C.c_int]),
    "params_set_int": (None, [C.c_char_p, C.c_int]),
    "params_set_float": (None, [C.c_char_p, C.c_double]),
    "params_set_color": (None, [C.c_char_p, C.c_float, C.c_float, C.c_float, C.c_float]),
    "params_set_matrix": (None, [C.c_char_p, C.POINTER(C.c_float), C.c_int]),
    "params_clear_all": (None, []),
    "params_start_list": (None, []),
    "params_push_list": (None, []),
    "params_end_list": (None, []),
    "create_light": (HANDLE, [C.c_char_p]),
    "create_material": (HANDLE, [C.c_char_p]),
    "create_camera": (HANDLE, [C.c_char_p]),
    "create_background": (HANDLE, [C.c_char_p]),
    "create_integrator": (HANDLE, [C.c_char_p]),
    "create_texture": (HANDLE, [C.c_char_p]),
    "create_volume_region": (HANDLE, [C.c_char_p]),
    "render": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "get_depth": (C.c_int, [C.c_void_p]),
    "abort": (None, []),
    "update_scene": (C.c_int, []),
    "trace_batch": (C.c_int, [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_int]),
}
# only libyb200.so has these (the multi-GPU film path, counters, stream)
_PRODUCT_ONLY = {
    "render_film": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    "render_film_pass": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "film_resolve": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int]),
    "register_image": (C.c_int, [C.c_char_p, C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "depth_film": (C.c_void_p, []),
    "depth_resolve": (C.c_int, [C.c_void_p, C.c_void_p]),
    "get_stats": (C.c_int, [C.POINTER(Stats)]),
    "cuda_stream": (C.c_void_p, []),
    "touch_geometry": (None, []),
    "trace_batch_static": (C.c_int, [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_int]),
    "selftest": (C.c_int, [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]),
    "probe_latency": (C.c_int, [C.c_size_t, C.c_int, C.POINTER(C.c_double)]),
    "set_devices": (C.c_int, [C.POINTER(C.c_int), C.c_int]),
}
EXPORTED_SYMBOLS = (["create", "version"] + sorted(_CTX_FUNCS) + sorted(_PRODUCT_ONLY))


class CLib:
    """One loaded library + prefix; attribute access returns the bound C function."""

    def __init__(self, path=LIB_PATH, prefix="yb200_", product=True):
        if not os.path.exists(path):
            raise RuntimeError(
                f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(core_b200 has no CPU fallback)")
        self.path, self.prefix, self.product = path, prefix, product
        self.lib = C.CDLL(path, mode=C.RTLD_GLOBAL if not product else C.RTLD_LOCAL)
        f = getattr(self.lib, prefix + "create")
        f.restype, f.argtypes = C.c_void_p, []
        f = getattr(self.lib, prefix + "version")
        f.restype, f.argtypes = C.c_char_p, []
        table = dict(_CTX_FUNCS)
        if product:
            table.update(_PRODUCT_ONLY)
        for name, (res, args) in table.items():
            f = getattr(self.lib, prefix + name)
            f.restype, f.argtypes = res, [C.c_void_p] + list(args)

    def fn(self, name):
        return getattr(self.lib, self.prefix + name)


_product = None


def product_lib():
    global _product
    if _product is None:
        _product = CLib()
    return _product
