"""Procedural scene generators for the five BASELINE.json configs.

Every generator only calls ``yafrayInterface_t`` methods on ``yi`` (core_b200.interface.YafrayInterface
over libyb200.so, or the same class bound to the reference driver in the tests), so both sides are
fed bit-identical inputs.  All randomness is seeded numpy (PCG64) — deterministic across machines.
"""
import numpy as np


# ----------------------------------------------------------------------------- helpers
def set_render(yi, width, height, spp=1, integrator="default", camera="cam", background="bg",
               threads=1, filter_type="box", pixelwidth=1.5, xstart=0, ystart=0, passes=1, **extra):
    yi.paramsClearAll()
    yi.paramsSetString("camera_name", camera)
    yi.paramsSetString("integrator_name", integrator)
    yi.paramsSetString("volintegrator_name", "volintegr")
    if background:
        yi.paramsSetString("background_name", background)
    yi.paramsSetInt("width", width)
    yi.paramsSetInt("height", height)
    yi.paramsSetInt("xstart", xstart)
    yi.paramsSetInt("ystart", ystart)
    yi.paramsSetInt("AA_minsamples", spp)
    yi.paramsSetInt("AA_passes", passes)
    yi.paramsSetFloat("AA_pixelwidth", pixelwidth)
    yi.paramsSetString("filter_type", filter_type)
    yi.paramsSetInt("threads", threads)
    yi.paramsSetFloat("gamma", 1.0)
    yi.paramsSetBool("clamp_rgb", False)
    yi.paramsSetBool("z_channel", False)
    for k, v in extra.items():
        _set(yi, k, v)


def _set(yi, k, v):
    if isinstance(v, bool):
        yi.paramsSetBool(k, v)
    elif isinstance(v, int):
        yi.paramsSetInt(k, v)
    elif isinstance(v, float):
        yi.paramsSetFloat(k, v)
    elif isinstance(v, str):
        yi.paramsSetString(k, v)
    elif isinstance(v, (tuple, list)) and len(v) == 3:
        yi.paramsSetPoint(k, *v)
    elif isinstance(v, (tuple, list)) and len(v) == 4:
        yi.paramsSetColor(k, *v)
    elif isinstance(v, dict) and "point" in v:
        yi.paramsSetPoint(k, *v["point"])
    else:
        raise TypeError(f"param {k}: {v!r}")


def create(yi, kind, name, nodes=None, **params):
    """kind in light|material|texture|camera|background|integrator; colours are 4-tuples, points 3-tuples.
    nodes: list of shader-node parameter dicts (each with name / type / ...), sent as the material's parameter list
    (paramsPushList per node + paramsEndList, what the XML <list_element> children become)."""
    yi.paramsClearAll()
    for k, v in params.items():
        _set(yi, k, v)
    if nodes:                                   # like the exporters: one paramsPushList per node (paramsStartList on an empty
        for nd in nodes:                        # list dereferences eparams->back() in the reference, yafrayinterface.cc:230-235)
            yi.paramsPushList()
            _set(yi, "element", "shader_node")
            for k, v in nd.items():
                _set(yi, k, v)
        yi.paramsEndList()
    h = getattr(yi, "create" + kind.capitalize())(name)
    if not h:
        raise RuntimeError(f"create{kind.capitalize()}({name!r}, {params.get('type')!r}) failed")
    return h


def add_mesh(yi, verts, tris, mat, normals=None, smooth_angle=None, oid=None, bulk=True):
    verts = np.asarray(verts, dtype=np.float32)
    tris = np.asarray(tris, dtype=np.int32)
    oid = yi.getNextFreeID() if oid is None else oid
    assert yi.startTriMesh(oid, len(verts), len(tris), False, False, 0)
    if bulk and hasattr(yi, "addVertices"):
        assert yi.addVertices(verts, normals)
        assert yi.addTriangles(tris, mat)
    else:
        for i, v in enumerate(verts):
            yi.addVertex(float(v[0]), float(v[1]), float(v[2]))
            if normals is not None:
                yi.addNormal(*[float(x) for x in normals[i]])
        for t in tris:
            assert yi.addTriangle(int(t[0]), int(t[1]), int(t[2]), mat)
    assert yi.endTriMesh()
    if smooth_angle is not None:
        assert yi.smoothMesh(oid, smooth_angle)
    return oid


def add_mesh_uv_orco(yi, verts, tris, mat, orco, uv, normals=None, smooth_angle=None, oid=None):
    """Mesh with original coordinates (has_orco) and one UV value per vertex (has_uv), through the single-item calls."""
    oid = yi.getNextFreeID() if oid is None else oid
    assert yi.startTriMesh(oid, len(verts), len(tris), True, True, 0)
    for i, v in enumerate(verts):
        yi.addVertex(float(v[0]), float(v[1]), float(v[2]), float(orco[i][0]), float(orco[i][1]), float(orco[i][2]))
        if normals is not None:
            yi.addNormal(*[float(x) for x in normals[i]])
    ids = [yi.addUV(float(u), float(v)) for u, v in uv]
    for t in tris:
        assert yi.addTriangle(int(t[0]), int(t[1]), int(t[2]), ids[int(t[0])], ids[int(t[1])], ids[int(t[2])], mat)
    assert yi.endTriMesh()
    if smooth_angle is not None:
        assert yi.smoothMesh(oid, smooth_angle)
    return oid


def test_image(w, h, seed=5):
    """Deterministic RGBA8 test picture (H, W, 4) uint8: colour gradients, a checker and an alpha ramp."""
    y, x = np.mgrid[0:h, 0:w]
    rng = np.random.Generator(np.random.PCG64(seed))
    r = (255 * x / max(1, w - 1)).astype(np.uint8)
    g = (255 * y / max(1, h - 1)).astype(np.uint8)
    b = ((((x // 4) + (y // 4)) & 1) * 200 + rng.integers(0, 56, (h, w))).astype(np.uint8)
    a = (64 + 191 * ((x + y) % 16) / 15).astype(np.uint8)
    return np.stack([r, g, b, a], 2)


def write_tga(path, img, top=False, rle=False, gray=False):
    """Write (H, W, 4) uint8 as a TGA file: 32-bit BGRA (24-bit if the alpha is all 255, 8-bit grey if gray), bottom or
    top origin, optionally run-length encoded with raw packets only."""
    import os
    os.makedirs(os.path.dirname(path), exist_ok=True)
    h, w = img.shape[:2]
    has_alpha = bool((img[..., 3] != 255).any()) and not gray
    bpp = 1 if gray else (4 if has_alpha else 3)
    rows = img if top else img[::-1]
    px = rows[..., :1] if gray else (rows[..., [2, 1, 0, 3]] if has_alpha else rows[..., [2, 1, 0]])
    data = np.ascontiguousarray(px).reshape(-1, bpp)
    itype = (3 if gray else 2) + (8 if rle else 0)
    hdr = bytes([0, 0, itype, 0, 0, 0, 0, 0, 0, 0, 0, 0, w & 255, w >> 8, h & 255, h >> 8, 8 * bpp, (0x20 if top else 0) | (8 if has_alpha else 0)])
    with open(path, "wb") as f:
        f.write(hdr)
        if not rle:
            f.write(data.tobytes())
        else:
            for i in range(0, len(data), 128):
                chunk = data[i:i + 128]
                f.write(bytes([len(chunk) - 1]))
                f.write(chunk.tobytes())
    return path


def quad(p0, p1, p2, p3):
    return np.array([p0, p1, p2, p3], dtype=np.float32), np.array([[0, 1, 2], [0, 2, 3]], dtype=np.int32)


def box(lo, hi, rot_z=0.0):
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    c = 0.5 * (lo + hi)
    corners = np.array([[x, y, z] for z in (lo[2], hi[2]) for y in (lo[1], hi[1]) for x in (lo[0], hi[0])])
    if rot_z:
        ca, sa = np.cos(rot_z), np.sin(rot_z)
        d = corners - c
        corners = np.stack([c[0] + ca * d[:, 0] - sa * d[:, 1], c[1] + sa * d[:, 0] + ca * d[:, 1],
                            corners[:, 2]], 1)
    # outward-facing winding
    f = [[0, 2, 3], [0, 3, 1], [4, 5, 7], [4, 7, 6], [0, 1, 5], [0, 5, 4],
         [2, 6, 7], [2, 7, 3], [0, 4, 6], [0, 6, 2], [1, 3, 7], [1, 7, 5]]
    return corners.astype(np.float32), np.array(f, dtype=np.int32)


def icosphere(subdiv):
    t = (1.0 + 5 ** 0.5) / 2
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t),
         (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2),
         (10, 7, 6), (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5),
         (2, 4, 11), (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    v = np.array(v, dtype=np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array(f, dtype=np.int64)
    for _ in range(subdiv):
        e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
        es = np.sort(e, axis=1)
        uniq, inv = np.unique(es, axis=0, return_inverse=True)
        mid = v[uniq[:, 0]] + v[uniq[:, 1]]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        base = len(v)
        v = np.concatenate([v, mid])
        n = len(f)
        ab, bc, ca = base + inv[:n], base + inv[n:2 * n], base + inv[2 * n:]
        f = np.concatenate([np.stack([f[:, 0], ab, ca], 1), np.stack([f[:, 1], bc, ab], 1),
                            np.stack([f[:, 2], ca, bc], 1), np.stack([ab, bc, ca], 1)])
    return v, f.astype(np.int32)


def std_volintegr(yi):
    create(yi, "integrator", "volintegr", type="none")


# ----------------------------------------------------------------------------- C1: Cornell box
def cornell_box(yi, width=512, height=512, spp=16, integrator="directlighting", light_samples=8,
                mirror_box=True, threads=1, integ_params=None, **render_extra):
    """~36 triangles, shinydiffuse walls/boxes (one box with specular_reflect), arealight bound to a
    light_mat quad, perspective camera (BASELINE.json configs[0])."""
    yi.startScene(0)
    white = create(yi, "material", "white", type="shinydiffusemat", color=(0.76, 0.75, 0.5, 1.0), diffuse_reflect=1.0)
    red = create(yi, "material", "red", type="shinydiffusemat", color=(0.63, 0.06, 0.04, 1.0), diffuse_reflect=1.0)
    green = create(yi, "material", "green", type="shinydiffusemat", color=(0.15, 0.48, 0.09, 1.0), diffuse_reflect=1.0)
    shiny = create(yi, "material", "shiny", type="shinydiffusemat", color=(0.7, 0.7, 0.75, 1.0),
                   diffuse_reflect=0.8, specular_reflect=0.35 if mirror_box else 0.0,
                   mirror_color=(1.0, 1.0, 1.0, 1.0))
    lmat = create(yi, "material", "lightmat", type="light_mat", color=(1.0, 0.9, 0.7, 1.0), power=12.0)
    yi.startGeometry()
    add_mesh(yi, *quad((-1, -1, -1), (1, -1, -1), (1, 1, -1), (-1, 1, -1)), white, bulk=False)   # floor
    add_mesh(yi, *quad((-1, -1, 1), (-1, 1, 1), (1, 1, 1), (1, -1, 1)), white, bulk=False)       # ceiling
    add_mesh(yi, *quad((-1, 1, -1), (1, 1, -1), (1, 1, 1), (-1, 1, 1)), white, bulk=False)       # back
    add_mesh(yi, *quad((-1, -1, -1), (-1, 1, -1), (-1, 1, 1), (-1, -1, 1)), red, bulk=False)     # left
    add_mesh(yi, *quad((1, -1, -1), (1, -1, 1), (1, 1, 1), (1, 1, -1)), green, bulk=False)       # right
    add_mesh(yi, *box((-0.7, -0.1, -1.0), (-0.1, 0.5, 0.2), rot_z=0.3), white)
    add_mesh(yi, *box((0.15, -0.55, -1.0), (0.7, 0.0, -0.4), rot_z=-0.3), shiny)
    lid = add_mesh(yi, *quad((-0.3, -0.3, 0.995), (0.3, -0.3, 0.995), (0.3, 0.3, 0.995), (-0.3, 0.3, 0.995)),
                   lmat, bulk=False)
    yi.endGeometry()
    create(yi, "light", "area", type="arealight", corner=(-0.3, -0.3, 0.995), point1=(-0.3, 0.3, 0.995),
           point2=(0.3, -0.3, 0.995), color=(1.0, 0.9, 0.7, 1.0), power=12.0, samples=light_samples,
           object=int(lid))
    create(yi, "camera", "cam", type="perspective", **{"from": (0.0, -3.9, 0.0)}, to=(0.0, 0.0, 0.0),
           up=(0.0, -3.9, 1.0), resx=width, resy=height, focal=1.4)
    create(yi, "background", "bg", type="constant", color=(0.02, 0.02, 0.03, 1.0))
    ip = dict(type=integrator, raydepth=5)
    ip.update(integ_params or {})
    create(yi, "integrator", "default", **ip)
    std_volintegr(yi)
    set_render(yi, width, height, spp=spp, threads=threads, **render_extra)


# ----------------------------------------------------------------------------- C2: icosphere field
def icosphere_field(yi, n_tris=1_000_000, width=1920, height=1080, spp=64, bounces=4, seed=1,
                    subdiv=3, smooth=True, threads=1, integrator="pathtracing", integ_params=None, cam_res=None,
                    background=None, **render_extra):
    """Subdivided icospheres (1280 tris each at subdiv=3) on a jittered grid over a ground quad;
    sun + point light; pathtracing (BASELINE.json configs[1])."""
    rng = np.random.Generator(np.random.PCG64(seed))
    v0, f0 = icosphere(subdiv)
    per = len(f0)
    n_sph = max(1, n_tris // per)
    g = int(np.ceil(np.sqrt(n_sph)))
    yi.startScene(0)
    mats = [create(yi, "material", f"m{i}", type="shinydiffusemat", color=(*c, 1.0), diffuse_reflect=0.9)
            for i, c in enumerate([(0.8, 0.3, 0.25), (0.3, 0.7, 0.3), (0.3, 0.4, 0.8), (0.8, 0.8, 0.7)])]
    ground = create(yi, "material", "ground", type="shinydiffusemat", color=(0.6, 0.6, 0.6, 1.0), diffuse_reflect=0.8)
    yi.startGeometry()
    ext = float(g)
    add_mesh(yi, *quad((-ext, -ext, 0), (ext, -ext, 0), (ext, ext, 0), (-ext, ext, 0)), ground, bulk=False)
    # one mesh per material: bake the spheres (1280 tris x n_sph) into 4 big meshes
    idx = np.arange(n_sph)
    cx = (idx % g - (g - 1) / 2) * 2.0 + rng.uniform(-0.25, 0.25, n_sph)
    cy = (idx // g - (g - 1) / 2) * 2.0 + rng.uniform(-0.25, 0.25, n_sph)
    r = rng.uniform(0.45, 0.8, n_sph)
    cz = r + rng.uniform(0.0, 0.6, n_sph)
    which = rng.integers(0, len(mats), n_sph)
    for m, mat in enumerate(mats):
        sel = np.nonzero(which == m)[0]
        if len(sel) == 0:
            continue
        V = (v0[None] * r[sel, None, None] + np.stack([cx[sel], cy[sel], cz[sel]], 1)[:, None]).reshape(-1, 3)
        N = np.broadcast_to(v0[None], (len(sel),) + v0.shape).reshape(-1, 3) if smooth else None
        F = (f0[None] + (np.arange(len(sel)) * len(v0))[:, None, None]).reshape(-1, 3)
        add_mesh(yi, V.astype(np.float32), F.astype(np.int32), mat,
                 normals=None if N is None else N.astype(np.float32), smooth_angle=181.0 if smooth else None)
    yi.endGeometry()
    create(yi, "light", "sun", type="sunlight", direction=(0.3, -0.4, 0.85), color=(1.0, 0.95, 0.85, 1.0),
           power=1.2, angle=1.0, samples=4)
    create(yi, "light", "pt", type="pointlight", **{"from": (0.0, 0.0, 6.0)}, color=(0.8, 0.85, 1.0, 1.0), power=30.0)
    d = 0.55 * ext
    create(yi, "camera", "cam", type="perspective", **{"from": (0.0, -d * 1.6, d * 0.9)}, to=(0.0, 0.0, 0.5),
           up=(0.0, -d * 1.6, d * 0.9 + 1.0), resx=cam_res[0] if cam_res else width, resy=cam_res[1] if cam_res else height, focal=1.1)
    create(yi, "background", "bg", **(background or dict(type="constant", color=(0.5, 0.6, 0.8, 1.0))))
    ip = dict(type=integrator, raydepth=5)
    if integrator == "pathtracing":
        ip.update(bounces=bounces, path_samples=1)
    ip.update(integ_params or {})
    create(yi, "integrator", "default", **ip)
    std_volintegr(yi)
    set_render(yi, width, height, spp=spp, threads=threads, **render_extra)
    return n_sph * per + 2


# ----------------------------------------------------------------------------- C3: triangle soup + AO
def soup_mesh(n_tris, seed=2):
    """(verts, tris) of the random soup: centres uniform in [-1,1]^3, edge vectors N(0, 0.01^2)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    c = rng.uniform(-1, 1, (n_tris, 3)).astype(np.float32)
    e = rng.normal(0, 0.01, (n_tris, 3, 3)).astype(np.float32)
    V = (c[:, None, :] + e).reshape(-1, 3)
    F = np.arange(3 * n_tris, dtype=np.int32).reshape(-1, 3)
    return V, F


def triangle_soup(yi, n_tris=10_000_000, width=512, height=512, spp=1, ao_samples=32, ao_distance=0.1,
                  seed=2, threads=1, cam_res=None, **render_extra):
    """Random triangles: centres uniform in [-1,1]^3, edges N(0, 0.01^2); directlighting with AO —
    the IntersectS any-hit stress (BASELINE.json configs[2])."""
    V, F = soup_mesh(n_tris, seed)
    yi.startScene(0)
    mat = create(yi, "material", "soup", type="shinydiffusemat", color=(0.8, 0.8, 0.8, 1.0), diffuse_reflect=1.0)
    yi.startGeometry()
    add_mesh(yi, V, F, mat)
    yi.endGeometry()
    create(yi, "light", "pt", type="pointlight", **{"from": (0.0, -3.0, 3.0)}, color=(1.0, 1.0, 1.0, 1.0), power=20.0)
    create(yi, "camera", "cam", type="perspective", **{"from": (0.0, -3.2, 0.4)}, to=(0.0, 0.0, 0.0),
           up=(0.0, -3.2, 1.4), resx=cam_res[0] if cam_res else width, resy=cam_res[1] if cam_res else height, focal=1.2)
    create(yi, "background", "bg", type="constant", color=(0.1, 0.1, 0.1, 1.0))
    create(yi, "integrator", "default", type="directlighting", raydepth=2, do_AO=True, AO_samples=ao_samples,
           AO_distance=float(ao_distance), AO_color=(0.9, 0.9, 0.9, 1.0))
    std_volintegr(yi)
    set_render(yi, width, height, spp=spp, threads=threads, **render_extra)
    return n_tris


def primary_rays(cam_from, cam_to, cam_up, width, height, focal, n=None, seed=0, aspect=1.0):
    """Host restatement of perspectiveCam_t::shootRay (perspectiveCamera.cc:57-71,127-149) at pixel
    centres — only used to make ray batches for the traversal parity tests/bench (float64 math; the
    rays themselves are the shared input, so exactness versus the reference camera is irrelevant)."""
    pos = np.asarray(cam_from, np.float64)
    camY = np.asarray(cam_up, np.float64) - pos
    camZ = np.asarray(cam_to, np.float64) - pos
    camX = np.cross(camZ, camY)
    camY = np.cross(camZ, camX)
    camX /= np.linalg.norm(camX); camY /= np.linalg.norm(camY); camZ /= np.linalg.norm(camZ)
    asp = aspect * height / width
    vright, vup = camX, asp * camY
    vto = camZ * focal - 0.5 * (vup + vright)
    vup = vup / height; vright = vright / width
    ys, xs = np.mgrid[0:height, 0:width]
    px = xs.reshape(-1) + 0.5; py = ys.reshape(-1) + 0.5
    if n is not None and n < px.size:
        sel = np.random.Generator(np.random.PCG64(seed)).choice(px.size, n, replace=False)
        px, py = px[sel], py[sel]
    d = vright[None] * px[:, None] + vup[None] * py[:, None] + vto[None]
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros((d.shape[0], 8), dtype=np.float32)
    rays[:, 0:3] = pos; rays[:, 3] = 0.0; rays[:, 4:7] = d; rays[:, 7] = -1.0
    return rays


# ----------------------------------------------------------------------------- C4: glossy / coated / glass materials
def materials_scene(yi, width=1920, height=1080, spp=256, bounces=8, subdiv=3, threads=1, integrator="pathtracing",
                    integ_params=None, area_samples=8, sun_samples=4, raydepth=5, cam_params=None, lights=None,
                    extra_materials=None, order=None, emitters=None, textures=None, uv_orco=False, image_files=None, cam_res=None,
                    background=None,
                    **render_extra):
    """Row of tessellated spheres over a ground plane with glossy (Blinn exp 50 and anisotropic 20/200),
    coated_glossy (IOR 1.4), glass (IOR 1.5, fake_shadows on and off), mirror and a transparent/mirror
    shinydiffuse; sun (1 degree, 4 samples) + point + area (8 samples) lights; pathtracing with 8 bounces
    (BASELINE.json configs[3]).  The area light has no emitter mesh (includeLights ordering)."""
    yi.startScene(0)
    M = {}
    M["ground"] = create(yi, "material", "ground", type="shinydiffusemat", color=(0.55, 0.55, 0.5, 1.0), diffuse_reflect=0.9)
    M["glossy"] = create(yi, "material", "glossy", type="glossy", color=(0.9, 0.6, 0.3, 1.0), diffuse_color=(0.5, 0.2, 0.1, 1.0),
                         diffuse_reflect=0.4, glossy_reflect=0.8, exponent=50.0, as_diffuse=False)
    M["aniso"] = create(yi, "material", "aniso", type="glossy", color=(0.7, 0.8, 0.9, 1.0), diffuse_color=(0.2, 0.3, 0.5, 1.0),
                        diffuse_reflect=0.3, glossy_reflect=0.9, anisotropic=True, exp_u=20.0, exp_v=200.0, as_diffuse=False)
    M["gdiff"] = create(yi, "material", "gdiff", type="glossy", color=(0.6, 0.9, 0.6, 1.0), diffuse_color=(0.3, 0.6, 0.3, 1.0),
                        diffuse_reflect=0.6, glossy_reflect=0.5, exponent=120.0, as_diffuse=True,
                        diffuse_brdf="Oren-Nayar", sigma=0.3)
    M["coated"] = create(yi, "material", "coated", type="coated_glossy", color=(0.8, 0.2, 0.2, 1.0),
                         diffuse_color=(0.6, 0.1, 0.1, 1.0), diffuse_reflect=0.5, glossy_reflect=0.6, exponent=80.0,
                         IOR=1.4, as_diffuse=False, mirror_color=(1.0, 1.0, 1.0, 1.0))
    M["glass"] = create(yi, "material", "glass", type="glass", IOR=1.5, filter_color=(0.8, 1.0, 0.9, 1.0), transmit_filter=0.6,
                        mirror_color=(1.0, 1.0, 1.0, 1.0), fake_shadows=True)
    M["glass2"] = create(yi, "material", "glass2", type="glass", IOR=1.33, filter_color=(1.0, 0.9, 0.8, 1.0), transmit_filter=0.4,
                         fake_shadows=False)
    M["mirror"] = create(yi, "material", "mirror", type="mirror", color=(0.9, 0.9, 1.0, 1.0), reflect=0.85)
    M["shiny"] = create(yi, "material", "shiny", type="shinydiffusemat", color=(0.3, 0.5, 0.9, 1.0), diffuse_reflect=0.7,
                        specular_reflect=0.3, transparency=0.4, transmit_filter=0.8, fresnel_effect=True, IOR=1.6,
                        mirror_color=(1.0, 0.95, 0.9, 1.0), diffuse_brdf="oren_nayar", sigma=0.25)
    for path, spec in (image_files or []):      # TGA files the image textures below name
        write_tga(path, test_image(spec["w"], spec["h"], spec.get("seed", 5)), top=spec.get("top", False), rle=spec.get("rle", False),
                  gray=spec.get("gray", False))
    for tname, tp in (textures or []):          # textures must exist before the texture_mapper nodes that name them
        create(yi, "texture", tname, **tp)
    for mname, mp in (extra_materials or []):   # e.g. rough_glass, blend_mat of two of the above, node materials
        M[mname] = create(yi, "material", mname, **mp)
    order = order or ["glossy", "aniso", "coated", "glass", "gdiff", "glass2", "mirror", "shiny"]
    v0, f0 = icosphere(subdiv)
    yi.startGeometry()
    add_mesh(yi, *quad((-9, -6, 0), (9, -6, 0), (9, 8, 0), (-9, 8, 0)), M["ground"], bulk=False)
    add_mesh(yi, *quad((-9, 8, 0), (9, 8, 0), (9, 8, 6), (-9, 8, 6)), M["ground"], bulk=False)
    for i, name in enumerate(order):
        cx = (i % 4 - 1.5) * 2.4
        cy = 0.0 if i < 4 else 3.2
        r = 1.0 if i < 4 else 0.9
        V = (v0 * r + np.array([cx, cy, r + 0.02])).astype(np.float32)
        if uv_orco:                             # orco = unit-sphere position, UV = longitude / latitude of the vertex
            uv = np.stack([np.arctan2(v0[:, 1], v0[:, 0]) / (2 * np.pi) + 0.5, np.arccos(np.clip(v0[:, 2], -1, 1)) / np.pi], 1)
            add_mesh_uv_orco(yi, V, f0, M[name], orco=v0.astype(np.float32), uv=uv.astype(np.float32),
                             normals=v0.astype(np.float32), smooth_angle=181.0)
        else:
            add_mesh(yi, V, f0, M[name], normals=v0.astype(np.float32), smooth_angle=181.0)
    for e in (emitters or []):              # visible emitter meshes for mesh lights: dict(id, centre, radius, subdiv, material params)
        M[e["name"]] = create(yi, "material", e["name"], **e["material"])
        ve, fe = icosphere(e.get("subdiv", 1))
        sc = np.asarray(e.get("scale", (1.0, 1.0, 1.0)), np.float32)
        add_mesh(yi, (ve * e["radius"] * sc + np.asarray(e["centre"])).astype(np.float32), fe, M[e["name"]], oid=e["id"])
    yi.endGeometry()
    if lights is None:
        create(yi, "light", "sun", type="sunlight", direction=(0.4, -0.5, 0.8), color=(1.0, 0.95, 0.9, 1.0), power=0.9,
               angle=1.0, samples=sun_samples)
        create(yi, "light", "pt", type="pointlight", **{"from": (-4.0, -3.0, 5.0)}, color=(0.9, 0.9, 1.0, 1.0), power=18.0)
        create(yi, "light", "area", type="arealight", corner=(2.0, -3.0, 5.0), point1=(2.0, -1.5, 5.0), point2=(3.5, -3.0, 5.0),
               color=(1.0, 0.85, 0.7, 1.0), power=20.0, samples=area_samples)
    else:                                   # explicit light set: list of (name, params)
        for lname, lp in lights:
            create(yi, "light", lname, **lp)
    cp = dict(type="perspective", focal=1.3)
    cp.update(cam_params or {})
    create(yi, "camera", "cam", **{"from": (0.0, -9.5, 4.2)}, to=(0.0, 1.0, 0.8),
           up=(0.0, -9.5, 5.2), resx=cam_res[0] if cam_res else width, resy=cam_res[1] if cam_res else height, **cp)
    create(yi, "background", "bg", **(background or dict(type="constant", color=(0.35, 0.45, 0.7, 1.0))))
    ip = dict(type=integrator, raydepth=raydepth)
    if integrator == "pathtracing":
        ip.update(bounces=bounces, path_samples=1)
    ip.update(integ_params or {})
    create(yi, "integrator", "default", **ip)
    std_volintegr(yi)
    set_render(yi, width, height, spp=spp, threads=threads, **render_extra)
    return len(order) * len(f0) + 4


# ----------------------------------------------------------------------------- instances + angle smoothing
def capped_cylinder(n=12, r=1.0, h=1.0):
    """Closed cylinder with shared vertices between side and caps: with smoothMesh(angle ~ 45) the side is smooth
    (360/n degrees between neighbouring faces) while the rim stays sharp (90 degrees) - both branches of the
    per-corner grouping of scene_t::smoothMesh (scene.cc:425-515) at every rim vertex."""
    a = np.arange(n) * (2 * np.pi / n)
    ring = np.stack([r * np.cos(a), r * np.sin(a)], 1)
    V = np.concatenate([np.c_[ring, np.zeros(n)], np.c_[ring, np.full(n, h)], [[0, 0, 0], [0, 0, h]]]).astype(np.float32)
    F = []
    for i in range(n):
        j = (i + 1) % n
        F += [(i, j, n + j), (i, n + j, n + i), (2 * n, j, i), (2 * n + 1, n + i, n + j)]
    return V, np.asarray(F, np.int32)


def _xform(scale=(1.0, 1.0, 1.0), rot_z=0.0, rot_x=0.0, shear_xy=0.0, translate=(0.0, 0.0, 0.0)):
    cz, sz, cx, sx = np.cos(rot_z), np.sin(rot_z), np.cos(rot_x), np.sin(rot_x)
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1.0]])
    Rx = np.array([[1.0, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Sh = np.array([[1.0, shear_xy, 0], [0, 1, 0], [0, 0, 1]])
    m = np.eye(4)
    m[:3, :3] = Rz @ Rx @ Sh @ np.diag(scale)
    m[:3, 3] = translate
    return m.astype(np.float32)


def instanced_scene(yi, width=96, height=54, spp=2, integrator="directlighting", bounces=3, threads=1, smooth_angle=45.0,
                    integ_params=None, **render_extra):
    """scene_t::addInstance (scene.cc:983-1008, triangleInstance_t) and angle-dependent scene_t::smoothMesh (scene.cc:425-515):
    an invisible BASEMESH icosphere with exported normals and an angle-smoothed cylinder, each instanced under rigid, uniformly
    scaled and non-uniformly scaled / sheared transforms, plus directly placed angle-smoothed cylinders."""
    yi.startScene(0)
    ground = create(yi, "material", "ground", type="shinydiffusemat", color=(0.6, 0.6, 0.55, 1.0), diffuse_reflect=0.9)
    red = create(yi, "material", "red", type="shinydiffusemat", color=(0.8, 0.25, 0.2, 1.0), diffuse_reflect=0.9)
    blue = create(yi, "material", "blue", type="glossy", color=(0.5, 0.6, 0.9, 1.0), diffuse_color=(0.2, 0.3, 0.6, 1.0),
                  diffuse_reflect=0.6, glossy_reflect=0.5, exponent=40.0, as_diffuse=True)
    shiny = create(yi, "material", "shiny", type="shinydiffusemat", color=(0.7, 0.7, 0.3, 1.0), diffuse_reflect=0.7,
                   specular_reflect=0.25, mirror_color=(1.0, 1.0, 1.0, 1.0))
    v0, f0 = icosphere(2)
    vc, fc = capped_cylinder(12, 0.8, 1.6)
    BASEMESH = 0x0200
    yi.startGeometry()
    add_mesh(yi, *quad((-8, -5, 0), (8, -5, 0), (8, 8, 0), (-8, 8, 0)), ground, bulk=False)
    # base objects: not rendered themselves (scene.cc:768)
    assert yi.startTriMesh(50, len(v0), len(f0), False, False, BASEMESH)
    for v in v0:
        yi.addVertex(float(v[0]), float(v[1]), float(v[2])); yi.addNormal(float(v[0]), float(v[1]), float(v[2]))
    for t in f0:
        assert yi.addTriangle(int(t[0]), int(t[1]), int(t[2]), red)
    assert yi.endTriMesh() and yi.smoothMesh(50, 181.0)
    assert yi.startTriMesh(51, len(vc), len(fc), False, False, BASEMESH)
    for v in vc:
        yi.addVertex(float(v[0]), float(v[1]), float(v[2]))
    for t in fc:
        assert yi.addTriangle(int(t[0]), int(t[1]), int(t[2]), blue)
    assert yi.endTriMesh() and yi.smoothMesh(51, smooth_angle)
    # directly placed cylinders: one angle-smoothed, one smoothed with a small angle (every corner keeps the flat normal)
    add_mesh(yi, vc * np.float32(0.9) + np.array([-4.5, 1.0, 0.01], np.float32), fc, shiny, smooth_angle=smooth_angle, oid=60, bulk=False)
    add_mesh(yi, vc * np.float32(0.7) + np.array([4.6, -1.5, 0.01], np.float32), fc, red, smooth_angle=5.0, oid=61, bulk=False)
    yi.endGeometry()
    for base, m in ((50, _xform(translate=(-2.2, 0.5, 1.0), rot_z=0.7, rot_x=0.4)),                       # rigid
                    (50, _xform(scale=(0.6, 0.6, 0.6), translate=(0.0, -2.0, 0.6))),                      # uniform scale
                    (50, _xform(scale=(1.6, 0.7, 0.45), rot_z=-0.5, translate=(2.4, 1.2, 0.45))),         # non-uniform scale
                    (51, _xform(rot_x=0.5, rot_z=1.1, translate=(0.2, 2.6, 0.6))),
                    (51, _xform(scale=(0.6, 1.4, 0.8), shear_xy=0.4, translate=(-2.0, -2.4, 0.013)))):    # sheared (cap not coplanar with the ground: t-ties)
        assert yi.addInstance(base, m)
    create(yi, "light", "sun", type="sunlight", direction=(0.35, -0.45, 0.8), color=(1.0, 0.95, 0.9, 1.0), power=1.0,
           angle=1.0, samples=2)
    create(yi, "light", "pt", type="pointlight", **{"from": (-3.0, -3.0, 5.0)}, color=(0.9, 0.9, 1.0, 1.0), power=25.0)
    create(yi, "camera", "cam", type="perspective", **{"from": (0.0, -9.0, 4.5)}, to=(0.0, 0.5, 0.6), up=(0.0, -9.0, 5.5),
           resx=width, resy=height, focal=1.25)
    create(yi, "background", "bg", type="constant", color=(0.4, 0.5, 0.7, 1.0))
    ip = dict(type=integrator, raydepth=4)
    if integrator == "pathtracing":
        ip.update(bounces=bounces, path_samples=1)
    ip.update(integ_params or {})
    create(yi, "integrator", "default", **ip)
    std_volintegr(yi)
    set_render(yi, width, height, spp=spp, threads=threads, **render_extra)
    return len(f0) * 3 + len(fc) * 4 + 2


# ----------------------------------------------------------------------------- C5: tessellated terrain
def _hash2(ix, iy, seed):
    """Integer lattice hash -> [0, 1) (uint32 arithmetic, identical on every machine)."""
    h = (ix.astype(np.uint32) * np.uint32(0x8da6b343)) ^ (iy.astype(np.uint32) * np.uint32(0xd8163841)) ^ np.uint32(seed * 0x9e3779b1 & 0xffffffff)
    h ^= h >> np.uint32(15); h *= np.uint32(0x2c1b3c6d)
    h ^= h >> np.uint32(12); h *= np.uint32(0x297a2d39)
    h ^= h >> np.uint32(15)
    return (h >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)


def value_noise(x, y, seed):
    x0 = np.floor(x); y0 = np.floor(y)
    fx = (x - x0).astype(np.float32); fy = (y - y0).astype(np.float32)
    ix = x0.astype(np.int64); iy = y0.astype(np.int64)
    sx = fx * fx * (3 - 2 * fx); sy = fy * fy * (3 - 2 * fy)
    a = _hash2(ix, iy, seed); b = _hash2(ix + 1, iy, seed); c = _hash2(ix, iy + 1, seed); d = _hash2(ix + 1, iy + 1, seed)
    return (a + (b - a) * sx) + ((c + (d - c) * sx) - (a + (b - a) * sx)) * sy


def terrain_heights(grid, extent=100.0, octaves=5, seed=3, amplitude=8.0):
    """Heights of a (grid+1)^2 vertex lattice over [-extent/2, extent/2]^2: fBm of value noise."""
    n = grid + 1
    u = np.linspace(0.0, 8.0, n, dtype=np.float32)
    z = np.zeros((n, n), np.float32)
    amp, freq = 0.5, 1.0
    for o in range(octaves):
        X, Y = np.meshgrid(u * np.float32(freq), u * np.float32(freq), indexing="xy")
        z += np.float32(amp) * value_noise(X, Y, seed + o)
        amp *= 0.5; freq *= 2.0
    return z * np.float32(amplitude)


def terrain_mesh(grid, extent=100.0, **kw):
    """(verts, tris, heights): 2*grid^2 triangles, cell (i,j) -> triangles 2*(j*grid+i) + {0,1}."""
    n = grid + 1
    z = terrain_heights(grid, extent, **kw)
    xs = np.linspace(-extent / 2, extent / 2, n, dtype=np.float32)
    V = np.empty((n, n, 3), np.float32)
    V[..., 0] = xs[None, :]; V[..., 1] = xs[:, None]; V[..., 2] = z
    i = np.arange(grid, dtype=np.int32)
    a = (i[None, :] + i[:, None] * n).reshape(-1)
    F = np.empty((grid * grid, 2, 3), np.int32)
    F[:, 0, 0] = a; F[:, 0, 1] = a + 1; F[:, 0, 2] = a + n + 1
    F[:, 1, 0] = a; F[:, 1, 1] = a + n + 1; F[:, 1, 2] = a + n
    return V.reshape(-1, 3), F.reshape(-1, 3), z


def terrain(yi, grid=5000, width=3840, height=2160, spp=1024, bounces=4, extent=100.0, seed=3, threads=1,
            integrator="pathtracing", cam_res=None, **render_extra):
    """Tessellated fBm terrain, 2*grid^2 triangles (grid=5000 -> 50 M), sun + constant background,
    pathtracing (BASELINE.json configs[4])."""
    V, F, _ = terrain_mesh(grid, extent, seed=seed)
    yi.startScene(0)
    mat = create(yi, "material", "soil", type="shinydiffusemat", color=(0.55, 0.5, 0.4, 1.0), diffuse_reflect=0.9)
    yi.startGeometry()
    add_mesh(yi, V, F, mat)
    yi.endGeometry()
    create(yi, "light", "sun", type="sunlight", direction=(0.4, -0.3, 0.6), color=(1.0, 0.96, 0.9, 1.0), power=1.5, angle=0.6, samples=2)
    e = extent
    create(yi, "camera", "cam", type="perspective", **{"from": (0.0, -0.62 * e, 0.35 * e)}, to=(0.0, 0.0, 0.02 * e),
           up=(0.0, -0.62 * e, 0.35 * e + 1.0), resx=cam_res[0] if cam_res else width, resy=cam_res[1] if cam_res else height, focal=1.0)
    create(yi, "background", "bg", type="constant", color=(0.55, 0.65, 0.85, 1.0))
    ip = dict(type=integrator, raydepth=3)
    if integrator == "pathtracing":
        ip.update(bounces=bounces, path_samples=1)
    create(yi, "integrator", "default", **ip)
    std_volintegr(yi)
    set_render(yi, width, height, spp=spp, threads=threads, **render_extra)
    return 2 * grid * grid
