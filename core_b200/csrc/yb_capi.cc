// yb_capi.cc — the C ABI of include/yb200.h over yb::Context.  Plain pointers and sizes only.
#include "../../include/yb200.h"
#include "yb_ctx.h"
#include "yb_trace.cuh"
#include <new>
#include <chrono>

using namespace yb;

#define CTX (ctx->c)
#define GUARD_BEGIN try {
#define GUARD_END(ret) } catch (const std::exception &e) { ctx->c.fail(e.what()); return ret; }

extern "C" {

yb200_ctx *yb200_create(void) { return new (std::nothrow) yb200_ctx(); }
void yb200_destroy(yb200_ctx *ctx) { delete ctx; }
const char *yb200_last_error(yb200_ctx *ctx) { return CTX.err.c_str(); }
const char *yb200_version(void)
{
#ifdef YB_HOSTEMU
    return "yb200 0.1 (HOST EMULATOR build - test infrastructure, not the product)";
#else
    return "yb200 0.1 (B200-native path for TheBounty 0.1.6 / YafaRay core)";
#endif
}
int yb200_set_device(yb200_ctx *ctx, int dev)
{
#ifndef YB_HOSTEMU
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || dev < 0 || dev >= n) return CTX.fail("set_device: no such CUDA device") ? 1 : 0;
    if (CTX.stream && CTX.device != dev) return CTX.fail("set_device: context already bound to another device") ? 1 : 0;
#endif
    CTX.device = dev;
    return 1;
}
int yb200_set_devices(yb200_ctx *ctx, const int *devices, int n)
{
#ifdef YB_HOSTEMU
    (void)devices;
    return n == 1 ? 1 : (CTX.fail("set_devices: several devices are not part of the host emulator") ? 1 : 0);
#else
    int have = 0;
    if (n < 1 || !devices) return CTX.fail("set_devices: empty device list") ? 1 : 0;
    if (cudaGetDeviceCount(&have) != cudaSuccess) return CTX.fail("set_devices: no CUDA device") ? 1 : 0;
    for (int i = 0; i < n; ++i) {
        if (devices[i] < 0 || devices[i] >= have) return CTX.fail("set_devices: no such CUDA device " + std::to_string(devices[i])) ? 1 : 0;
        for (int j = 0; j < i; ++j) if (devices[j] == devices[i]) return CTX.fail("set_devices: device listed twice") ? 1 : 0;
    }
    if (CTX.stream && CTX.device != devices[0]) return CTX.fail("set_devices: context already bound to another device (the first entry is the leader)") ? 1 : 0;
    CTX.device = devices[0];
    for (size_t i = 0; i < CTX.helpers.size(); ++i) delete CTX.helpers[i];
    CTX.helpers.clear();
    for (int i = 1; i < n; ++i) { Context *h = new Context(); h->device = devices[i]; CTX.helpers.push_back(h); }
    return 1;
#endif
}
void yb200_set_verbosity(yb200_ctx *ctx, int level) { CTX.verbosity = level; }
void yb200_clear_all(yb200_ctx *ctx) { CTX.clearAll(); }

int yb200_start_scene(yb200_ctx *ctx, int type) { return CTX.startScene(type); }
int yb200_start_geometry(yb200_ctx *ctx) { return CTX.startGeometry(); }
int yb200_end_geometry(yb200_ctx *ctx) { return CTX.endGeometry(); }
unsigned yb200_get_next_free_id(yb200_ctx *ctx) { return CTX.getNextFreeID(); }
int yb200_start_trimesh(yb200_ctx *ctx, unsigned id, int v, int t, int orco, int uv, int type) { return CTX.startTriMesh(id, v, t, orco != 0, uv != 0, type); }
int yb200_end_trimesh(yb200_ctx *ctx) { return CTX.endTriMesh(); }
int yb200_add_vertex(yb200_ctx *ctx, double x, double y, double z) { return CTX.addVertex(x, y, z); }
int yb200_add_vertex_orco(yb200_ctx *ctx, double x, double y, double z, double ox, double oy, double oz) { return CTX.addVertexOrco(x, y, z, ox, oy, oz); }
void yb200_add_normal(yb200_ctx *ctx, double x, double y, double z) { CTX.addNormal(x, y, z); }
int yb200_add_triangle(yb200_ctx *ctx, int a, int b, int c, yb200_handle mat) { return CTX.addTriangle(a, b, c, mat); }
int yb200_add_triangle_uv(yb200_ctx *ctx, int a, int b, int c, int ua, int ub, int uc, yb200_handle mat) { return CTX.addTriangleUV(a, b, c, ua, ub, uc, mat); }
int yb200_add_uv(yb200_ctx *ctx, float u, float v) { return CTX.addUV(u, v); }
int yb200_smooth_mesh(yb200_ctx *ctx, unsigned id, double angle) { return CTX.smoothMesh(id, angle); }
int yb200_add_instance(yb200_ctx *ctx, unsigned base, const float m[16]) { return CTX.addInstance(base, m); }
int yb200_add_vertices(yb200_ctx *ctx, const float *xyz, const float *nrm, int n)
{
    for (int i = 0; i < n; ++i) {
        if (CTX.addVertex(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]) < 0) return 0;
        if (nrm) CTX.addNormal(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]);
    }
    return 1;
}
int yb200_add_triangles(yb200_ctx *ctx, const int *abc, int n, yb200_handle mat)
{
    for (int i = 0; i < n; ++i)
        if (!CTX.addTriangle(abc[3 * i], abc[3 * i + 1], abc[3 * i + 2], mat)) return 0;
    return 1;
}

static Param &slot(yb200_ctx *ctx, const char *n) { return ctx->c.cparams->dicc[n]; }
void yb200_params_set_point(yb200_ctx *ctx, const char *n, double x, double y, double z)
{ Param &p = slot(ctx, n); p = Param(); p.type = Param::T_POINT; p.v[0] = (float)x; p.v[1] = (float)y; p.v[2] = (float)z; }
void yb200_params_set_string(yb200_ctx *ctx, const char *n, const char *s) { Param &p = slot(ctx, n); p = Param(); p.type = Param::T_STRING; p.sval = s; }
void yb200_params_set_bool(yb200_ctx *ctx, const char *n, int b) { Param &p = slot(ctx, n); p = Param(); p.type = Param::T_BOOL; p.bval = b != 0; }
void yb200_params_set_int(yb200_ctx *ctx, const char *n, int i) { Param &p = slot(ctx, n); p = Param(); p.type = Param::T_INT; p.ival = i; }
void yb200_params_set_float(yb200_ctx *ctx, const char *n, double f) { Param &p = slot(ctx, n); p = Param(); p.type = Param::T_FLOAT; p.fval = f; }
void yb200_params_set_color(yb200_ctx *ctx, const char *n, float r, float g, float b, float a)
{ Param &p = slot(ctx, n); p = Param(); p.type = Param::T_COLOR; p.v[0] = r; p.v[1] = g; p.v[2] = b; p.v[3] = a; }
void yb200_params_set_matrix(yb200_ctx *ctx, const char *n, const float m[16], int transpose)
{
    std::vector<float> mm(16);
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) mm[i * 4 + j] = transpose ? m[j * 4 + i] : m[i * 4 + j];
    ctx->c.cparams->mdicc[n] = mm;
}
void yb200_params_clear_all(yb200_ctx *ctx) { CTX.params.clear(); CTX.eparams.clear(); CTX.cparams = &CTX.params; }
void yb200_params_start_list(yb200_ctx *ctx) { if (!CTX.eparams.empty()) CTX.eparams.push_back(ParaMap()); else CTX.eparams.push_back(ParaMap()); CTX.cparams = &CTX.eparams.back(); }
void yb200_params_push_list(yb200_ctx *ctx) { CTX.eparams.push_back(ParaMap()); CTX.cparams = &CTX.eparams.back(); }
void yb200_params_end_list(yb200_ctx *ctx) { CTX.cparams = &CTX.params; }

yb200_handle yb200_create_light(yb200_ctx *ctx, const char *n) { return CTX.createLight(n); }
yb200_handle yb200_create_material(yb200_ctx *ctx, const char *n) { return CTX.createMaterial(n); }
yb200_handle yb200_create_camera(yb200_ctx *ctx, const char *n) { return CTX.createCamera(n); }
yb200_handle yb200_create_background(yb200_ctx *ctx, const char *n) { return CTX.createBackground(n); }
yb200_handle yb200_create_integrator(yb200_ctx *ctx, const char *n) { return CTX.createIntegrator(n); }
int yb200_register_image(yb200_ctx *ctx, const char *file_name, const float *rgba, int width, int height, int is_hdr)
{
    GUARD_BEGIN
    return CTX.registerImage(file_name, rgba, width, height, is_hdr != 0) ? 1 : 0;
    GUARD_END(0)
}
yb200_handle yb200_create_texture(yb200_ctx *ctx, const char *n) { GUARD_BEGIN return CTX.createTexture(n); GUARD_END(0) }
yb200_handle yb200_create_volume_region(yb200_ctx *ctx, const char *n) { CTX.fail(std::string("createVolumeRegion('") + n + "'): participating media are not supported by the B200 path"); return 0; }

int yb200_update_scene(yb200_ctx *ctx)
{
    GUARD_BEGIN
    return CTX.updateScene() ? 1 : 0;
    GUARD_END(0)
}

void yb200_touch_geometry(yb200_ctx *ctx) { CTX.geomChanged = true; }

int yb200_render_film(yb200_ctx *ctx, float *dev_film, int rank, int world, const yb200_callbacks *cb)
{
    GUARD_BEGIN
    CTX.abortFlag = 0;
    if (!CTX.updateScene()) return 0;
    return render_film_device(CTX, dev_film, 0, rank, world, -1, cb ? cb->should_abort : 0, cb ? cb->progress : 0, cb ? cb->user : 0) ? 1 : 0;
    GUARD_END(0)
}
int yb200_render_film_pass(yb200_ctx *ctx, float *dev_film, const float *dev_film_all, int rank, int world, int pass, const yb200_callbacks *cb)
{
    GUARD_BEGIN
    if (pass == 0) CTX.abortFlag = 0;
    if (!CTX.updateScene()) return 0;
    return render_film_device(CTX, dev_film, dev_film_all, rank, world, pass, cb ? cb->should_abort : 0, cb ? cb->progress : 0, cb ? cb->user : 0) ? 1 : 0;
    GUARD_END(0)
}
int yb200_film_resolve(yb200_ctx *ctx, const float *dev_film, float *rgba, int on_device)
{
    GUARD_BEGIN
    if (!CTX.sceneReady) return CTX.fail("film_resolve: scene not updated") ? 1 : 0;
    const size_t npx = (size_t)CTX.sc.film.w * CTX.sc.film.h;
    if (on_device) { film_resolve_device(CTX, dev_film, rgba); rt_sync(CTX.stream); return 1; }
    ctx->io.rgba.reserve(npx * 4);
    film_resolve_device(CTX, dev_film, ctx->io.rgba.p);
    rt_d2h(rgba, ctx->io.rgba.p, npx * 4 * sizeof(float), CTX.stream);
    return 1;
    GUARD_END(0)
}
int yb200_render(yb200_ctx *ctx, float *rgba, int width, int height, const yb200_callbacks *cb)
{
    GUARD_BEGIN
    CTX.abortFlag = 0;
    if (!CTX.sceneStarted) return CTX.fail("render: no scene") ? 1 : 0;
    if (!CTX.setupRender(width, height)) return 0;
    if (!CTX.updateScene()) return 0;
    if (!CTX.helpers.empty())       // yb200_set_devices: tiles over all devices, films summed on the leader over NVLink
        return render_group(CTX, rgba, cb ? cb->should_abort : 0, cb ? cb->progress : 0, cb ? cb->user : 0) ? 1 : 0;
    const size_t npx = (size_t)CTX.sc.film.w * CTX.sc.film.h;
    ctx->io.film.reserve(npx * 5);
    rt_memset(ctx->io.film.p, 0, npx * 5 * sizeof(float), CTX.stream);
    if (!render_film_device(CTX, ctx->io.film.p, 0, 0, 1, -1, cb ? cb->should_abort : 0, cb ? cb->progress : 0, cb ? cb->user : 0)) return 0;
    return yb200_film_resolve(ctx, ctx->io.film.p, rgba, 0);
    GUARD_END(0)
}
float *yb200_depth_film(yb200_ctx *ctx) { return depth_film_device(CTX); }
int yb200_depth_resolve(yb200_ctx *ctx, const float *dev_depth_film, float *z)
{
    GUARD_BEGIN
    return depth_resolve_device(CTX, dev_depth_film, z) ? 1 : 0;
    GUARD_END(0)
}
int yb200_get_depth(yb200_ctx *ctx, float *z) { return yb200_depth_resolve(ctx, depth_film_device(CTX), z); }
void yb200_abort(yb200_ctx *ctx) { CTX.abortFlag = 1; for (size_t i = 0; i < CTX.helpers.size(); ++i) CTX.helpers[i]->abortFlag = 1; }

int yb200_trace_batch(yb200_ctx *ctx, const float *rays, size_t n, int mode, int shadow_depth, void *out, int on_device)
{
    GUARD_BEGIN
    if (!CTX.sceneReady) return CTX.fail("trace_batch: call update_scene first") ? 1 : 0;
    if (mode < 0 || mode > 2) return CTX.fail("trace_batch: bad mode") ? 1 : 0;
    if (n == 0) return 1;
    const size_t osz = mode == YB200_TRACE_SHADOW ? 4 : 16;
    auto t0 = std::chrono::steady_clock::now();
    if (on_device) trace_batch_device(CTX, rays, n, mode, shadow_depth, out);
    else {
        ctx->io.in.reserve(n * 32); ctx->io.out.reserve(n * osz);
        rt_h2d(ctx->io.in.p, rays, n * 32, CTX.stream);
        trace_batch_device(CTX, (const float *)ctx->io.in.p, n, mode, shadow_depth, ctx->io.out.p);
        rt_d2h(out, ctx->io.out.p, n * osz, CTX.stream);
    }
    (void)t0;
    if (mode == YB200_TRACE_CLOSEST) CTX.stats.rays_closest = n; else CTX.stats.rays_shadow = n;
    return 1;
    GUARD_END(0)
}

int yb200_get_stats(yb200_ctx *ctx, yb200_stats *o)
{
    const Stats &s = CTX.stats;
    o->n_triangles = s.n_triangles; o->n_bvh_nodes = s.n_bvh_nodes; o->rays_closest = s.rays_closest; o->rays_shadow = s.rays_shadow;
    o->camera_samples = s.camera_samples; o->kernel_launches = s.kernel_launches;
    o->build_ms = s.build_ms; o->render_ms = s.render_ms; o->trace_closest_ms = s.trace_closest_ms; o->trace_shadow_ms = s.trace_shadow_ms;
    o->shade_ms = s.shade_ms; o->film_ms = s.film_ms; o->bvh_bytes = s.bvh_bytes; o->tri_bytes = s.tri_bytes;
    return 1;
}
void *yb200_cuda_stream(yb200_ctx *ctx) { return (void *)CTX.stream; }

}  // extern "C"
