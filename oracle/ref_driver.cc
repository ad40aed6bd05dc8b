// oracle/ref_driver.cc — TEST INFRASTRUCTURE ONLY.
//
// Exposes the UNMODIFIED reference (inferrna/Core; compiled from /root/reference by oracle/Makefile
// into oracle/_ref/) behind the same C entry points as include/yb200.h, with the prefix `yref_`,
// so that one scene-building script can drive either library and the outputs can be compared.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference arm load it.
// Nothing in the product (core_b200/, libyb200.so) links or calls this file.
//
// It calls yafrayInterface_t (src/interface/yafrayinterface.cc) for everything; ray batches go to
// triKdTree_t::Intersect via the same pre-processing as scene_t::intersect (scene.cc:852-879) and to
// scene_t::isShadowed (scene.cc:881-928).
#define protected public   // scene_t::tree / meshes and yafrayInterface_t::scene are protected
#define private public
#include <interface/yafrayinterface.h>
#include <core_api/scene.h>
#include <core_api/environment.h>
#include <core_api/params.h>
#include <core_api/matrix4.h>
#include <core_api/material.h>
#include <core_api/object3d.h>
#include <yafraycore/kdtree.h>
#include <yafraycore/memoryIO.h>
#include <yafraycore/timer.h>
#include <yafraycore/monitor.h>
#include <yafraycore/scr_halton.h>
#include <utilities/mcqmc.h>
#include <utilities/mathOptimizations.h>
#undef protected
#undef private
#include <map>
#include <vector>
#include <string>
#include <limits>
#include <cstring>
#include <cstdint>
#include <dlfcn.h>
#include <libgen.h>

using namespace yafaray;

struct silentBar_t : public yafaray::progressBar_t {
    void init(int) {} void update(int) {} void done() {} void setTag(const char *) {}
};

struct yref_ctx {
    std::vector<float> depth;      // z values the film handed to colorOutput_t::putPixel in the last render (z_channel)
    yafrayInterface_t *yi;
    std::string err;
    std::map<const triangle_t *, int> primIndex;
    std::vector<float> scratch;
    double render_s, kd_s;
};

static std::string plugin_dir()
{
    Dl_info info;
    dladdr((void *)&plugin_dir, &info);
    std::string p(info.dli_fname);
    size_t k = p.rfind('/');
    return (k == std::string::npos ? std::string(".") : p.substr(0, k)) + "/plugins";
}

extern "C" {

yref_ctx *yref_create(void)
{
    yref_ctx *c = new yref_ctx();
    c->yi = new yafrayInterface_t();
    c->yi->setVerbosityMute();
    c->yi->loadPlugins(plugin_dir().c_str());
    c->render_s = c->kd_s = 0;
    return c;
}
void yref_destroy(yref_ctx *c) { if (c) { delete c->yi; delete c; } }
const char *yref_last_error(yref_ctx *c) { return c->err.c_str(); }
const char *yref_version(void) { return "reference TheBounty 0.1.6 (oracle/_ref)"; }
int  yref_set_device(yref_ctx *, int) { return 1; }
void yref_set_verbosity(yref_ctx *c, int level) { c->yi->setVerbosityLevel(level); }
void yref_clear_all(yref_ctx *c) { c->yi->clearAll(); }

int yref_start_scene(yref_ctx *c, int type) { return c->yi->startScene(type); }
int yref_start_geometry(yref_ctx *c) { return c->yi->startGeometry(); }
int yref_end_geometry(yref_ctx *c) { return c->yi->endGeometry(); }
unsigned yref_get_next_free_id(yref_ctx *c) { return c->yi->getNextFreeID(); }
int yref_start_trimesh(yref_ctx *c, unsigned id, int v, int t, int orco, int uv, int type)
{ return c->yi->startTriMesh(id, v, t, orco != 0, uv != 0, type); }
int yref_end_trimesh(yref_ctx *c) { return c->yi->endTriMesh(); }
int yref_add_vertex(yref_ctx *c, double x, double y, double z) { return c->yi->addVertex(x, y, z); }
int yref_add_vertex_orco(yref_ctx *c, double x, double y, double z, double ox, double oy, double oz)
{ return c->yi->addVertex(x, y, z, ox, oy, oz); }
void yref_add_normal(yref_ctx *c, double x, double y, double z) { c->yi->addNormal(x, y, z); }
int yref_add_triangle(yref_ctx *c, int a, int b, int cc, uintptr_t mat)
{ return c->yi->addTriangle(a, b, cc, (const material_t *)mat); }
int yref_add_triangle_uv(yref_ctx *c, int a, int b, int cc, int ua, int ub, int uc, uintptr_t mat)
{ return c->yi->addTriangle(a, b, cc, ua, ub, uc, (const material_t *)mat); }
int yref_add_uv(yref_ctx *c, float u, float v) { return c->yi->addUV(u, v); }
int yref_smooth_mesh(yref_ctx *c, unsigned id, double angle) { return c->yi->smoothMesh(id, angle); }
int yref_add_instance(yref_ctx *c, unsigned base, const float m[16])
{
    float mm[4][4];
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) mm[i][j] = m[i * 4 + j];
    return c->yi->addInstance(base, matrix4x4_t(mm));
}
int yref_add_vertices(yref_ctx *c, const float *xyz, const float *nrm, int n)
{
    for (int i = 0; i < n; ++i) {
        if (c->yi->addVertex(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]) < 0) return 0;
        if (nrm) c->yi->addNormal(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]);
    }
    return 1;
}
int yref_add_triangles(yref_ctx *c, const int *abc, int n, uintptr_t mat)
{
    for (int i = 0; i < n; ++i)
        if (!c->yi->addTriangle(abc[3 * i], abc[3 * i + 1], abc[3 * i + 2], (const material_t *)mat)) return 0;
    return 1;
}

void yref_params_set_point(yref_ctx *c, const char *n, double x, double y, double z) { c->yi->paramsSetPoint(n, x, y, z); }
void yref_params_set_string(yref_ctx *c, const char *n, const char *s) { c->yi->paramsSetString(n, s); }
void yref_params_set_bool(yref_ctx *c, const char *n, int b) { c->yi->paramsSetBool(n, b != 0); }
void yref_params_set_int(yref_ctx *c, const char *n, int i) { c->yi->paramsSetInt(n, i); }
void yref_params_set_float(yref_ctx *c, const char *n, double f) { c->yi->paramsSetFloat(n, f); }
void yref_params_set_color(yref_ctx *c, const char *n, float r, float g, float b, float a) { c->yi->paramsSetColor(n, r, g, b, a); }
void yref_params_set_matrix(yref_ctx *c, const char *n, const float m[16], int transpose)
{ c->yi->paramsSetMemMatrix(n, const_cast<float *>(m), transpose != 0); }
void yref_params_clear_all(yref_ctx *c) { c->yi->paramsClearAll(); }
void yref_params_start_list(yref_ctx *c) { c->yi->paramsStartList(); }
void yref_params_push_list(yref_ctx *c) { c->yi->paramsPushList(); }
void yref_params_end_list(yref_ctx *c) { c->yi->paramsEndList(); }

uintptr_t yref_create_light(yref_ctx *c, const char *n) { return (uintptr_t)c->yi->createLight(n); }
uintptr_t yref_create_material(yref_ctx *c, const char *n) { return (uintptr_t)c->yi->createMaterial(n); }
uintptr_t yref_create_camera(yref_ctx *c, const char *n) { return (uintptr_t)c->yi->createCamera(n); }
uintptr_t yref_create_background(yref_ctx *c, const char *n) { return (uintptr_t)c->yi->createBackground(n); }
uintptr_t yref_create_integrator(yref_ctx *c, const char *n) { return (uintptr_t)c->yi->createIntegrator(n); }
uintptr_t yref_create_texture(yref_ctx *c, const char *n) { return (uintptr_t)c->yi->createTexture(n); }
uintptr_t yref_create_volume_region(yref_ctx *c, const char *n) { return (uintptr_t)c->yi->createVolumeRegion(n); }

struct yref_callbacks { void *user; int (*should_abort)(void *); void (*progress)(void *, int, int); };

// memoryIO_t (src/yafraycore/memoryIO.cc:27-43) drops the depth argument of putPixel; this output keeps it
struct depthIO_t : public colorOutput_t {
    int w, h; float *mem; std::vector<float> *z;
    depthIO_t(int w_, int h_, float *m, std::vector<float> *z_) : w(w_), h(h_), mem(m), z(z_) { z->assign((size_t)w * h, 0.f); }
    virtual bool putPixel(int x, int y, const float *c, bool alpha = true, bool depth = false, float zz = 0.f)
    {
        for (int i = 0; i < 4; ++i) mem[(x + w * y) * 4 + i] = (!alpha && i == 3) ? 1.f : c[i];
        if (depth) (*z)[(size_t)x + (size_t)w * y] = zz;
        return true;
    }
    virtual void flush() {}
    virtual void flushArea(int, int, int, int) {}
};
int yref_get_depth(yref_ctx *c, float *z)
{
    if (c->depth.empty()) { c->err = "no depth values (render with z_channel = true first)"; return 0; }
    memcpy(z, c->depth.data(), c->depth.size() * sizeof(float));
    return 1;
}

int yref_render(yref_ctx *c, float *rgba, int w, int h, const yref_callbacks *)
{
    depthIO_t out(w, h, rgba, &c->depth);
    memset(rgba, 0, sizeof(float) * 4 * (size_t)w * h);
    gTimer.addEvent("yref_render");
    gTimer.start("yref_render");
    // imageFilm_t takes ownership of the progress bar and deletes it (imagefilm.cc:175,637-641): heap-allocate
    c->yi->render(out, new silentBar_t());
    gTimer.stop("yref_render");
    c->render_s = gTimer.getTime("rendert");           // integrator.cc:146-167
    if (!c->yi->film) { c->err = "reference render failed (setupScene)"; return 0; }
    return 1;
}
// seconds spent inside tiledIntegrator_t::render of the last yref_render (gTimer "rendert")
double yref_last_render_seconds(yref_ctx *c) { return c->render_s; }
void yref_abort(yref_ctx *c) { c->yi->abort(); }

int yref_update_scene(yref_ctx *c)
{
    scene_t *scene = c->yi->scene;
    int w = 4, h = 4;
    c->yi->params->getParam("width", w);
    c->yi->params->getParam("height", h);
    c->scratch.assign((size_t)w * h * 4, 0.f);
    static memoryIO_t *dummy = 0;   // film keeps a pointer to the output; leak one per call deliberately
    dummy = new memoryIO_t(w, h, c->scratch.data());
    if (!c->yi->env->setupScene(*scene, *c->yi->params, *dummy, 0)) { c->err = "setupScene failed"; return 0; }
    if (!scene->update()) { c->err = "scene update failed"; return 0; }
    // global primitive order of scene.cc:760-781
    c->primIndex.clear();
    int idx = 0;
    for (std::map<objID_t, objData_t>::iterator i = scene->meshes.begin(); i != scene->meshes.end(); ++i) {
        objData_t &dat = i->second;
        if (!dat.obj->isVisible() || dat.obj->isBaseObject() || dat.type != TRIM) continue;
        int n = dat.obj->numPrimitives();
        std::vector<const triangle_t *> tmp(n);
        dat.obj->getPrimitives(tmp.data());
        for (int k = 0; k < n; ++k) c->primIndex[tmp[k]] = idx++;
    }
    return 1;
}

int yref_trace_batch(yref_ctx *c, const float *rays, size_t n, int mode, int shadow_depth, void *out, int)
{
    scene_t *scene = c->yi->scene;
    if (!scene || !scene->tree) { c->err = "no kd-tree (call update_scene)"; return 0; }
    renderState_t state;
    unsigned char udat[USER_DATA_SIZE + 8];
    state.userdata = udat;
    state.cam = scene->getCamera();
    for (size_t i = 0; i < n; ++i) {
        const float *r = rays + 8 * i;
        ray_t ray(point3d_t(r[0], r[1], r[2]), vector3d_t(r[4], r[5], r[6]), r[3], r[7]);
        if (mode == 0) {
            struct { int32_t prim; float t, b1, b2; } *o = (decltype(o))out + i;
            PFLOAT dis = ray.tmax < 0 ? std::numeric_limits<PFLOAT>::infinity() : ray.tmax, Z;
            intersectData_t data;
            triangle_t *hitt = 0;
            if (scene->tree->Intersect(ray, dis, &hitt, Z, data)) {
                o->prim = c->primIndex[hitt]; o->t = Z; o->b1 = data.b1; o->b2 = data.b2;
            } else { o->prim = -1; o->t = 0; o->b1 = o->b2 = 0; }
        } else if (mode == 1) {
            ((int32_t *)out)[i] = scene->isShadowed(state, ray) ? 1 : 0;
        } else {
            struct { int32_t s; float r, g, b; } *o = (decltype(o))out + i;
            color_t filt(1.f);
            o->s = scene->isShadowed(state, ray, shadow_depth, filt) ? 1 : 0;
            o->r = filt.R; o->g = filt.G; o->b = filt.B;
        }
    }
    return 1;
}

// ---- known-answer hooks into the reference's inline math (used to pin oracle/yb_oracle.c and the
//      device math library): include/utilities/mcqmc.h, include/yafraycore/scr_halton.h,
//      include/utilities/mathOptimizations.h
double yref_scr_halton(int dim, unsigned n) { return scrHalton(dim, n); }
float  yref_ri_vdc(unsigned b, unsigned r) { return RI_vdC(b, r); }
float  yref_ri_s(unsigned b, unsigned r) { return RI_S(b, r); }
float  yref_ri_lp(unsigned b, unsigned r) { return RI_LP(b, r); }
unsigned yref_fnv(unsigned v) { return fnv_32a_buf(v); }
void   yref_halton_seq(int base, unsigned start, int n, float *out)
{ Halton h(base); h.setStart(start); for (int i = 0; i < n; ++i) out[i] = h.getNext(); }
float yref_fsin(float x) { return fSin(x); }
float yref_fcos(float x) { return fCos(x); }
float yref_fpow(float a, float b) { return fPow(a, b); }
float yref_fexp(float a) { return fExp(a); }
float yref_flog(float a) { return fLog(a); }
float yref_fsqrt(float a) { return fSqrt(a); }
// BSDF known answers: initBSDF + eval + pdf + sample of a material created through the interface, at a surface point with the
// given shading / geometric normal (the record layout of yb200_selftest's YB200_KAT_BSDF table; word 0 is ignored here)
int yref_bsdf_kat(yref_ctx *c, const char *matName, const float *in, size_t n, float *out)
{
    const material_t *mat = c->yi->env->getMaterial(matName);
    if (!mat) { c->err = std::string("no material ") + matName; return 0; }
    for (size_t i = 0; i < n; ++i) {
        const float *f = in + 16 * i;
        float *o = out + 16 * i;
        surfacePoint_t sp;
        memset((void *)&sp, 0, sizeof(sp));
        sp.material = mat;
        sp.N = vector3d_t(f[1], f[2], f[3]); sp.Ng = vector3d_t(f[4], f[5], f[6]);
        sp.orcoNg = sp.Ng;
        createCS(sp.N, sp.NU, sp.NV);
        const vector3d_t wo(f[7], f[8], f[9]), wl(f[10], f[11], f[12]);
        uint32_t flags; memcpy(&flags, f + 15, 4);
        renderState_t state;
        unsigned char udat[USER_DATA_SIZE + 8];
        state.userdata = udat;
        state.includeLights = true;
        BSDF_t bsdfs = 0;
        mat->initBSDF(state, sp, bsdfs);
        const color_t ev = mat->eval(state, sp, wo, wl, (BSDF_t)flags);
        const float pdf = mat->pdf(state, sp, wo, wl, (BSDF_t)flags);
        sample_t s(f[13], f[14], (BSDF_t)flags);
        vector3d_t wi(0.f, 0.f, 0.f); float W = 0.f;
        const color_t sc = mat->sample(state, sp, wo, wi, s, W);
        o[0] = ev.R; o[1] = ev.G; o[2] = ev.B; o[3] = pdf;
        o[4] = wi.x; o[5] = wi.y; o[6] = wi.z; o[7] = sc.R; o[8] = sc.G; o[9] = sc.B; o[10] = W; o[11] = s.pdf;
        uint32_t sf = (uint32_t)s.sampledFlags, bf = (uint32_t)bsdfs;
        memcpy(o + 12, &sf, 4); memcpy(o + 13, &bf, 4); o[14] = 0.f; o[15] = 0.f;
    }
    return 1;
}
int   yref_faure(int dim, int *out) { int b = prims[dim]; for (int i = 0; i < b; ++i) out[i] = faure[dim][i]; return b; }

} // extern "C"
