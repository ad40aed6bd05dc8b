// yb200_interface.hpp — header-only C++ host side over the C ABI of libyb200.so.
//
// Mirrors the method set of the reference façade yafrayInterface_t
// (include/interface/yafrayinterface.h:46-144, src/interface/yafrayinterface.cc) so that an exporter, or the
// reference's own xml loader, can swap `getYafray()` for this class without touching its call sites: same method
// names, same argument meaning, same `bool`/NULL error behaviour.  The handles returned by create*() are borrowed
// tokens exactly like the reference's material_t* / light_t* (only ever passed back to addTriangle).
//
// The types below (colorOutput_t, progressBar_t, matrix4x4_t, memoryIO_t) restate the *interfaces* of
// include/core_api/output.h:31-39, include/yafraycore/monitor.h:24-36, include/core_api/matrix4.h and
// src/yafraycore/memoryIO.cc:27-43 so this header does not depend on the reference tree.  When building inside
// the reference tree define YB200_USE_REFERENCE_TYPES before including and the reference's own headers are used.
//
// Define YB200_DEFINE_GETYAFRAY in exactly one translation unit to emit `extern "C" getYafray()`
// (yafrayinterface.cc:386-392).
#pragma once
#include "yb200.h"
#include <string>
#include <vector>
#include <cstring>
#include <cstdio>

#ifdef YB200_USE_REFERENCE_TYPES
#include <core_api/output.h>
#include <core_api/matrix4.h>
#include <yafraycore/monitor.h>
namespace yb200 {
using yafaray::colorOutput_t; using yafaray::progressBar_t; using yafaray::matrix4x4_t;
using yafaray::material_t; using yafaray::light_t; using yafaray::texture_t; using yafaray::camera_t;
using yafaray::background_t; using yafaray::integrator_t; using yafaray::VolumeRegion; using yafaray::imageHandler_t;
}
#else
namespace yb200 {
// opaque token types: never dereferenced on either side of the interface
struct material_t; struct light_t; struct texture_t; struct camera_t; struct background_t; struct integrator_t;
struct VolumeRegion; struct imageHandler_t;

class colorOutput_t {                      // include/core_api/output.h:31-39
public:
    virtual ~colorOutput_t() {}
    virtual bool putPixel(int x, int y, const float *c, bool alpha = true, bool depth = false, float z = 0.f) = 0;
    virtual void flush() = 0;
    virtual void flushArea(int x0, int y0, int x1, int y1) = 0;
    virtual void highliteArea(int, int, int, int) {}
};

class progressBar_t {                      // include/yafraycore/monitor.h:24-36
public:
    virtual ~progressBar_t() {}
    virtual void init(int totalSteps) = 0;
    virtual void update(int steps = 1) = 0;
    virtual void done() = 0;
    virtual void setTag(const char *text) = 0;
};

struct matrix4x4_t {                       // row-major m[row][col], include/core_api/matrix4.h
    float m[4][4];
    matrix4x4_t() { std::memset(m, 0, sizeof(m)); m[0][0] = m[1][1] = m[2][2] = m[3][3] = 1.f; }
    explicit matrix4x4_t(const float *src) { std::memcpy(m, src, sizeof(m)); }
    float *operator[](int i) { return m[i]; }
    const float *operator[](int i) const { return m[i]; }
};
}  // namespace yb200
#endif

namespace yb200 {

// memoryIO_t (src/yafraycore/memoryIO.cc:27-43): writes into a caller-owned float RGBA buffer, (x + W*y)*4
class memoryIO_t : public colorOutput_t {
public:
    memoryIO_t(int resx, int resy, float *mem) : sizex(resx), sizey(resy), imageMem(mem) {}
    bool putPixel(int x, int y, const float *c, bool alpha = true, bool = false, float = 0.f) override {
        float *p = imageMem + (size_t)(x + sizex * y) * 4;
        p[0] = c[0]; p[1] = c[1]; p[2] = c[2]; p[3] = alpha ? c[3] : 1.f;
        return true;
    }
    void flush() override {}
    void flushArea(int, int, int, int) override {}
protected:
    int sizex, sizey; float *imageMem;
};

class yafrayInterface_t {
public:
    yafrayInterface_t() : ctx(yb200_create()), width(0), height(0), xstart(0), ystart(0), drawParams(false) {}
    virtual ~yafrayInterface_t() { yb200_destroy(ctx); }
    yafrayInterface_t(const yafrayInterface_t &) = delete;
    yafrayInterface_t &operator=(const yafrayInterface_t &) = delete;

    // ---- scene_t side (yafrayinterface.cc:66-144)
    virtual void loadPlugins(const char *) {}          // the supported plugin types are built in; nothing to load
    virtual bool startScene(int type = 0) { return yb200_start_scene(ctx, type) != 0; }
    virtual bool startGeometry() { return yb200_start_geometry(ctx) != 0; }
    virtual bool endGeometry() { return yb200_end_geometry(ctx) != 0; }
    virtual unsigned int getNextFreeID() { return yb200_get_next_free_id(ctx); }
    virtual bool startTriMesh(unsigned int id, int vertices, int triangles, bool hasOrco, bool hasUV = false, int type = 0) {
        return yb200_start_trimesh(ctx, id, vertices, triangles, hasOrco, hasUV, type) != 0;
    }
    virtual bool startTriMeshPtr(unsigned int *id, int vertices, int triangles, bool hasOrco, bool hasUV = false, int type = 0) {
        unsigned int nid = getNextFreeID();            // yafrayinterface.cc:91-96 hands a fresh id back
        if (!startTriMesh(nid, vertices, triangles, hasOrco, hasUV, type)) return false;
        if (id) *id = nid;
        return true;
    }
    virtual bool startCurveMesh(unsigned int, int) { return unsupported("startCurveMesh"); }
    virtual bool endCurveMesh(const material_t *, float, float, float) { return unsupported("endCurveMesh"); }
    virtual bool endTriMesh() { return yb200_end_trimesh(ctx) != 0; }
    virtual int addVertex(double x, double y, double z) { return yb200_add_vertex(ctx, x, y, z); }
    virtual int addVertex(double x, double y, double z, double ox, double oy, double oz) {
        return yb200_add_vertex_orco(ctx, x, y, z, ox, oy, oz);
    }
    virtual void addNormal(double nx, double ny, double nz) { yb200_add_normal(ctx, nx, ny, nz); }
    virtual bool addTriangle(int a, int b, int c, const material_t *mat) {
        return yb200_add_triangle(ctx, a, b, c, (yb200_handle)mat) != 0;
    }
    virtual bool addTriangle(int a, int b, int c, int uv_a, int uv_b, int uv_c, const material_t *mat) {
        return yb200_add_triangle_uv(ctx, a, b, c, uv_a, uv_b, uv_c, (yb200_handle)mat) != 0;
    }
    virtual int addUV(float u, float v) { return yb200_add_uv(ctx, u, v); }
    virtual bool smoothMesh(unsigned int id, double angle) { return yb200_smooth_mesh(ctx, id, angle) != 0; }
    virtual bool addInstance(unsigned int baseObjectId, matrix4x4_t objToWorld) {
        float m[16];
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) m[4 * i + j] = objToWorld[i][j];
        return yb200_add_instance(ctx, baseObjectId, m) != 0;
    }
    // bulk forms (new; same semantics as n single calls)
    bool addVertices(const float *xyz, const float *normalsOrNull, int n) { return yb200_add_vertices(ctx, xyz, normalsOrNull, n) != 0; }
    bool addTriangles(const int *abc, int n, const material_t *mat) { return yb200_add_triangles(ctx, abc, n, (yb200_handle)mat) != 0; }

    // ---- paraMap_t side (yafrayinterface.cc:146-267)
    virtual void paramsSetPoint(const char *name, double x, double y, double z) { yb200_params_set_point(ctx, name, x, y, z); }
    virtual void paramsSetString(const char *name, const char *s) { yb200_params_set_string(ctx, name, s); }
    virtual void paramsSetBool(const char *name, bool b) { yb200_params_set_bool(ctx, name, b); }
    virtual void paramsSetInt(const char *name, int i) {
        if (!std::strcmp(name, "width")) width = i; else if (!std::strcmp(name, "height")) height = i;
        else if (!std::strcmp(name, "xstart")) xstart = i; else if (!std::strcmp(name, "ystart")) ystart = i;
        yb200_params_set_int(ctx, name, i);
    }
    virtual void paramsSetFloat(const char *name, double f) { yb200_params_set_float(ctx, name, f); }
    virtual void paramsSetColor(const char *name, float r, float g, float b, float a = 1.f) { yb200_params_set_color(ctx, name, r, g, b, a); }
    virtual void paramsSetColor(const char *name, float *rgb, bool with_alpha = false) {
        yb200_params_set_color(ctx, name, rgb[0], rgb[1], rgb[2], with_alpha ? rgb[3] : 1.f);
    }
    virtual void paramsSetMatrix(const char *name, float m[4][4], bool transpose = false) { yb200_params_set_matrix(ctx, name, &m[0][0], transpose); }
    virtual void paramsSetMatrix(const char *name, double m[4][4], bool transpose = false) {
        float f[16]; for (int i = 0; i < 16; ++i) f[i] = (float)(&m[0][0])[i];
        yb200_params_set_matrix(ctx, name, f, transpose);
    }
    virtual void paramsSetMemMatrix(const char *name, float *matrix, bool transpose = false) { yb200_params_set_matrix(ctx, name, matrix, transpose); }
    virtual void paramsSetMemMatrix(const char *name, double *matrix, bool transpose = false) {
        float f[16]; for (int i = 0; i < 16; ++i) f[i] = (float)matrix[i];
        yb200_params_set_matrix(ctx, name, f, transpose);
    }
    virtual void paramsClearAll() { width = height = xstart = ystart = 0; yb200_params_clear_all(ctx); }
    virtual void paramsStartList() { yb200_params_start_list(ctx); }
    virtual void paramsPushList() { yb200_params_push_list(ctx); }
    virtual void paramsEndList() { yb200_params_end_list(ctx); }

    // ---- renderEnvironment_t side (yafrayinterface.cc:269-311); NULL on failure, message in lastError()
    virtual light_t *createLight(const char *name) { return (light_t *)yb200_create_light(ctx, name); }
    virtual texture_t *createTexture(const char *name) { return (texture_t *)yb200_create_texture(ctx, name); }
    virtual material_t *createMaterial(const char *name) { return (material_t *)yb200_create_material(ctx, name); }
    virtual camera_t *createCamera(const char *name) { return (camera_t *)yb200_create_camera(ctx, name); }
    virtual background_t *createBackground(const char *name) { return (background_t *)yb200_create_background(ctx, name); }
    virtual integrator_t *createIntegrator(const char *name) { return (integrator_t *)yb200_create_integrator(ctx, name); }
    virtual VolumeRegion *createVolumeRegion(const char *name) { return (VolumeRegion *)yb200_create_volume_region(ctx, name); }
    virtual imageHandler_t *createImageHandler(const char *, bool = true) { unsupported("createImageHandler"); return 0; }
    virtual unsigned int createObject(const char *) { unsupported("createObject"); return 0; }
    virtual void clearAll() { yb200_clear_all(ctx); image.clear(); }

    // ---- render (yafrayinterface.cc:330-346): on setup failure the reference returns silently; so does this,
    //      with the reason in lastError().  Pixels reach `output` through putPixel + one flush(), as
    //      imageFilm_t::flush does (imagefilm.cc:383-443); per-tile flushArea is folded into the final flush.
    virtual void render(colorOutput_t &output, progressBar_t *pb = 0) {
        if (width <= 0 || height <= 0) { unsupported("render: width/height not set"); return; }
        localErr.clear(); renderOk = false;
        image.assign((size_t)width * height * 4, 0.f);
        Progress st = { pb, 0, false };
        yb200_callbacks cb = { &st, 0, &yafrayInterface_t::onProgress };
        int ok = yb200_render(ctx, image.data(), width, height, &cb);
        if (pb) { if (!st.inited) pb->init(1); pb->done(); delete pb; }   // imageFilm_t owns and deletes the bar (imagefilm.cc:167-176)
        if (!ok) { image.clear(); return; }
        renderOk = true;
        getRenderedImage(output);
    }
    virtual bool getRenderedImage(colorOutput_t &output) {
        if (image.empty()) return false;
        for (int y = 0; y < height; ++y)
            for (int x = 0; x < width; ++x) output.putPixel(x, y, &image[(size_t)(x + width * y) * 4], true, false, 0.f);
        output.flush();
        return true;
    }
    virtual void abort() { yb200_abort(ctx); }
    virtual void setInputGamma(float gammaVal, bool enable) {
        // yafrayinterface.cc:348-358 linearises colours passed to paramsSetColor; with gamma 1 / disabled it is the
        // identity.  Anything else is not reproduced here, so refuse instead of silently rendering different colours.
        if (enable && gammaVal != 1.f) unsupported("setInputGamma != 1");
    }
    virtual void setVerbosityLevel(int vlevel) { yb200_set_verbosity(ctx, vlevel); }
    virtual void setVerbosityInfo() { yb200_set_verbosity(ctx, 3); }
    virtual void setVerbosityWarning() { yb200_set_verbosity(ctx, 2); }
    virtual void setVerbosityError() { yb200_set_verbosity(ctx, 1); }
    virtual void setVerbosityMute() { yb200_set_verbosity(ctx, 0); }
    virtual void setDrawParams(bool on = true) { drawParams = on; }
    virtual bool getDrawParams() { return drawParams; }
    virtual char *getVersion() const { return const_cast<char *>(yb200_version()); }
    virtual std::vector<std::string> listImageHandlers() { return std::vector<std::string>(); }
    virtual std::vector<std::string> listImageHandlersFullName() { return std::vector<std::string>(); }

    // ---- additions
    const char *lastError() const { return localErr.empty() ? yb200_last_error(ctx) : localErr.c_str(); }
    bool renderSucceeded() const { return renderOk; }   // render() is void like the reference's; this tells how it went
    bool setDevice(int cudaDevice) { return yb200_set_device(ctx, cudaDevice) != 0; }
    bool getStats(yb200_stats &s) { return yb200_get_stats(ctx, &s) != 0; }
    yb200_ctx *context() { return ctx; }

protected:
    struct Progress { progressBar_t *pb; int last; bool inited; };
    static void onProgress(void *user, int done, int total) {
        Progress *p = (Progress *)user;
        if (!p->pb) return;
        if (!p->inited) { p->pb->init(total); p->inited = true; }
        if (done > p->last) { p->pb->update(done - p->last); p->last = done; }
    }
    bool unsupported(const char *what) {
        localErr = std::string("yb200: unsupported through this interface: ") + what;
        std::fprintf(stderr, "ERROR: %s\n", localErr.c_str());
        return false;
    }
    std::string localErr;
    yb200_ctx *ctx;
    int width, height, xstart, ystart;
    bool drawParams;
    bool renderOk = false;
    std::vector<float> image;
};

typedef yafrayInterface_t *interfaceConstructor();

}  // namespace yb200

#ifdef YB200_DEFINE_GETYAFRAY
extern "C" yb200::yafrayInterface_t *getYafray() { return new yb200::yafrayInterface_t(); }
#endif
