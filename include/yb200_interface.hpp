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
#include <core_api/params.h>
#include <yafraycore/monitor.h>
namespace yb200 {
using yafaray::colorOutput_t; using yafaray::progressBar_t; using yafaray::matrix4x4_t; using yafaray::paraMap_t;
using yafaray::material_t; using yafaray::light_t; using yafaray::texture_t; using yafaray::camera_t;
using yafaray::background_t; using yafaray::integrator_t; using yafaray::VolumeRegion; using yafaray::imageHandler_t;
}
#else
namespace yb200 {
// opaque token types: never dereferenced on either side of the interface
struct material_t; struct light_t; struct texture_t; struct camera_t; struct background_t; struct integrator_t;
struct VolumeRegion; struct imageHandler_t; struct paraMap_t;

class colorOutput_t {                      // include/core_api/output.h:31-39 (same virtual order)
public:
    virtual ~colorOutput_t() {}
    virtual bool putPixel(int x, int y, const float *c, bool alpha = true, bool depth = false, float z = 0.f) = 0;
    virtual void flush() = 0;
    virtual void flushArea(int x0, int y0, int x1, int y1) = 0;
    virtual void highliteArea(int, int, int, int) {}
};

class progressBar_t {                      // include/yafraycore/monitor.h:24-36 (same virtual order)
public:
    virtual ~progressBar_t() {}
    virtual void init(int totalSteps) = 0;
    virtual void update(int steps = 1) = 0;
    virtual void done() = 0;
    virtual void setTag(const char *text) = 0;
};

// include/core_api/matrix4.h:33-69: row-major matrix[row][col] followed by the `_invalid` flag.  The reference class has a
// user-provided copy constructor and destructor, so the C++ ABI passes it BY INVISIBLE REFERENCE (addInstance takes it by
// value): this restatement keeps both, and the member layout, so that a caller compiled against either header passes the
// same thing.
class matrix4x4_t {
public:
    matrix4x4_t() : _invalid(0) { std::memset(matrix, 0, sizeof(matrix)); matrix[0][0] = matrix[1][1] = matrix[2][2] = matrix[3][3] = 1.f; }
    explicit matrix4x4_t(const float *src) : _invalid(0) { std::memcpy(matrix, src, sizeof(matrix)); }
    matrix4x4_t(const matrix4x4_t &o) : _invalid(o._invalid) { std::memcpy(matrix, o.matrix, sizeof(matrix)); }
    matrix4x4_t &operator=(const matrix4x4_t &o) { std::memcpy(matrix, o.matrix, sizeof(matrix)); _invalid = o._invalid; return *this; }
    ~matrix4x4_t() {}
    float *operator[](int i) { return matrix[i]; }
    const float *operator[](int i) const { return matrix[i]; }
protected:
    float matrix[4][4];
    int _invalid;
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

// The virtual functions below are declared in EXACTLY the order of include/interface/yafrayinterface.h:47-129 (destructor first),
// and none is added in between: the vtable of this class is laid out like the reference's, so an exporter or binding compiled
// against the reference's header that obtains the object from getYafray() of libyb200_yafray.so dispatches to the right
// methods (checked by tests/test_reference_header_client.py with a client compiled against the reference header only).
// Anything new is non-virtual and comes after.
class yafrayInterface_t {
public:
    yafrayInterface_t() : ctx(yb200_create()), width(0), height(0), xstart(0), ystart(0), tileSize(32), zChannel(false), drawParams(false) {}
    virtual ~yafrayInterface_t() { yb200_destroy(ctx); }                                                    // :48
    // ---- scene_t side (yafrayinterface.cc:66-144)
    virtual void loadPlugins(const char *) {}          // :50  the supported plugin types are built in; nothing to load
    virtual bool startGeometry() { return yb200_start_geometry(ctx) != 0; }                                 // :51
    virtual bool endGeometry() { return yb200_end_geometry(ctx) != 0; }                                     // :52
    virtual unsigned int getNextFreeID() { return yb200_get_next_free_id(ctx); }                            // :53
    virtual bool startTriMesh(unsigned int id, int vertices, int triangles, bool hasOrco, bool hasUV = false, int type = 0) {   // :59
        return yb200_start_trimesh(ctx, id, vertices, triangles, hasOrco, hasUV, type) != 0;
    }
    virtual bool startCurveMesh(unsigned int, int) { return unsupported("startCurveMesh"); }                // :62
    virtual bool startTriMeshPtr(unsigned int *id, int vertices, int triangles, bool hasOrco, bool hasUV = false, int type = 0) {   // :63
        unsigned int nid = getNextFreeID();            // yafrayinterface.cc:91-96 hands a fresh id back
        if (!startTriMesh(nid, vertices, triangles, hasOrco, hasUV, type)) return false;
        if (id) *id = nid;
        return true;
    }
    virtual bool endTriMesh() { return yb200_end_trimesh(ctx) != 0; }                                       // :64
    virtual bool endCurveMesh(const material_t *, float, float, float) { return unsupported("endCurveMesh"); }   // :65
    virtual int addVertex(double x, double y, double z) { return yb200_add_vertex(ctx, x, y, z); }          // :66
    virtual int addVertex(double x, double y, double z, double ox, double oy, double oz) {                  // :67
        return yb200_add_vertex_orco(ctx, x, y, z, ox, oy, oz);
    }
    virtual void addNormal(double nx, double ny, double nz) { yb200_add_normal(ctx, nx, ny, nz); }          // :68
    virtual bool addTriangle(int a, int b, int c, const material_t *mat) {                                  // :70
        return yb200_add_triangle(ctx, a, b, c, (yb200_handle)mat) != 0;
    }
    virtual bool addTriangle(int a, int b, int c, int uv_a, int uv_b, int uv_c, const material_t *mat) {    // :72
        return yb200_add_triangle_uv(ctx, a, b, c, uv_a, uv_b, uv_c, (yb200_handle)mat) != 0;
    }
    virtual int addUV(float u, float v) { return yb200_add_uv(ctx, u, v); }                                 // :74
    virtual bool smoothMesh(unsigned int id, double angle) { return yb200_smooth_mesh(ctx, id, angle) != 0; }   // :76
    virtual bool addInstance(unsigned int baseObjectId, matrix4x4_t objToWorld) {                           // :77
        float m[16];
        for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) m[4 * i + j] = objToWorld[i][j];
        return yb200_add_instance(ctx, baseObjectId, m) != 0;
    }
    // ---- paraMap_t side (yafrayinterface.cc:146-267)
    virtual void paramsSetPoint(const char *name, double x, double y, double z) { yb200_params_set_point(ctx, name, x, y, z); }   // :80
    virtual void paramsSetString(const char *name, const char *s) { yb200_params_set_string(ctx, name, s); }                     // :81
    virtual void paramsSetBool(const char *name, bool b) {                                                                       // :82
        if (!std::strcmp(name, "z_channel")) zChannel = b;
        yb200_params_set_bool(ctx, name, b);
    }
    virtual void paramsSetInt(const char *name, int i) {                                                                         // :83
        if (!std::strcmp(name, "width")) width = i; else if (!std::strcmp(name, "height")) height = i;
        else if (!std::strcmp(name, "xstart")) xstart = i; else if (!std::strcmp(name, "ystart")) ystart = i;
        else if (!std::strcmp(name, "tile_size")) tileSize = i > 0 ? i : 32;
        yb200_params_set_int(ctx, name, i);
    }
    virtual void paramsSetFloat(const char *name, double f) { yb200_params_set_float(ctx, name, f); }                            // :84
    virtual void paramsSetColor(const char *name, float r, float g, float b, float a = 1.f) { yb200_params_set_color(ctx, name, r, g, b, a); }   // :85
    virtual void paramsSetColor(const char *name, float *rgb, bool with_alpha = false) {                                         // :86
        yb200_params_set_color(ctx, name, rgb[0], rgb[1], rgb[2], with_alpha ? rgb[3] : 1.f);
    }
    virtual void paramsSetMatrix(const char *name, float m[4][4], bool transpose = false) { yb200_params_set_matrix(ctx, name, &m[0][0], transpose); }   // :87
    virtual void paramsSetMatrix(const char *name, double m[4][4], bool transpose = false) {                                     // :88
        float f[16]; for (int i = 0; i < 16; ++i) f[i] = (float)(&m[0][0])[i];
        yb200_params_set_matrix(ctx, name, f, transpose);
    }
    virtual void paramsSetMemMatrix(const char *name, float *matrix, bool transpose = false) { yb200_params_set_matrix(ctx, name, matrix, transpose); }   // :89
    virtual void paramsSetMemMatrix(const char *name, double *matrix, bool transpose = false) {                                  // :90
        float f[16]; for (int i = 0; i < 16; ++i) f[i] = (float)matrix[i];
        yb200_params_set_matrix(ctx, name, f, transpose);
    }
    virtual void paramsClearAll() { width = height = xstart = ystart = 0; tileSize = 32; zChannel = false; yb200_params_clear_all(ctx); }   // :91
    virtual void paramsStartList() { yb200_params_start_list(ctx); }                                                             // :92
    virtual void paramsPushList() { yb200_params_push_list(ctx); }                                                               // :93
    virtual void paramsEndList() { yb200_params_end_list(ctx); }                                                                 // :94
    // ---- renderEnvironment_t side (yafrayinterface.cc:269-311); NULL on failure, message in lastError()
    virtual light_t *createLight(const char *name) { return (light_t *)yb200_create_light(ctx, name); }                          // :96
    virtual texture_t *createTexture(const char *name) { return (texture_t *)yb200_create_texture(ctx, name); }                  // :97
    virtual material_t *createMaterial(const char *name) { return (material_t *)yb200_create_material(ctx, name); }              // :98
    virtual camera_t *createCamera(const char *name) { return (camera_t *)yb200_create_camera(ctx, name); }                      // :99
    virtual background_t *createBackground(const char *name) { return (background_t *)yb200_create_background(ctx, name); }      // :100
    virtual integrator_t *createIntegrator(const char *name) { return (integrator_t *)yb200_create_integrator(ctx, name); }      // :101
    virtual VolumeRegion *createVolumeRegion(const char *name) { return (VolumeRegion *)yb200_create_volume_region(ctx, name); } // :102
    virtual imageHandler_t *createImageHandler(const char *, bool = true) { unsupported("createImageHandler"); return 0; }       // :104
    virtual unsigned int createObject(const char *) { unsupported("createObject"); return 0; }                                   // :105
    virtual void clearAll() { yb200_clear_all(ctx); image.clear(); depth.clear(); }                                              // :106
    // ---- render (yafrayinterface.cc:330-346): on setup failure the reference returns silently; so does this, with the reason
    //      in lastError().  The frame reaches `output` the way imageFilm_t delivers it: tile by tile (imageSpliter_t order,
    //      putPixel per pixel + flushArea per tile, imagefilm.cc:339-381), then the whole frame once more + flush()
    //      (imagefilm.cc:383-443); with z_channel every pixel arrives as putPixel(x, y, col, true, depth = true, z)
    //      (imagefilm.cc:360-362,425-427), and a putPixel that returns false stops the delivery (abort, imagefilm.cc:362-366).
    virtual void render(colorOutput_t &output, progressBar_t *pb = 0) {                                                          // :107
        if (width <= 0 || height <= 0) { unsupported("render: width/height not set"); if (pb) delete pb; return; }
        localErr.clear(); renderOk = false;
        image.assign((size_t)width * height * 4, 0.f);
        depth.clear();
        Progress st = { pb, 0, false };
        yb200_callbacks cb = { &st, 0, &yafrayInterface_t::onProgress };
        int ok = yb200_render(ctx, image.data(), width, height, &cb);
        if (ok && zChannel) { depth.assign((size_t)width * height, 0.f); ok = yb200_get_depth(ctx, depth.data()); }
        if (pb) { if (!st.inited) pb->init(1); pb->done(); delete pb; }   // imageFilm_t owns and deletes the bar (imagefilm.cc:167-176)
        if (!ok) { image.clear(); depth.clear(); return; }
        renderOk = true;
        bool go = true;
        for (int y0 = 0; y0 < height && go; y0 += tileSize)
            for (int x0 = 0; x0 < width && go; x0 += tileSize) {
                const int x1 = x0 + tileSize < width ? x0 + tileSize : width, y1 = y0 + tileSize < height ? y0 + tileSize : height;
                go = putArea(output, x0, y0, x1, y1);
                if (go) output.flushArea(x0, y0, x1, y1);
            }
        if (go) getRenderedImage(output);
    }
    virtual bool startScene(int type = 0) { return yb200_start_scene(ctx, type) != 0; }                                          // :108
    virtual void setInputGamma(float gammaVal, bool enable) {                                                                    // :109
        // yafrayinterface.cc:348-358 linearises colours passed to paramsSetColor; with gamma 1 / disabled it is the
        // identity.  Anything else is not reproduced here, so refuse instead of silently rendering different colours.
        if (enable && gammaVal != 1.f) unsupported("setInputGamma != 1");
    }
    virtual void abort() { yb200_abort(ctx); }                                                                                   // :110
    // :111 the reference hands out its paraMap_t; the parameters live behind the C ABI here, there is no such object (only the
    // reference's Qt GUI reads it, src/gui/mywindow.cc:178)
    virtual paraMap_t *getRenderParameters() { return 0; }
    virtual bool getRenderedImage(colorOutput_t &output) {                                                                       // :113
        if (image.empty()) return false;
        if (!putArea(output, 0, 0, width, height)) return false;
        output.flush();
        return true;
    }
    virtual std::vector<std::string> listImageHandlers() { return std::vector<std::string>(); }                                  // :114
    virtual std::vector<std::string> listImageHandlersFullName() { return std::vector<std::string>(); }                          // :115
    virtual std::string getImageFormatFromFullName(const std::string &) { return std::string(); }                                // :116  no image handlers
    virtual std::string getImageFullNameFromFormat(const std::string &) { return std::string(); }                                // :117
    virtual void setVerbosityLevel(int vlevel) { yb200_set_verbosity(ctx, vlevel); }                                             // :119
    virtual void setVerbosityInfo() { yb200_set_verbosity(ctx, 3); }                                                             // :120
    virtual void setVerbosityWarning() { yb200_set_verbosity(ctx, 2); }                                                          // :121
    virtual void setVerbosityError() { yb200_set_verbosity(ctx, 1); }                                                            // :122
    virtual void setVerbosityMute() { yb200_set_verbosity(ctx, 0); }                                                             // :123
    virtual void setDrawParams(bool on = true) { drawParams = on; }                                                              // :125
    virtual bool getDrawParams() { return drawParams; }                                                                          // :126
    virtual char *getVersion() const { return const_cast<char *>(yb200_version()); }                                             // :128

    // ---- additions (non-virtual: they must not move the slots above)
    yafrayInterface_t(const yafrayInterface_t &) = delete;
    yafrayInterface_t &operator=(const yafrayInterface_t &) = delete;
    bool addVertices(const float *xyz, const float *normalsOrNull, int n) { return yb200_add_vertices(ctx, xyz, normalsOrNull, n) != 0; }
    bool addTriangles(const int *abc, int n, const material_t *mat) { return yb200_add_triangles(ctx, abc, n, (yb200_handle)mat) != 0; }
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
    // pixels of [x0, x1) x [y0, y1) to the output; false as soon as the output refuses one
    bool putArea(colorOutput_t &output, int x0, int y0, int x1, int y1) {
        for (int y = y0; y < y1; ++y)
            for (int x = x0; x < x1; ++x) {
                const float *c = &image[(size_t)(x + width * y) * 4];
                const bool ok = depth.empty() ? output.putPixel(x, y, c, true, false, 0.f)
                                              : output.putPixel(x, y, c, true, true, depth[(size_t)x + (size_t)width * y]);
                if (!ok) return false;
            }
        return true;
    }
    bool unsupported(const char *what) {
        localErr = std::string("yb200: unsupported through this interface: ") + what;
        std::fprintf(stderr, "ERROR: %s\n", localErr.c_str());
        return false;
    }
    std::string localErr;
    yb200_ctx *ctx;
    int width, height, xstart, ystart, tileSize;
    bool zChannel, drawParams;
    bool renderOk = false;
    std::vector<float> image, depth;
};

typedef yafrayInterface_t *interfaceConstructor();

}  // namespace yb200

#ifdef YB200_DEFINE_GETYAFRAY
extern "C" yb200::yafrayInterface_t *getYafray() { return new yb200::yafrayInterface_t(); }
#endif
