// yb_scene.h — host-side state behind the C ABI: parameter maps, the factories of renderEnvironment_t for the
// supported plugin types, the scene_t geometry state machine, the flattener to SoA device buffers, and the
// entry points of the GPU stages (BVH build, ray batches, wavefront render).
#pragma once
#include "yb_rt.h"
#include "yb_types.h"
#include <map>
#include <list>
#include <vector>
#include <string>

namespace yb {

// ---- paraMap_t / parameter_t (include/core_api/params.h:40-125): typed values, typed (strict) reads
struct Param {
    enum Type { T_NONE = -1, T_INT = 1, T_BOOL, T_FLOAT, T_STRING, T_POINT, T_COLOR };
    int type;
    int ival; bool bval; double fval; std::string sval; float v[4];
    Param() : type(T_NONE), ival(0), bval(false), fval(0) { v[0] = v[1] = v[2] = v[3] = 0; }
};
struct ParaMap {
    std::map<std::string, Param> dicc;
    std::map<std::string, std::vector<float> > mdicc;   // 16 floats row-major
    bool get(const std::string &n, int &i) const;
    bool get(const std::string &n, bool &b) const;
    bool get(const std::string &n, float &f) const;
    bool get(const std::string &n, double &f) const;
    bool get(const std::string &n, std::string &s) const;
    bool getPoint(const std::string &n, v3 &p) const;
    bool getColor(const std::string &n, col3 &c) const;
    bool has(const std::string &n) const { return dicc.count(n) != 0; }
    void clear() { dicc.clear(); mdicc.clear(); }
};

// ---- geometry as ingested (scene.cc:266-319,522-662): one record per mesh object
struct MeshObj {
    std::vector<v3> points;          // with has_orco: interleaved (p, orco) like the reference (scene.cc:550-552)
    std::vector<v3> normals;
    std::vector<int32_t> tri_idx;    // 3 per triangle: pa, pb, pc (already doubled for orco)
    std::vector<int32_t> tri_nidx;   // 3 per triangle: na, nb, nc (-1 = none)
    std::vector<uint32_t> tri_mat;   // material index per triangle
    std::vector<float> uv_values;    // uv_t pairs (scene.cc:648-662)
    std::vector<int32_t> uv_offsets; // 3 per triangle (scene.cc:628-646)
    bool has_orco, has_uv, is_smooth, normals_exported, visible, is_base;
    int n_uv_offsets;
    // instance (scene.cc:983-1008): base != 0 -> triangles come from *base transformed by objToWorld
    const MeshObj *base;
    float m[16];
    unsigned lightId;                // light bound through `object` (arealight.cc:51-58); informational
    MeshObj() : has_orco(false), has_uv(false), is_smooth(false), normals_exported(false), visible(true), is_base(false),
                n_uv_offsets(0), base(0), lightId(0) {}
    size_t numTris() const { return base ? base->tri_mat.size() : tri_mat.size(); }
};

struct IntegratorDesc { bool surface; bool volume_none; Integrator p; };
struct CameraDesc { Camera cam; };
struct BackgroundDesc { col3 color; int gradient; col3 gzenith, ghoriz, szenith, shoriz; };

struct Stats {
    uint64_t n_triangles, n_bvh_nodes, rays_closest, rays_shadow, camera_samples, kernel_launches;
    double build_ms, render_ms, trace_closest_ms, trace_shadow_ms, shade_ms, film_ms;
    uint64_t bvh_bytes, tri_bytes;
    uint64_t caustic_photons;         // photons stored in the caustic map of the last render
};

struct RenderBuffers;   // wavefront queues (yb_render.cu)
struct PhotonBuffers;   // caustic photon map (yb_photon.cu)

// Scratch of the BVH build, kept across builds so that a rebuild (animation, the bench's host-to-host step) never
// goes through cudaMalloc / cudaFree, which serialise with the whole device.
struct BvhScratch {
    DevBuf<int32_t> ob, left, right, cnt, parent, ref, ref2, nn;
    DevBuf<uint64_t> keys, keys2;
    DevBuf<uint32_t> idx, idx2, flag, pos, nodeCounter, counters;
    DevBuf<v3> nlo, nhi, lo, hi, lo2, hi2;
    DevBuf<char> sortTmp, scanTmp;
    DevBuf<uint64_t> qa, qb;          // collapse work items (2 x int32 each)
};

struct RaySortScratch { DevBuf<uint32_t> keys, keys2, idx, idx2; DevBuf<char> tmp; };

struct Context {
    std::string err;
    int verbosity;
    int device;
    rt_stream stream;
    bool streamOwned;
    // parameter state (yafrayinterface.cc:146-267)
    ParaMap params; std::list<ParaMap> eparams; ParaMap *cparams;
    // environment tables (environment.cc:189-479)
    std::vector<Material> materials; std::map<std::string, int> materialNames;
    std::vector<Texture> textures; std::map<std::string, int> textureNames;
    // image textures: decoded pixels (RGBA float, row 0 on top) by file name - registered through the C ABI or read from a TGA file;
    // texPixels = pool of the pixels of all image textures, texPixOffset[i] = first pixel of texture i (or npos)
    struct HostImage { int w, h; bool hdr; std::vector<f4> px; };
    std::map<std::string, HostImage> images;
    std::vector<f4> texPixels; std::vector<size_t> texPixOffset;
    std::vector<ShaderNode> shaderNodes;
    std::vector<Light> lights; std::map<std::string, int> lightNames; std::vector<unsigned> lightObj;
    std::map<std::string, CameraDesc> cameras;
    std::map<std::string, BackgroundDesc> backgrounds;
    std::map<std::string, IntegratorDesc> integrators;
    // scene state machine (scene.cc:53,95-135)
    enum { ST_READY = 0, ST_GEOMETRY, ST_OBJECT };
    std::vector<int> stateStack;
    bool sceneStarted; int sceneMode;
    std::map<unsigned, MeshObj> meshes;
    MeshObj *curObj; unsigned curId; int lastVertId; bool curOrco;
    unsigned nextFreeID;
    bool geomChanged;
    volatile int abortFlag;
    // device-side scene
    DevBuf<BvhNode> d_nodes; DevBuf<f4> d_tri_v; DevBuf<v3> d_tri_n; DevBuf<Material> d_materials; DevBuf<Light> d_lights;
    DevBuf<f4> d_tex_pixels;
    DevBuf<Texture> d_textures; DevBuf<ShaderNode> d_shaderNodes; DevBuf<float> d_prim_uv; DevBuf<v3> d_prim_orco, d_prim_ng;
    DevBuf<f4> d_light_tris; DevBuf<float> d_light_cdf;   // mesh light emitters (yb_host.cc uploadTables)
    DevBuf<uint32_t> d_traceCounter;   // ray-fetch counter of the persistent traversal kernels
    DevBuf<f4> d_in_v; DevBuf<v3> d_in_n;             // flattened geometry in scene order (input of the build)
    HostBuf<f4> h_in_v; HostBuf<v3> h_in_n;           // its pinned host staging
    BvhScratch bvhScratch;
    RaySortScratch raySort;
    uint32_t nTris, nNodes, bvhDepth; bool haveNormals;
    v3 boundLo, boundHi;
    SceneDev sc;                     // filled by updateScene / setupRender
    bool sceneReady;
    RenderBuffers *rb;
    PhotonBuffers *photonBufs;
    Stats stats;
    // ---- several GPUs behind ONE context (yb200_set_devices): this context is the leader on `device`; every further device
    //      has a helper context that owns the device-side copy of the scene there.  The leader flattens, uploads and builds
    //      once; helpers receive nodes / triangles / normals by peer copy over NVLink and only upload the small tables.
    std::vector<Context *> helpers;
    const std::map<unsigned, MeshObj> *meshSrc;   // helper: the leader's meshes (mesh-light emitters are read from there)
    DevBuf<float> groupFilm, groupFilmAll, groupTmp;   // helper: its film; leader: sum of all films / staging for the peer copies
    bool adoptScene(Context &leader);              // helper: host tables + device geometry of the leader -> this device

    Context();
    ~Context();
    bool fail(const std::string &msg);           // records + logs a Y_ERROR-style message, returns false
    void info(const std::string &msg);
    void warn(const std::string &msg);
    void clearAll();
    // scene
    bool startScene(int type);
    bool startGeometry(); bool endGeometry();
    unsigned getNextFreeID();
    bool startTriMesh(unsigned id, int vertices, int triangles, bool hasOrco, bool hasUV, int type);
    bool endTriMesh();
    int addVertex(double x, double y, double z);
    int addVertexOrco(double x, double y, double z, double ox, double oy, double oz);
    void addNormal(double x, double y, double z);
    bool addTriangle(int a, int b, int c, uintptr_t mat);
    bool addTriangleUV(int a, int b, int c, int ua, int ub, int uc, uintptr_t mat);
    int addUV(float u, float v);
    bool smoothMesh(unsigned id, double angle);
    bool addInstance(unsigned baseId, const float m[16]);
    // factories
    uintptr_t createLight(const char *name);
    uintptr_t createMaterial(const char *name);
    uintptr_t createTexture(const char *name);
    bool registerImage(const char *fileName, const float *rgba, int w, int h, bool hdr);
    bool loadTga(const std::string &fileName, HostImage &img);
    bool loadMaterialNodes(const char *name, Material &m, const char *const *slotKeys, const int *slotIds, int nSlots);
    uintptr_t createCamera(const char *name);
    uintptr_t createBackground(const char *name);
    uintptr_t createIntegrator(const char *name);
    // render
    bool setupRender(int width, int height);    // renderEnvironment_t::setupScene (environment.cc:596-694) + createImageFilm
    bool updateScene();                          // scene_t::update (scene.cc:748-850): flatten + BVH build
    bool flattenAndBuild();
    bool uploadTables();
    bool ensureDevice();
};

// handles: material handle = index + 1 tagged; see yb_host.cc
static const uintptr_t HANDLE_MAT_TAG = 0x4d000000u;

// ---- GPU stages
void bvh_build(const f4 *in_v, const v3 *in_n, uint32_t n, DevBuf<BvhNode> &nodes, DevBuf<f4> &out_v, DevBuf<v3> &out_n,
               uint32_t &nNodes, uint32_t &maxDepth, v3 &boundLo, v3 &boundHi, BvhScratch &w, rt_stream s);
// coherence sort of a wave of rays (yb_bvh.cu): sorted ray indices on the device, or 0
const uint32_t *sort_rays(const char *rays, uint32_t stride, uint32_t n, v3 lo, v3 hi, RaySortScratch &w, rt_stream s);
// rays/out are DEVICE pointers; mode = TRACE_*; out layout as documented in include/yb200.h
void trace_batch_device(Context &c, const float *d_rays, size_t n, int mode, int shadow_depth, void *d_out);
// accumulate the tiles of (rank, world) of one frame into d_film (w*h*5 floats, raw sums)
// accumulate the tiles of (rank, world) into d_film.  onlyPass < 0: all AA passes (single rank only when AA_passes > 1);
// onlyPass >= 0: just that pass, with d_filmAll = raw film sums of ALL ranks after the previous pass (0: use d_film)
bool render_film_device(Context &c, float *d_film, const float *d_filmAll, int rank, int world, int onlyPass,
                        int (*should_abort)(void *), void (*progress)(void *, int, int), void *user);
void film_resolve_device(Context &c, const float *d_film, float *d_rgba);
// z_channel: this rank's raw depth film (w*h*2 floats, device; 0 when off) and its normalisation into a HOST buffer of w*h floats
float *depth_film_device(Context &c);
bool depth_resolve_device(Context &c, const float *d_depthFilm, float *z_host);
void render_release(RenderBuffers *rb);
// mcIntegrator_t::createCausticMap: shoot the caustic photons of the scene's lights and sort them into the gather grid
// (c.sc.photonGrid); a no-op unless the integrator asks for photon caustics.  Runs once per frame before the first wave.
bool build_caustic_map(Context &c);
void photon_release(PhotonBuffers *p);
struct RenderArgs;
void launch_caustic_gather(uint32_t n, const RenderArgs &a, rt_stream s);   // estimateCausticPhotons for the integrate() tasks of a wave
// yafrayInterface_t::render over all devices of the context (leader + helpers): tiles dealt round-robin over the devices, films
// summed on the leader's device by peer copies, resolved into the HOST buffer rgba (w*h*4 floats)
bool render_group(Context &c, float *rgba_host, int (*should_abort)(void *), void (*progress)(void *, int, int), void *user);
void film_add_device(float *d_dst, const float *d_src, size_t n, rt_stream s);

}  // namespace yb
