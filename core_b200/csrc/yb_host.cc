// yb_host.cc — host side of the drop-in boundary: parameter maps, the scene_t geometry state machine
// (src/yafraycore/scene.cc:95-135,266-319,362-662,950-1008), the factories of the supported plugin types
// (renderEnvironment_t::create*, src/yafraycore/environment.cc:189-479, and each plugin's factory()),
// render setup (environment.cc:481-531,596-694) and the flattener that turns meshes into device SoA buffers
// in the global primitive order of scene_t::update (scene.cc:760-781).
// Anything unsupported fails loudly (Y_ERROR-style message + failed call); there is no CPU render path here.
#include "yb_scene.h"
#include <stdio.h>
#include <math.h>
#include <limits.h>
#include <algorithm>
#include <chrono>
#include <thread>
#include <atomic>

namespace yb {

// ------------------------------------------------------------------------------------------------ ParaMap
#define YB_GET(T, field, TYPE)                                                       \
    bool ParaMap::get(const std::string &n, T &o) const                              \
    {                                                                                \
        std::map<std::string, Param>::const_iterator i = dicc.find(n);               \
        if (i == dicc.end() || i->second.type != Param::TYPE) return false;          \
        o = (T)i->second.field;                                                      \
        return true;                                                                 \
    }
YB_GET(int, ival, T_INT)
YB_GET(bool, bval, T_BOOL)
YB_GET(float, fval, T_FLOAT)
YB_GET(double, fval, T_FLOAT)
bool ParaMap::get(const std::string &n, std::string &s) const
{
    std::map<std::string, Param>::const_iterator i = dicc.find(n);
    if (i == dicc.end() || i->second.type != Param::T_STRING) return false;
    s = i->second.sval;
    return true;
}
bool ParaMap::getPoint(const std::string &n, v3 &p) const
{
    std::map<std::string, Param>::const_iterator i = dicc.find(n);
    if (i == dicc.end() || i->second.type != Param::T_POINT) return false;
    p = v3(i->second.v[0], i->second.v[1], i->second.v[2]);
    return true;
}
bool ParaMap::getColor(const std::string &n, col3 &c) const
{
    std::map<std::string, Param>::const_iterator i = dicc.find(n);
    if (i == dicc.end() || i->second.type != Param::T_COLOR) return false;
    c = col3(i->second.v[0], i->second.v[1], i->second.v[2]);
    return true;
}

// ------------------------------------------------------------------------------------------------ Context
Context::Context()
    : verbosity(2), device(-1), stream(0), streamOwned(false), cparams(&params), sceneStarted(false), sceneMode(0),
      curObj(0), curId(0), lastVertId(0), curOrco(false), nextFreeID(INT_MAX), geomChanged(true), abortFlag(0),
      nTris(0), nNodes(0), bvhDepth(0), haveNormals(false), sceneReady(false), rb(0), photonBufs(0), meshSrc(0)
{
    memset(&sc, 0, sizeof(sc));
    memset(&stats, 0, sizeof(stats));
    stateStack.push_back(ST_READY);
}
Context::~Context()
{
    for (size_t i = 0; i < helpers.size(); ++i) delete helpers[i];
#ifndef YB_HOSTEMU
    int callerDevice = -1;
    cudaGetDevice(&callerDevice);
    if (device >= 0) cudaSetDevice(device);        // the buffers of this context are freed with the device that owns them current
#endif
    render_release(rb); rb = 0;
    photon_release(photonBufs); photonBufs = 0;
#ifndef YB_HOSTEMU
    if (streamOwned && stream) cudaStreamDestroy(stream);
    d_nodes.release(); d_tri_v.release(); d_tri_n.release(); d_materials.release(); d_lights.release(); d_tex_pixels.release();
    d_textures.release(); d_shaderNodes.release(); d_prim_uv.release(); d_prim_orco.release(); d_prim_ng.release();
    d_light_tris.release(); d_light_cdf.release(); d_traceCounter.release(); d_in_v.release(); d_in_n.release();
    groupFilm.release(); groupFilmAll.release(); groupTmp.release();
    if (callerDevice >= 0) cudaSetDevice(callerDevice);      // leave the caller's current device as it was
#endif
}
bool Context::fail(const std::string &msg)
{
    err = msg;
    if (verbosity >= 1) fprintf(stderr, "ERROR: yb200: %s\n", msg.c_str());
    return false;
}
void Context::warn(const std::string &msg) { if (verbosity >= 2) fprintf(stderr, "WARNING: yb200: %s\n", msg.c_str()); }
void Context::info(const std::string &msg) { if (verbosity >= 3) fprintf(stderr, "INFO: yb200: %s\n", msg.c_str()); }

void Context::clearAll()   // yafrayinterface.cc:47-64
{
    params.clear(); eparams.clear(); cparams = &params;
    materials.clear(); materialNames.clear(); textures.clear(); textureNames.clear(); texPixels.clear(); texPixOffset.clear(); shaderNodes.clear(); lights.clear(); lightNames.clear(); lightObj.clear();
    cameras.clear(); backgrounds.clear(); integrators.clear();
    meshes.clear(); curObj = 0; stateStack.clear(); stateStack.push_back(ST_READY);
    nextFreeID = INT_MAX; geomChanged = true; sceneReady = false; sceneStarted = false;
}

// ------------------------------------------------------------------------------------------------ scene state machine
bool Context::startScene(int type)   // yafrayinterface.cc:66-73
{
    if (type != 0) return fail("startScene: universal scene mode (type 1) is not supported (triangle mode only)");
    meshes.clear(); curObj = 0; stateStack.clear(); stateStack.push_back(ST_READY);
    nextFreeID = INT_MAX; geomChanged = true; sceneReady = false;
    sceneStarted = true; sceneMode = 0;
    return true;
}
bool Context::startGeometry()
{
    if (stateStack.back() != ST_READY) return false;
    stateStack.push_back(ST_GEOMETRY);
    return true;
}
bool Context::endGeometry()
{
    if (stateStack.back() != ST_GEOMETRY) return false;
    stateStack.pop_back();
    return true;
}
unsigned Context::getNextFreeID()   // scene.cc:950-966
{
    unsigned id = nextFreeID;
    while (meshes.find(id) != meshes.end()) { --nextFreeID; id = nextFreeID; }
    --nextFreeID;
    return id;
}
bool Context::startTriMesh(unsigned id, int vertices, int triangles, bool hasOrco, bool hasUV, int type)   // scene.cc:266-294
{
    if (stateStack.back() != ST_GEOMETRY) return false;
    int ptype = type & 0xFF;
    if (ptype != 0) { fail("startTriMesh: only TRIM (type 0) meshes are supported (VTRIM/MTRIM need universal mode)"); return false; }
    MeshObj &o = meshes[id];
    o = MeshObj();
    o.has_orco = hasOrco; o.has_uv = hasUV;
    o.visible = !(type & 0x0100);     // INVISIBLEM (scene.h:44-46)
    o.is_base = (type & 0x0200) != 0; // BASEMESH
    if (vertices > 0) o.points.reserve((size_t)vertices * (hasOrco ? 2 : 1));
    if (triangles > 0) { o.tri_idx.reserve(3 * (size_t)triangles); o.tri_nidx.reserve(3 * (size_t)triangles); o.tri_mat.reserve(triangles); }
    stateStack.push_back(ST_OBJECT);
    geomChanged = true;
    curOrco = hasOrco; curObj = &o; curId = id; lastVertId = 0;
    return true;
}
bool Context::endTriMesh()   // scene.cc:296-319
{
    if (stateStack.back() != ST_OBJECT) return false;
    if (curObj->has_uv && (size_t)curObj->n_uv_offsets != 3 * curObj->tri_mat.size()) { fail("Scene: UV-offsets mismatch!"); return false; }
    stateStack.pop_back();
    return true;
}
int Context::addVertex(double x, double y, double z)   // scene.cc:522-542
{
    if (stateStack.back() != ST_OBJECT) return -1;
    curObj->points.push_back(v3((float)x, (float)y, (float)z));
    lastVertId = (int)curObj->points.size() - 1;
    return lastVertId;
}
int Context::addVertexOrco(double x, double y, double z, double ox, double oy, double oz)   // scene.cc:544-566
{
    if (stateStack.back() != ST_OBJECT) return -1;
    curObj->points.push_back(v3((float)x, (float)y, (float)z));
    curObj->points.push_back(v3((float)ox, (float)oy, (float)oz));
    lastVertId = ((int)curObj->points.size() - 1) / 2;
    return lastVertId;
}
void Context::addNormal(double x, double y, double z)   // scene.cc:568-583
{
    if (!curObj) return;
    MeshObj &o = *curObj;
    if (o.points.size() > (size_t)lastVertId && o.points.size() > o.normals.size()) {
        if (o.normals.size() < o.points.size()) o.normals.resize(o.points.size(), v3(0.f));
        o.normals[lastVertId] = v3((float)x, (float)y, (float)z);
        o.normals_exported = true;
    }
}
bool Context::addTriangle(int a, int b, int c, uintptr_t mat)   // scene.cc:585-626
{
    if (stateStack.back() != ST_OBJECT) return false;
    if ((mat & 0xff000000u) != HANDLE_MAT_TAG || (mat & 0xffffffu) == 0 || (mat & 0xffffffu) > materials.size()) {
        fail("addTriangle: invalid material handle");
        return false;
    }
    MeshObj &o = *curObj;
    if (curOrco) { a *= 2; b *= 2; c *= 2; }
    o.tri_idx.push_back(a); o.tri_idx.push_back(b); o.tri_idx.push_back(c);
    if (o.normals_exported) {
        if (curOrco) { o.tri_nidx.push_back(a >> 1); o.tri_nidx.push_back(b >> 1); o.tri_nidx.push_back(c >> 1); }
        else { o.tri_nidx.push_back(a); o.tri_nidx.push_back(b); o.tri_nidx.push_back(c); }
    } else { o.tri_nidx.push_back(-1); o.tri_nidx.push_back(-1); o.tri_nidx.push_back(-1); }
    o.tri_mat.push_back((uint32_t)((mat & 0xffffffu) - 1));
    return true;
}
bool Context::addTriangleUV(int a, int b, int c, int uv_a, int uv_b, int uv_c, uintptr_t mat)   // scene.cc:628-646
{
    if (!addTriangle(a, b, c, mat)) return false;
    curObj->n_uv_offsets += 3;
    curObj->uv_offsets.push_back(uv_a); curObj->uv_offsets.push_back(uv_b); curObj->uv_offsets.push_back(uv_c);
    return true;
}
int Context::addUV(float u, float v)   // scene.cc:648-662
{
    if (stateStack.back() != ST_OBJECT) return 0;
    curObj->uv_values.push_back(u); curObj->uv_values.push_back(v);
    return (int)(curObj->uv_values.size() / 2) - 1;
}
bool Context::addInstance(unsigned baseId, const float m[16])   // scene.cc:983-1008
{
    std::map<unsigned, MeshObj>::iterator it = meshes.find(baseId);
    if (it == meshes.end()) { fail("Base mesh for instance doesn't exist"); return false; }
    unsigned id = getNextFreeID();
    if ((int)id <= 0) return false;
    MeshObj &o = meshes[id];
    o = MeshObj();
    o.base = &meshes[baseId];
    memcpy(o.m, m, sizeof(o.m));
    o.has_orco = o.base->has_orco; o.has_uv = o.base->has_uv; o.is_smooth = o.base->is_smooth;
    o.normals_exported = o.base->normals_exported;
    geomChanged = true;
    return true;
}

// ------------------------------------------------------------------------------------------------ smoothMesh
static inline v3 tri_rec_normal(const MeshObj &o, size_t t)   // triangle_t::recNormal (triangle_inline.h:100-107)
{
    const v3 &a = o.points[o.tri_idx[3 * t]], &b = o.points[o.tri_idx[3 * t + 1]], &c = o.points[o.tri_idx[3 * t + 2]];
    return normalize(cross(b - a, c - a));
}
static inline float sinFromVectors(v3 a, v3 b)   // vector3d.h:263-268
{
    float div = (length(a) * length(b)) * 0.99999f + 0.00001f;
    float ret = (length(cross(a, b)) / div) * 0.99999f;
    if (ret > 1.f) ret = 1.f;
    if (ret < -1.f) ret = -1.f;
    return asinf(ret);
}
// scene_t::smoothMesh (scene.cc:362-520), restated over a corner table.
//
// A "corner" is (triangle t, slot c) -> vertex tri_idx[3t+c]; corner id = 3t+c.  The corners are grouped by vertex with a
// counting sort (ascending corner id inside a group - the order in which the reference meets the faces of a vertex, which
// fixes the float summation order and therefore the normals bit for bit).  Per corner the interior angle weight is
// asin(|e1 x e2| / |e1||e2|) (vector3d.h:263-268).
//   angle >= 180 : one normal per vertex = normalised sum of weight * face normal over the vertex' corners;
//   0.1 < angle  : one normal per smoothing group: a corner takes its own face plus every other face at the vertex whose
//                  normal is within `angle` (dot > fCos(angle)); corners without such a neighbour keep the flat normal
//                  (index -1); groups of one vertex whose normals agree to dot > 0.999 share the first one created.
bool Context::smoothMesh(unsigned id, double angle_)
{
    if (stateStack.back() != ST_GEOMETRY) return false;
    MeshObj *o = curObj;
    if (id) {
        std::map<unsigned, MeshObj>::iterator it = meshes.find(id);
        if (it == meshes.end()) return false;
        o = &it->second;
    }
    if (!o || o->base) return false;
    if (o->normals_exported && o->points.size() == o->normals.size()) { o->is_smooth = true; return true; }
    const float angle = (float)angle_;
    const size_t nPoints = o->points.size(), nCorners = 3 * o->tri_mat.size();
    o->normals.resize(nPoints, v3(0.f));
    if (angle > 0.1f) {
        for (size_t k = 0; k < nCorners; ++k)
            if (o->tri_idx[k] < 0 || (size_t)o->tri_idx[k] >= nPoints) { fail("smoothMesh: triangle references a vertex that does not exist"); return false; }
        // corner table: cornersOf[first[v] .. first[v + 1]) = the corners at vertex v
        std::vector<uint32_t> first(nPoints + 1, 0u), cornersOf(nCorners);
        for (size_t k = 0; k < nCorners; ++k) first[(size_t)o->tri_idx[k] + 1]++;
        for (size_t v = 0; v < nPoints; ++v) first[v + 1] += first[v];
        {
            std::vector<uint32_t> fill(first.begin(), first.end() - 1);
            for (size_t k = 0; k < nCorners; ++k) cornersOf[fill[(size_t)o->tri_idx[k]]++] = (uint32_t)k;
        }
        std::vector<float> weight(nCorners);
        std::vector<v3> faceN(nCorners / 3);
        for (size_t t = 0; t < nCorners / 3; ++t) {
            faceN[t] = tri_rec_normal(*o, t);
            const v3 P[3] = {o->points[o->tri_idx[3 * t]], o->points[o->tri_idx[3 * t + 1]], o->points[o->tri_idx[3 * t + 2]]};
            // edges in the reference's order: from the corner to the remaining two vertices, lower slot first
            weight[3 * t + 0] = sinFromVectors(P[1] - P[0], P[2] - P[0]);
            weight[3 * t + 1] = sinFromVectors(P[0] - P[1], P[2] - P[1]);
            weight[3 * t + 2] = sinFromVectors(P[0] - P[2], P[1] - P[2]);
        }
        if (angle >= 180.f) {
            for (size_t v = 0; v < nPoints; ++v) {
                v3 sum = o->normals[v];
                for (uint32_t q = first[v]; q < first[v + 1]; ++q) sum += faceN[cornersOf[q] / 3] * weight[cornersOf[q]];
                o->normals[v] = normalize(sum);
            }
            for (size_t k = 0; k < nCorners; ++k) o->tri_nidx[k] = o->tri_idx[k];
        } else {
            const float thresh = fCos((float)(angle * 0.01745329251994329576922));
            std::vector<int> groupIdx;                               // normals created for the current vertex
            for (size_t v = 0; v < nPoints; ++v) {
                groupIdx.clear();
                for (uint32_t q = first[v]; q < first[v + 1]; ++q) {
                    const uint32_t k = cornersOf[q];
                    // a triangle that names the vertex twice: the reference resolves both visits to the first slot holding it
                    uint32_t slot = 3 * (k / 3);
                    while (o->tri_idx[slot] != (int32_t)v) ++slot;
                    const v3 own = faceN[k / 3];
                    v3 sum = own * weight[k];
                    bool grouped = false;
                    for (uint32_t r = first[v]; r < first[v + 1]; ++r) {
                        const uint32_t k2 = cornersOf[r];
                        if (k2 / 3 == k / 3) continue;               // triangle_t::operator== is identity of the face (triangle.h:68-71)
                        if (dot(own, faceN[k2 / 3]) > thresh) { grouped = true; sum += faceN[k2 / 3] * weight[k2]; }
                    }
                    int n = -1;
                    if (grouped) {
                        sum = normalize(sum);
                        for (size_t g = 0; g < groupIdx.size() && n < 0; ++g)
                            if (dot(sum, o->normals[groupIdx[g]]) > 0.999) n = groupIdx[g];
                        if (n < 0) { n = (int)o->normals.size(); o->normals.push_back(sum); groupIdx.push_back(n); }
                    }
                    o->tri_nidx[slot] = n;
                }
            }
        }
    }
    o->is_smooth = true;
    geomChanged = true;
    return true;
}

// ------------------------------------------------------------------------------------------------ factories
static bool has_shader_params(const ParaMap &p, const std::list<ParaMap> &ep, std::string &which)
{
    static const char *keys[] = {"diffuse_shader", "mirror_color_shader", "bump_shader", "mirror_shader", "transparency_shader",
                                 "translucency_shader", "glossy_shader", "glossy_reflect_shader", 0};
    for (int i = 0; keys[i]; ++i) if (p.has(keys[i])) { which = keys[i]; return true; }
    if (!ep.empty()) { which = "shader node list"; return true; }
    return false;
}

// ------------------------------------------------------------------------------------------------ image textures
// The reference decodes image files in its image-handler plugins (src/image_handlers/*, a loader subsystem outside this path) and
// textureImage_t only calls imageHandler_t::getPixel.  Here the decoded pixels come in through yb200_register_image (what a plugin
// living inside the reference passes on from its handler); files are only read for the one self-contained format, TGA.
bool Context::registerImage(const char *fileName, const float *rgba, int w, int h, bool hdr)
{
    if (!fileName || !rgba || w <= 0 || h <= 0) return fail("register_image: bad arguments");
    HostImage &im = images[fileName];
    im.w = w; im.h = h; im.hdr = hdr;
    im.px.resize((size_t)w * h);
    memcpy((void *)im.px.data(), rgba, (size_t)w * h * sizeof(f4));
    return true;
}
// tgaHandler_t::loadFromFile (src/image_handlers/tgaHandler.cc:433-581) for true-colour 24 / 32 bit and 8-bit grey images,
// uncompressed or RLE, any origin; colour-mapped and 15 / 16 bit files are not read here.
bool Context::loadTga(const std::string &fileName, HostImage &img)
{
    FILE *fp = fopen(fileName.c_str(), "rb");
    if (!fp) return fail("TGAHandler: Cannot open file " + fileName);
    unsigned char hd[18];
    if (fread(hd, 1, 18, fp) != 18) { fclose(fp); return fail("TGAHandler: truncated file " + fileName); }
    const int idLength = hd[0], cmType = hd[1], imageType = hd[2], width = hd[12] | (hd[13] << 8), height = hd[14] | (hd[15] << 8), bitDepth = hd[16], desc = hd[17];
    const bool isRLE = imageType == 10 || imageType == 11, isGray = imageType == 3 || imageType == 11;
    const bool trueCol = imageType == 2 || imageType == 10;
    if (cmType != 0 || !(trueCol || isGray) || !((trueCol && (bitDepth == 24 || bitDepth == 32)) || (isGray && bitDepth == 8)) || width <= 0 || height <= 0) {
        fclose(fp);
        return fail("TGAHandler: " + fileName + ": only 24/32-bit true-colour and 8-bit grey TGA files are read by the B200 path (pass other images through yb200_register_image)");
    }
    fseek(fp, idLength, SEEK_CUR);
    const bool fromTop = ((desc & 0x20) >> 5) != 0, fromLeft = ((desc & 0x10) >> 4) != 0;   // sic: the reference mirrors in x when bit 4 is set
    img.w = width; img.h = height; img.hdr = false;
    img.px.assign((size_t)width * height, f4{0.f, 0.f, 0.f, 0.f});
    const int bpp = bitDepth / 8;
    long x = fromLeft ? width - 1 : 0, y = fromTop ? 0 : height - 1;
    const long minX = x, maxX = fromLeft ? -1 : width, stepX = fromLeft ? -1 : 1, maxY = fromTop ? height : -1, stepY = fromTop ? 1 : -1;
    const double inv255 = 0.00392156862745098039;
    auto put = [&](const unsigned char *c) {
        f4 p;
        if (bpp == 1) { p.x = p.y = p.z = p.w = (float)(c[0] * inv255); }                   // colorA_t(float): all four channels
        else { p.z = (float)(c[0] * inv255); p.y = (float)(c[1] * inv255); p.x = (float)(c[2] * inv255); p.w = bpp == 4 ? (float)(c[3] * inv255) : 1.f; }
        img.px[(size_t)y * width + x] = p;
        x += stepX;
        if (x == maxX) { x = minX; y += stepY; }
    };
    unsigned char c[4] = {0, 0, 0, 0};
    if (!isRLE) {
        for (size_t i = 0; i < (size_t)width * height && y != maxY; ++i) { if (fread(c, 1, bpp, fp) != (size_t)bpp) break; put(c); }
    } else {
        while (!feof(fp) && y != maxY) {
            unsigned char pack = 0;
            if (fread(&pack, 1, 1, fp) != 1) break;
            const bool rlePack = (pack & 0x80) != 0;
            const int rep = (int)(pack & 0x7f) + 1;
            if (rlePack && fread(c, 1, bpp, fp) != (size_t)bpp) break;
            for (int i = 0; i < rep && y != maxY; ++i) { if (!rlePack && fread(c, 1, bpp, fp) != (size_t)bpp) break; put(c); }
        }
    }
    fclose(fp);
    return true;
}

// ------------------------------------------------------------------------------------------------ textures + shader nodes
// newNoise (basictex.cc:26-50): generator id of yb_types.h for a noise_type string; anything unknown means newPerlin_t
static bool noise_kind(const std::string &ntype, int &kind)
{
    kind = 0;
    if (ntype == "blender") kind = 1;
    else if (ntype == "stdperlin") kind = 9;
    else if (ntype.find("voronoi") != std::string::npos)
        kind = ntype == "voronoi_f2" ? 4 : (ntype == "voronoi_f3" ? 5 : (ntype == "voronoi_f4" ? 6 : (ntype == "voronoi_f2f1" ? 7 : (ntype == "voronoi_crackle" ? 8 : 3))));
    else if (ntype == "cellnoise") kind = 2;
    return true;
}
uintptr_t Context::createTexture(const char *name)   // environment.cc:252-290 + src/textures/basictex.cc factories
{
    std::string type;
    if (!params.get("type", type)) { fail(std::string("Environment: Type of texture not specified for ") + name); return 0; }
    if (textureNames.count(name)) { warn(std::string("Environment: Sorry, texture already in table: ") + name); return 0; }
    const ParaMap &p = params;
    Texture t;
    memset((void *)&t, 0, sizeof(t));
    t.color1 = col3(0.f); t.color2 = col3(1.f); t.depth = 2; t.size = 1.f; t.turb = 1.f; t.sharpness = 1.f;
    std::string ntype, s;
    if (p.get("noise_type", ntype) && !noise_kind(ntype, t.noiseType)) { fail("texture '" + std::string(name) + "': noise_type 'stdperlin' is not supported"); return 0; }
    if (type == "rgb_cube") t.type = TEX_RGB_CUBE;
    else if (type == "blend") t.type = TEX_BLEND;
    else if (type == "clouds") {                     // basictex.cc:82-98
        bool hard = false;
        t.type = TEX_CLOUDS;
        p.getColor("color1", t.color1); p.getColor("color2", t.color2); p.get("depth", t.depth); p.get("size", t.size); p.get("hard", hard);
        t.hard = hard;
        if (p.get("bias", s)) t.bias = s == "positive" ? 1 : (s == "negative" ? 2 : 0);
    } else if (type == "marble" || type == "wood") { // basictex.cc:132-152,200-224
        bool hard = false; float shp = 1.f;
        t.type = type == "marble" ? TEX_MARBLE : TEX_WOOD;
        p.getColor("color1", t.color1); p.getColor("color2", t.color2); p.get("depth", t.depth); p.get("turbulence", t.turb); p.get("size", t.size);
        p.get("hard", hard); t.hard = hard;
        if (t.type == TEX_MARBLE) { p.get("sharpness", shp); if (shp > 1) t.sharpness = (float)(1.0 / (double)shp); }
        else if (p.get("wood_type", s)) t.rings = s == "rings";
        if (p.get("shape", s)) t.wshape = s == "saw" ? 1 : (s == "tri" ? 2 : 0);
    } else if (type == "voronoi") {                  // basictex.cc:320-344
        std::string cltype, dname; float fw1 = 1, fw2 = 0, fw3 = 0, fw4 = 0, mex = 2.5f, isc = 1, sz = 1;
        p.getColor("color1", t.color1); p.getColor("color2", t.color2);
        if (p.get("color_type", cltype)) t.coltype = cltype == "col1" ? 1 : (cltype == "col2" ? 2 : (cltype == "col3" ? 3 : 0));
        p.get("weight1", fw1); p.get("weight2", fw2); p.get("weight3", fw3); p.get("weight4", fw4);
        p.get("mk_exponent", mex); p.get("intensity", isc); p.get("size", sz); p.get("distance_metric", dname);
        t.type = TEX_VORONOI; t.size = sz; t.mkExp = mex;
        t.distMetric = (dname == "squared" || dname == "manhattan") ? 1 : (dname == "chebychev" ? 2 : (dname == "minkovsky_half" ? 3 : (dname == "minkovsky_four" ? 4 : (dname == "minkovsky" ? 5 : 0))));
        t.vw[0] = fw1; t.vw[1] = fw2; t.vw[2] = fw3; t.vw[3] = fw4;
        for (int k = 0; k < 4; ++k) t.vaw[k] = fabsf(t.vw[k]);
        t.iscale = t.vaw[0] + t.vaw[1] + t.vaw[2] + t.vaw[3];
        if (t.iscale != 0) t.iscale = isc / t.iscale;
    } else if (type == "musgrave") {                 // basictex.cc:385-407
        std::string mtype; float H = 1, lacu = 2, octs = 2, offs = 1, gain = 1, size = 1, iscale = 1;
        p.getColor("color1", t.color1); p.getColor("color2", t.color2); p.get("musgrave_type", mtype);
        p.get("H", H); p.get("lacunarity", lacu); p.get("octaves", octs); p.get("offset", offs); p.get("gain", gain); p.get("size", size); p.get("intensity", iscale);
        t.type = TEX_MUSGRAVE; t.size = size; t.iscale = iscale;
        t.musType = mtype == "multifractal" ? 1 : (mtype == "heteroterrain" ? 2 : (mtype == "hybridmf" ? 3 : (mtype == "ridgedmf" ? 4 : 0)));
        t.musH = H; t.musLacu = lacu; t.musOcts = octs; t.musOffs = offs; t.musGain = gain;
    } else if (type == "distorted_noise") {          // basictex.cc:443-458
        std::string n1, n2; float dist = 1, size = 1; int k1 = 0, k2 = 0;
        p.getColor("color1", t.color1); p.getColor("color2", t.color2); p.get("distort", dist); p.get("size", size);
        if (p.get("noise_type1", n1) && !noise_kind(n1, k1)) { fail("texture '" + std::string(name) + "': noise_type1 'stdperlin' is not supported"); return 0; }
        if (p.get("noise_type2", n2) && !noise_kind(n2, k2)) { fail("texture '" + std::string(name) + "': noise_type2 'stdperlin' is not supported"); return 0; }
        t.type = TEX_DISTNOISE; t.size = size; t.distort = dist; t.noiseType = k1; t.noiseType2 = k2;
    } else if (type == "image") {                    // imagetex.cc:241-340
        std::string fname, intpstr, clip;
        double gamma = 1.0, minx = 0.0, miny = 0.0, maxx = 1.0, maxy = 1.0, cdist = 0.0; bool normalmap = false, rot90 = false, even = false, odd = true;
        int xrep = 1, yrep = 1;
        if (!p.get("filename", fname)) { fail("ImageTexture: Required argument filename not found for image texture"); return 0; }
        std::map<std::string, HostImage>::iterator ii = images.find(fname);
        if (ii == images.end()) {
            std::string ext = fname.substr(fname.rfind('.') + 1);
            for (size_t i = 0; i < ext.size(); ++i) ext[i] = (char)tolower(ext[i]);
            if (ext != "tga" && ext != "tpic") {
                fail("ImageTexture: '" + fname + "': only TGA files are decoded by the B200 path; register the decoded pixels with yb200_register_image");
                return 0;
            }
            HostImage im;
            if (!loadTga(fname, im)) return 0;
            ii = images.insert(std::make_pair(fname, im)).first;
        }
        const HostImage &im = ii->second;
        p.get("interpolate", intpstr); p.get("gamma", gamma); p.get("normalmap", normalmap);
        p.get("xrepeat", xrep); p.get("yrepeat", yrep); p.get("cropmin_x", minx); p.get("cropmin_y", miny); p.get("cropmax_x", maxx); p.get("cropmax_y", maxy);
        p.get("rot90", rot90); p.get("clipping", clip); p.get("even_tiles", even); p.get("odd_tiles", odd); p.get("checker_dist", cdist);
        t.type = TEX_IMAGE; t.resx = im.w; t.resy = im.h; t.isHDR = im.hdr; t.normalmap = normalmap;
        t.intp = intpstr == "none" ? 0 : (intpstr == "bicubic" ? 2 : 1);
        t.gamma = (float)gamma; t.xrepeat = xrep; t.yrepeat = yrep; t.rot90 = rot90;
        t.cropminx = (float)minx; t.cropmaxx = (float)maxx; t.cropminy = (float)miny; t.cropmaxy = (float)maxy;
        t.cropx = (t.cropminx != 0.0f) || (t.cropmaxx != 1.0f); t.cropy = (t.cropminy != 0.0f) || (t.cropmaxy != 1.0f);
        t.clipmode = clip == "extend" ? TCL_EXTEND : (clip == "clip" ? TCL_CLIP : (clip == "clipcube" ? TCL_CLIPCUBE : (clip == "checker" ? TCL_CHECKER : TCL_REPEAT)));
        t.checkerEven = even; t.checkerOdd = odd; t.checkerDist = (float)cdist;
        texPixOffset.resize(textures.size() + 1, (size_t)-1);
        texPixOffset[textures.size()] = texPixels.size();
        texPixels.insert(texPixels.end(), im.px.begin(), im.px.end());
    } else {
        fail("Environment: texture type '" + type + "' is not supported by the B200 path (supported: rgb_cube, blend, clouds, marble, wood, voronoi, musgrave, distorted_noise, image)");
        return 0;
    }
    texPixOffset.resize(textures.size() + 1, (size_t)-1);
    textures.push_back(t);
    textureNames[name] = (int)textures.size() - 1;
    return 0x54000000u | (uintptr_t)textures.size();
}

// nodeMaterial_t::loadNodes + parseNodes + solveNodesOrder (nodematerial.cc:77-243) for the node types of yb_types.h.  The
// material's nodes are appended to the scene's node table in evaluation order; m.slot[] gets the index of each bound root.
bool Context::loadMaterialNodes(const char *name, Material &m, const char *const *slotKeys, const int *slotIds, int nSlots)
{
    m.nodeBase = (int32_t)shaderNodes.size(); m.nNodes = 0;
    for (int i = 0; i < 8; ++i) m.slot[i] = -1;
    m.bumpMask = 0;
    if (eparams.empty()) {
        for (int k = 0; k < nSlots; ++k) { std::string nm; if (params.get(slotKeys[k], nm)) warn("Shader node " + std::string(slotKeys[k]) + " '" + nm + "' does not exist!"); }
        return true;
    }
    struct Raw { ShaderNode n; std::string name, in1, in2, factor, input, upper; };
    std::vector<Raw> raw;
    std::map<std::string, int> table;
    for (std::list<ParaMap>::const_iterator it = eparams.begin(); it != eparams.end(); ++it) {
        const ParaMap &p = *it;
        std::string element, nname, type;
        if (p.get("element", element) && element != "shader_node") continue;
        if (!p.get("name", nname)) return fail("NodeMaterial: Name of shader node not specified!");
        if (table.count(nname)) return fail("NodeMaterial: Multiple nodes with identically names!");
        if (!p.get("type", type)) return fail("NodeMaterial: Type of shader node not specified!");
        Raw r;
        memset((void *)&r.n, 0, sizeof(r.n));
        r.name = nname;
        ShaderNode &n = r.n;
        n.in1 = n.in2 = n.inFactor = n.input = n.upper = -1;
        if (type == "texture_mapper") {                 // basicnodes.cc:236-304
            std::string texname, opt;
            if (!p.get("texture", texname)) return fail("TextureMapper: No texture given for texture mapper!");
            std::map<std::string, int>::const_iterator ti = textureNames.find(texname);
            if (ti == textureNames.end()) return fail("TextureMapper: texture '" + texname + "' does not exist!");
            n.type = NODE_TEXMAP; n.tex = ti->second; n.texCoords = TXC_GLOB;
            if (p.get("texco", opt)) {
                if (opt == "uv") n.texCoords = TXC_UV;
                else if (opt == "orco") n.texCoords = TXC_ORCO;
                else if (opt == "transformed") n.texCoords = TXC_TRAN;
                else if (opt == "window") return fail("TextureMapper: texco 'window' is not supported");
                else n.texCoords = TXC_GLOB;             // normal / reflect / stick / stress / tangent fall back to global (basicnodes.cc:139-146)
            }
            float bump = 1.f; bool scalar = true; int map[3] = {1, 2, 3}; v3 offset(0.f), scale(1.f);
            p.getPoint("scale", scale); p.getPoint("offset", offset); p.get("do_scalar", scalar); p.get("bump_strength", bump);
            p.get("proj_x", map[0]); p.get("proj_y", map[1]); p.get("proj_z", map[2]);
            for (int i = 0; i < 3; ++i) map[i] = std::min(3, std::max(0, map[i]));
            n.mapX = map[0]; n.mapY = map[1]; n.mapZ = map[2]; n.scale = scale; n.offset = 2.f * offset; n.doScalar = scalar;
            const Texture &tx = textures[n.tex];
            n.mapType = TXP_PLAIN;
            if (p.get("mapping", opt) && tx.type == TEX_IMAGE) n.mapType = opt == "cube" ? TXP_CUBE : (opt == "tube" ? TXP_TUBE : (opt == "sphere" ? TXP_SPHERE : TXP_PLAIN));
            if (tx.type == TEX_IMAGE) { n.dU = 1.f / (float)tx.resx; n.dV = 1.f / (float)tx.resy; }   // setup() (basicnodes.cc:12-38)
            else n.dU = n.dV = 0.0002f;
            n.bumpStr = bump / length(scale);
            if (!(tx.type == TEX_IMAGE && tx.normalmap)) n.bumpStr /= 100.0f;
            const float ident[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
            memcpy(n.mtx, ident, sizeof(ident));
            std::map<std::string, std::vector<float> >::const_iterator mi = p.mdicc.find("transform");
            if (mi != p.mdicc.end() && mi->second.size() >= 12) memcpy(n.mtx, mi->second.data(), 12 * sizeof(float));
        } else if (type == "value") {                   // basicnodes.cc:322-333
            col3 col(1.f); float alpha = 1.f, val = 1.f;
            p.getColor("color", col); p.get("alpha", alpha); p.get("scalar", val);
            n.type = NODE_VALUE; n.col1[0] = col.x; n.col1[1] = col.y; n.col1[2] = col.z; n.col1[3] = alpha; n.val1 = val;
        } else if (type == "mix") {                     // basicnodes.cc:383-436,580-600
            float val = 0.5f; int mode = 0;
            p.get("cfactor", val); p.get("mode", mode);
            n.type = NODE_MIX; n.mode = mode; n.cfactor = (mode == 0 || mode == 5 || mode < 0 || mode > 9) ? val : 0.f;   // subclasses start from cfactor(0)
            std::map<std::string, Param>::const_iterator ci;
            if (!p.get("input1", r.in1)) {
                ci = p.dicc.find("color1");
                if (ci == p.dicc.end() || ci->second.type != Param::T_COLOR) return fail("MixNode: Color1 not set");
                for (int k = 0; k < 4; ++k) n.col1[k] = ci->second.v[k];
            }
            if (!p.get("input2", r.in2)) {
                ci = p.dicc.find("color2");
                if (ci == p.dicc.end() || ci->second.type != Param::T_COLOR) return fail("MixNode: Color2 not set");
                for (int k = 0; k < 4; ++k) n.col2[k] = ci->second.v[k];
            }
            if (!p.get("factor", r.factor)) { if (!p.get("value", n.cfactor)) return fail("MixNode: Value not set"); }
            // val1 / val2 (the scalars of constant inputs) are never initialised by the reference: 0 here
        } else if (type == "layer") {                   // layernode.cc:190-225
            col3 def_col(1.f); bool do_color = true, do_scalar = false, color_input = true, use_alpha = false, stencil = false, noRGB = false, negative = false;
            double def_val = 1.0, colfac = 1.0, valfac = 1.0; int mode = 0;
            p.get("mode", mode); p.getColor("def_col", def_col); p.get("colfac", colfac); p.get("def_val", def_val); p.get("valfac", valfac);
            p.get("do_color", do_color); p.get("do_scalar", do_scalar); p.get("color_input", color_input); p.get("use_alpha", use_alpha);
            p.get("noRGB", noRGB); p.get("stencil", stencil); p.get("negative", negative);
            n.type = NODE_LAYER; n.mode = mode;
            n.texflag = (noRGB ? TXF_RGBTOINT : 0u) | (stencil ? TXF_STENCIL : 0u) | (negative ? TXF_NEGATIVE : 0u) | (use_alpha ? TXF_ALPHAMIX : 0u);
            n.colfac = (float)colfac; n.valfac = (float)valfac; n.defVal = (float)def_val; n.defCol = def_col;
            n.doColor = do_color; n.doScalarL = do_scalar; n.colorInput = color_input; n.useAlpha = use_alpha;
            if (!p.get("input", r.input)) return fail("LayerNode: input not set");
            if (!p.get("upper_layer", r.upper)) {
                std::map<std::string, Param>::const_iterator ci = p.dicc.find("upper_color");
                if (ci != p.dicc.end() && ci->second.type == Param::T_COLOR) for (int k = 0; k < 4; ++k) n.upperCol[k] = ci->second.v[k];
                else { n.upperCol[0] = n.upperCol[1] = n.upperCol[2] = 0.f; n.upperCol[3] = 1.f; }
                float uv = 0.f; p.get("upper_value", uv); n.upperVal = uv;
            }
        } else return fail("NodeMaterial: Don't know how to create shader node of type '" + type + "'!");
        table[nname] = (int)raw.size();
        raw.push_back(r);
    }
    // configInputs: names -> raw indices
    std::vector<std::vector<int> > deps(raw.size());
    auto link = [&](const std::string &nm, int &dst, size_t self, const char *what) -> bool {
        if (nm.empty()) return true;
        std::map<std::string, int>::const_iterator i = table.find(nm);
        if (i == table.end()) { fail(std::string(what) + ": Couldn't get input " + nm); return false; }
        dst = i->second; deps[self].push_back(i->second);
        return true;
    };
    for (size_t i = 0; i < raw.size(); ++i) {
        Raw &r = raw[i];
        // dependency order of getDependencies: mix = input1, input2, factor; layer = input, upper layer
        if (!link(r.in1, r.n.in1, i, "MixNode") || !link(r.in2, r.n.in2, i, "MixNode") || !link(r.factor, r.n.inFactor, i, "MixNode")) return false;
        if (!link(r.input, r.n.input, i, "LayerNode") || !link(r.upper, r.n.upper, i, "LayerNode")) return false;
    }
    // roots in the order of the material's nodeList map (std::map<std::string, ...>: alphabetical by slot key), then
    // recursiveSolver: depth-first, dependencies before the node
    std::vector<std::pair<std::string, int> > roots;   // (slot key, raw index)
    std::vector<int> rootSlot;
    {
        std::vector<std::pair<std::string, int> > keys;
        for (int k = 0; k < nSlots; ++k) keys.push_back(std::make_pair(std::string(slotKeys[k]), slotIds[k]));
        std::sort(keys.begin(), keys.end());
        for (size_t k = 0; k < keys.size(); ++k) {
            std::string nm;
            if (!params.get(keys[k].first, nm)) continue;
            std::map<std::string, int>::const_iterator i = table.find(nm);
            if (i == table.end()) { warn("Shader node " + keys[k].first + " '" + nm + "' does not exist!"); continue; }
            roots.push_back(std::make_pair(keys[k].first, i->second)); rootSlot.push_back(keys[k].second);
        }
    }
    std::vector<int> order, newIndex(raw.size(), -1);
    std::vector<char> seen(raw.size(), 0);
    struct Rec { static void go(int n, const std::vector<std::vector<int> > &deps, std::vector<char> &seen, std::vector<int> &order) {
        if (seen[n]) return; seen[n] = 1;
        for (size_t k = 0; k < deps[n].size(); ++k) if (!seen[deps[n][k]]) go(deps[n][k], deps, seen, order);
        order.push_back(n); } };
    for (size_t k = 0; k < roots.size(); ++k) Rec::go(roots[k].second, deps, seen, order);
    if (order.size() != raw.size()) warn("NodeMaterial: Unreachable nodes!");
    if (order.size() > YB_MAX_NODES) return fail("material '" + std::string(name) + "': more than " + std::to_string(YB_MAX_NODES) + " shader nodes are not supported");
    for (size_t k = 0; k < order.size(); ++k) newIndex[order[k]] = (int)k;
    for (size_t k = 0; k < order.size(); ++k) {
        ShaderNode n = raw[order[k]].n;
        if (n.in1 >= 0) n.in1 = newIndex[n.in1];
        if (n.in2 >= 0) n.in2 = newIndex[n.in2];
        if (n.inFactor >= 0) n.inFactor = newIndex[n.inFactor];
        if (n.input >= 0) n.input = newIndex[n.input];
        if (n.upper >= 0) n.upper = newIndex[n.upper];
        shaderNodes.push_back(n);
    }
    m.nNodes = (int32_t)order.size();
    for (size_t k = 0; k < roots.size(); ++k) m.slot[rootSlot[k]] = newIndex[roots[k].second];
    if (m.slot[SLOT_BUMP] >= 0) {           // getNodeList(bumpS, bumpNodes): the bump root and everything it depends on
        std::vector<int> todo(1, order[m.slot[SLOT_BUMP]]);
        while (!todo.empty()) {
            const int n = todo.back(); todo.pop_back();
            if (m.bumpMask & (1u << newIndex[n])) continue;
            m.bumpMask |= 1u << newIndex[n];
            for (size_t k = 0; k < deps[n].size(); ++k) todo.push_back(deps[n][k]);
        }
    }
    return true;
}

// glass.cc:275-304 / roughglass.cc:288-316: an `absorption` colour below white gives the material a "beer" volume handler for its inside
// (beer.cc:24-33: sigma_a = -log(colour) / absorption_dist, distance 1 unless given) and BSDF_VOLUMETRIC
static void setBeerAbsorption(const ParaMap &p, Material &m)
{
    col3 absorp(1.f);
    if (!p.getColor("absorption", absorp) || !(absorp.x < 1.f || absorp.y < 1.f || absorp.z < 1.f)) return;
    double dist = 1.f;
    p.get("absorption_dist", dist);
    const float maxlog = (float)log(1e38);
    col3 sigma;
    sigma.x = (absorp.x > 1e-38f) ? (float)-log((double)absorp.x) : maxlog;
    sigma.y = (absorp.y > 1e-38f) ? (float)-log((double)absorp.y) : maxlog;
    sigma.z = (absorp.z > 1e-38f) ? (float)-log((double)absorp.z) : maxlog;
    if (dist != 0.0) sigma = sigma * (float)(1.f / dist);
    m.hasBeer = 1; m.beerSigma = sigma;
    m.bsdfFlags |= BSDF_VOLUMETRIC;
}

uintptr_t Context::createMaterial(const char *name)   // environment.cc:218-250 + the plugins' factory()
{
    std::string type;
    if (!params.get("type", type)) { fail(std::string("Environment: Type of material not specified for ") + name); return 0; }
    if (materialNames.count(name)) { warn(std::string("Environment: Sorry, material already in table: ") + name); return 0; }
    Material m;
    memset((void *)&m, 0, sizeof(m));
    for (int i = 0; i < 8; ++i) m.slot[i] = -1;
    const size_t nodesBefore = shaderNodes.size();
    const ParaMap &p = params;
    if (type == "shinydiffusemat") {              // shinydiffuse.cc:470-515 + ctor :5-14 + config :32-80
        col3 diffuseColor(1.f), mirrorColor(1.f);
        float diffuseStrength = 1.f, transparency = 0.f, translucency = 0.f, mirrorStrength = 0.f, emit = 0.f;
        bool fresnelEffect = false; double IOR = 1.33, filt = 1.0;
        p.getColor("color", diffuseColor); p.getColor("mirror_color", mirrorColor);
        p.get("transparency", transparency); p.get("translucency", translucency); p.get("diffuse_reflect", diffuseStrength);
        p.get("specular_reflect", mirrorStrength); p.get("emit", emit); p.get("IOR", IOR); p.get("fresnel_effect", fresnelEffect);
        p.get("transmit_filter", filt);
        m.type = MAT_SHINYDIFFUSE;
        m.diffuseColor = diffuseColor; m.mirrorColor = mirrorColor;
        m.mirrorStrength = mirrorStrength; m.transparencyStrength = transparency; m.translucencyStrength = translucency;
        m.diffuseStrength = diffuseStrength; m.transmitFilterStrength = (float)filt;
        m.emitColor = emit * diffuseColor;
        m.bsdfFlags = BSDF_NONE;
        if (emit > 0.f) m.bsdfFlags |= BSDF_EMIT;
        if (fresnelEffect) { m.iorSquared = (float)(IOR * IOR); m.hasFresnel = 1; }
        std::string brdf;
        if (p.get("diffuse_brdf", brdf) && brdf == "oren_nayar") {
            double sigma = 0.1; p.get("sigma", sigma);
            double s2 = sigma * sigma;
            m.orenA = (float)(1.0 - 0.5 * (s2 / (s2 + 0.33)));
            m.orenB = (float)(0.45 * s2 / (s2 + 0.09));
            m.useOrenNayar = 1;
        }
        m.emitStrength = emit;
        {
            static const char *const keys[] = {"diffuse_shader", "mirror_color_shader", "mirror_shader", "transparency_shader", "translucency_shader", "bump_shader"};
            static const int ids[] = {SLOT_DIFFUSE, SLOT_MIRROR_COLOR, SLOT_MIRROR, SLOT_TRANSPARENCY, SLOT_TRANSLUCENCY, SLOT_BUMP};
            if (!loadMaterialNodes(name, m, keys, ids, 6)) { shaderNodes.resize(nodesBefore); return 0; }
        }
        const bool shMirror = m.slot[SLOT_MIRROR] >= 0, shTransp = m.slot[SLOT_TRANSPARENCY] >= 0, shTransl = m.slot[SLOT_TRANSLUCENCY] >= 0;
        // config() (shinydiffuse.cc:27-73): a bound shader switches the component on and leaves `acc` alone
        float acc = 1.f; int n = 0;
        if (m.mirrorStrength > 0.00001f || shMirror) {
            m.isMirror = 1;
            if (shMirror) {} else if (!m.hasFresnel) acc = 1.f - m.mirrorStrength;
            m.bsdfFlags |= BSDF_SPECULAR | BSDF_REFLECT; m.cFlags[n] = BSDF_SPECULAR | BSDF_REFLECT; m.cIndex[n] = 0; ++n;
        }
        if (m.transparencyStrength * acc > 0.00001f || shTransp) {
            m.isTransparent = 1; if (!shTransp) acc *= 1.f - m.transparencyStrength;
            m.bsdfFlags |= BSDF_TRANSMIT | BSDF_FILTER; m.cFlags[n] = BSDF_TRANSMIT | BSDF_FILTER; m.cIndex[n] = 1; ++n;
        }
        if (m.translucencyStrength * acc > 0.00001f || shTransl) {
            m.isTranslucent = 1; if (!shTransl) acc *= 1.f - m.transparencyStrength;   // sic (shinydiffuse.cc:57)
            m.bsdfFlags |= BSDF_DIFFUSE | BSDF_TRANSMIT; m.cFlags[n] = BSDF_DIFFUSE | BSDF_TRANSMIT; m.cIndex[n] = 2; ++n;
        }
        if (m.diffuseStrength * acc > 0.00001f) {
            m.isDiffuse = 1;
            m.bsdfFlags |= BSDF_DIFFUSE | BSDF_REFLECT; m.cFlags[n] = BSDF_DIFFUSE | BSDF_REFLECT; m.cIndex[n] = 3; ++n;
        }
        m.nBSDF = n;
    } else if (type == "glossy" || type == "coated_glossy") {   // glossy.cc:353-395, coatedglossy.cc:420-470
        const bool coated = type == "coated_glossy";
        col3 col(1.f), dcol(1.f), mirCol(1.f);
        float refl = 1.f, diff = 0.f, exponent = 50.f; double ior = 1.4; bool as_diff = true, aniso = false;
        p.getColor("color", col); p.getColor("diffuse_color", dcol); p.get("diffuse_reflect", diff); p.get("glossy_reflect", refl);
        p.get("as_diffuse", as_diff); p.get("exponent", exponent); p.get("anisotropic", aniso);
        if (coated) { p.get("IOR", ior); p.getColor("mirror_color", mirCol); if (ior == 1.f) ior = 1.0000001f; }
        m.type = coated ? MAT_COATED_GLOSSY : MAT_GLOSSY;
        m.glossColor = col; m.diffColor = dcol; m.mirrorColor = mirCol; m.reflectivity = refl; m.mDiffuse = diff;
        m.exponent = exponent; m.asDiffuse = as_diff; m.ior = (float)ior;
        if (aniso) {
            if (coated) { double eu = 50.0, ev = 50.0; p.get("exp_u", eu); p.get("exp_v", ev); m.expU = (float)eu; m.expV = (float)ev; }
            else { float eu = 50.f, ev = 50.f; p.get("exp_u", eu); p.get("exp_v", ev); m.expU = eu; m.expV = ev; }
            m.anisotropic = 1;
        }
        std::string brdf;
        if (p.get("diffuse_brdf", brdf) && brdf == "Oren-Nayar") {
            double sigma = 0.1; p.get("sigma", sigma);
            double s2 = sigma * sigma;
            m.orenA = (float)(1.0 - 0.5 * (s2 / (s2 + 0.33)));
            m.orenB = (float)(0.45 * s2 / (s2 + 0.09));
            m.orenNayar = 1;
        }
        {
            static const char *const keys[] = {"diffuse_shader", "glossy_shader", "glossy_reflect_shader", "bump_shader"};
            static const int ids[] = {SLOT_DIFFUSE, SLOT_GLOSSY, SLOT_GLOSSY_REFLECT, SLOT_BUMP};
            if (!loadMaterialNodes(name, m, keys, ids, 4)) { shaderNodes.resize(nodesBefore); return 0; }
        }
        const uint32_t glossyFlags = as_diff ? (BSDF_DIFFUSE | BSDF_REFLECT) : (BSDF_GLOSSY | BSDF_REFLECT);
        if (coated) {
            m.cFlags[0] = BSDF_SPECULAR | BSDF_REFLECT; m.cFlags[1] = glossyFlags;
            if (diff > 0) { m.cFlags[2] = BSDF_DIFFUSE | BSDF_REFLECT; m.withDiffuse = 1; m.nBSDF = 3; }
            else { m.cFlags[2] = BSDF_NONE; m.nBSDF = 2; }
            m.bsdfFlags = m.cFlags[0] | m.cFlags[1] | m.cFlags[2];
        } else {
            m.bsdfFlags = BSDF_NONE;
            if (diff > 0) { m.bsdfFlags = BSDF_DIFFUSE | BSDF_REFLECT; m.withDiffuse = 1; }
            m.bsdfFlags |= glossyFlags;
        }
    } else if (type == "glass") {                 // glass.cc:255-300
        double IOR = 1.4, filt = 0.f, disp_power = 0.0; col3 filtCol(1.f), absorp(1.f), srCol(1.f); bool fake = false;
        p.get("IOR", IOR); p.getColor("filter_color", filtCol); p.get("transmit_filter", filt); p.getColor("mirror_color", srCol);
        p.get("dispersion_power", disp_power); p.get("fake_shadows", fake);
        if (disp_power > 0.0) { fail("glass '" + std::string(name) + "': dispersion is not supported"); return 0; }
        {
            static const char *const keys[] = {"mirror_color_shader", "bump_shader"};
            static const int ids[] = {SLOT_MIRROR_COLOR, SLOT_BUMP};
            if (!loadMaterialNodes(name, m, keys, ids, 2)) { shaderNodes.resize(nodesBefore); return 0; }
        }
        m.type = MAT_GLASS;
        m.ior = (float)IOR;
        m.filterCol = (float)filt * filtCol + col3((float)(1.f - filt));
        m.specRefCol = srCol; m.fakeShadow = fake;
        m.bsdfFlags = BSDF_ALL_SPECULAR;
        if (fake) m.bsdfFlags |= BSDF_FILTER;
        setBeerAbsorption(p, m);
    } else if (type == "rough_glass") {           // roughglass.cc:257-285
        float IOR = 1.4f, filt = 0.f, alpha = 0.5f, disp_power = 0.f; col3 filtCol(1.f), absorp(1.f), srCol(1.f); bool fake = false;
        p.get("IOR", IOR); p.getColor("filter_color", filtCol); p.get("transmit_filter", filt); p.getColor("mirror_color", srCol);
        p.get("alpha", alpha); p.get("dispersion_power", disp_power); p.get("fake_shadows", fake);
        if (disp_power > 0.0) { fail("rough_glass '" + std::string(name) + "': dispersion is not supported"); return 0; }
        alpha = std::max(1e-4f, std::min(alpha * 0.5f, 1.f));
        {
            static const char *const keys[] = {"mirror_color_shader", "bump_shader"};
            static const int ids[] = {SLOT_MIRROR_COLOR, SLOT_BUMP};
            if (!loadMaterialNodes(name, m, keys, ids, 2)) { shaderNodes.resize(nodesBefore); return 0; }
        }
        m.type = MAT_ROUGH_GLASS;
        m.ior = IOR; m.rgAlpha2 = alpha * alpha;
        m.filterCol = filt * filtCol + col3(1.f - filt);
        m.specRefCol = srCol; m.fakeShadow = fake;
        m.bsdfFlags = BSDF_ALL_GLOSSY;
        if (fake) m.bsdfFlags |= BSDF_FILTER;
        setBeerAbsorption(p, m);
    } else if (type == "blend_mat") {             // blend.cc:395-431 + ctor :32-41
        std::string n1, n2; double blend_val = 0.5;
        if (!p.get("material1", n1) || !p.get("material2", n2)) { fail("blend_mat '" + std::string(name) + "': material1 / material2 missing"); return 0; }
        std::map<std::string, int>::const_iterator i1 = materialNames.find(n1), i2 = materialNames.find(n2);
        if (i1 == materialNames.end() || i2 == materialNames.end()) { fail("blend_mat '" + std::string(name) + "': material '" + (i1 == materialNames.end() ? n1 : n2) + "' does not exist"); return 0; }
        if (materials[i1->second].nNodes > 0 || materials[i2->second].nNodes > 0) { fail("blend_mat '" + std::string(name) + "': blending materials with shader nodes is not supported"); return 0; }
        if (materials[i1->second].type == MAT_BLEND || materials[i2->second].type == MAT_BLEND) { fail("blend_mat '" + std::string(name) + "': blending a blend_mat is not supported (one level only)"); return 0; }
        p.get("blend_value", blend_val);
        const int self = (int)materials.size();
        m.type = MAT_BLEND; m.blendOff1 = i1->second - self; m.blendOff2 = i2->second - self; m.blendVal = (float)blend_val;
        m.bsdfFlags = materials[i1->second].bsdfFlags | materials[i2->second].bsdfFlags;
        if (p.has("mask")) {                        // blend.cc:406-427: the scalar of this node is the blend value at every point
            const size_t nodesBefore = shaderNodes.size();
            static const char *const keys[] = {"mask"};
            static const int ids[] = {SLOT_BLEND_MASK};
            if (!loadMaterialNodes(name, m, keys, ids, 1)) { shaderNodes.resize(nodesBefore); return 0; }
        }
    } else if (type == "mirror") {                // glass.cc:336-371
        col3 col(1.f); float refl = 1.f;
        p.getColor("color", col); p.get("reflect", refl);
        m.type = MAT_MIRROR; m.specRefCol = col * refl; m.mirrorRefl = refl > 1.f ? 1.f : refl;
        m.bsdfFlags = BSDF_SPECULAR;
    } else if (type == "null") {
        m.type = MAT_NULL; m.bsdfFlags = BSDF_NONE;
    } else if (type == "light_mat") {             // simple.cc:77-86
        col3 col(1.f); double power = 1.0; bool ds = false;
        p.getColor("color", col); p.get("power", power); p.get("double_sided", ds);
        m.type = MAT_LIGHT; m.lightCol = col * (float)power; m.doubleSided = ds; m.bsdfFlags = BSDF_EMIT;
    } else {
        fail("Environment: material type '" + type + "' is not supported by the B200 path (supported: shinydiffusemat, glossy, coated_glossy, glass, rough_glass, blend_mat, mirror, null, light_mat)");
        return 0;
    }
    if (materials.size() >= TRI_MAT_MASK) { fail("too many materials"); return 0; }
    materials.push_back(m);
    materialNames[name] = (int)materials.size() - 1;
    return HANDLE_MAT_TAG | (uintptr_t)materials.size();
}

uintptr_t Context::createLight(const char *name)   // environment.cc:189-216 + light factories
{
    std::string type;
    if (!params.get("type", type)) { fail(std::string("Environment: Type of light not specified for ") + name); return 0; }
    if (lightNames.count(name)) { warn(std::string("Environment: Sorry, light already in table: ") + name); return 0; }
    const ParaMap &p = params;
    Light l;
    memset(&l, 0, sizeof(l));
    unsigned obj = 0;
    if (type == "pointlight") {                    // pointlight.cc:43-52,117-129
        v3 from(0.f); col3 color(1.f); float power = 1.0f;
        p.getPoint("from", from); p.getColor("color", color); p.get("power", power);
        l.type = LIGHT_POINT; l.position = from; l.color = color * power; l.samples = 1;
    } else if (type == "sunlight") {               // sunlight.cc:52-63,121-136
        v3 dir(0.f, 0.f, 1.f); col3 color(1.f); float power = 1.0f, angle = 0.27f; int samples = 4;
        p.getPoint("direction", dir); p.getColor("color", color); p.get("power", power); p.get("angle", angle); p.get("samples", samples);
        l.type = LIGHT_SUN; l.samples = samples;
        l.color = color * power;
        l.direction = normalize(dir);
        createCS(dir, l.du, l.dv);                 // sic: from the un-normalised vector (sunlight.cc:57)
        if (angle > 80.f) angle = 80.f;
        l.cosAngle = fCos((float)(angle * 0.01745329251994329576922));
        l.invpdf = (float)(YB_M_2PI * (1.f - l.cosAngle));
        l.pdf = (float)(1.0 / l.invpdf);
        l.colPdf = l.color * l.pdf;
    } else if (type == "arealight") {              // arealight.cc:33-49,181-199
        v3 corner(0.f), p1(0.f), p2(0.f); col3 color(1.f); float power = 1.0f; int samples = 4, object = 0;
        p.getPoint("corner", corner); p.getPoint("point1", p1); p.getPoint("point2", p2); p.getColor("color", color);
        p.get("power", power); p.get("samples", samples); p.get("object", object);
        l.type = LIGHT_AREA; l.samples = samples;
        l.corner = corner; l.toX = p1 - corner; l.toY = p2 - corner;
        l.fnormal = cross(l.toY, l.toX);
        l.color = (color * power) * (float)YB_M_PI;
        l.area = normLen(l.fnormal);
        l.invArea = (float)(1.0 / l.area);
        l.normal = -l.fnormal;
        l.du = normalize(l.toX); l.dv = cross(l.normal, l.du);
        l.c2 = corner + l.toX; l.c3 = corner + (l.toX + l.toY); l.c4 = corner + l.toY;
        obj = (unsigned)object;
    } else if (type == "meshlight") {              // meshlight.cc:216-233; geometry is bound at scene update (meshLight_t::init)
        col3 color(1.f); double power = 1.0; int samples = 4, object = 0; bool ds = false;
        p.get("object", object); p.getColor("color", color); p.get("power", power); p.get("samples", samples); p.get("double_sided", ds);
        l.type = LIGHT_MESH; l.samples = samples; l.doubleSided = ds ? 1 : 0;
        l.color = (color * (float)power) * (float)YB_M_PI;
        obj = (unsigned)object;
    } else if (type == "directional") {            // directional.cc:52-61,142-160
        v3 from(0.f), dir(0.f, 0.f, 1.f); col3 color(1.f); float power = 1.0f, rad = 1.f; bool inf = true;
        p.getPoint("direction", dir); p.getColor("color", color); p.get("power", power); p.get("infinite", inf);
        if (!inf) {
            if (!p.getPoint("from", from)) { if (p.getPoint("position", from)) info("DirectionalLight: Deprecated parameter 'position', use 'from' instead"); }
            p.get("radius", rad);
        }
        l.type = LIGHT_DIRECTIONAL; l.samples = 1;
        l.color = color * power; l.position = from; l.radius = rad; l.infinite = inf ? 1 : 0;
        l.direction = normalize(dir);
        createCS(l.direction, l.du, l.dv);
    } else if (type == "spotlight") {              // spotlight.cc:64-76,282-306
        v3 from(0.f), to(0.f, 0.f, -1.f); col3 color(1.f); float power = 1.0f, angle = 45.f, falloff = 0.15f, ssfuzzy = 1.f;
        bool pOnly = false, soft = false; int smpl = 8;
        p.getPoint("from", from); p.getPoint("to", to); p.getColor("color", color); p.get("power", power); p.get("cone_angle", angle);
        p.get("blend", falloff); p.get("photon_only", pOnly); p.get("soft_shadows", soft); p.get("shadowFuzzyness", ssfuzzy); p.get("samples", smpl);
        l.type = LIGHT_SPOT; l.samples = smpl; l.position = from;
        l.ndir = normalize(from - to); l.direction = -l.ndir;
        l.color = color * power;
        createCS(l.direction, l.du, l.dv);
        const double rad_angle = (double)angle * 0.01745329251994329576922;
        const double rad_inner_angle = rad_angle * (double)(1.f - falloff);
        l.cosStart = fCos((float)rad_inner_angle);
        l.cosEnd = fCos((float)rad_angle);
        l.icosDiff = (float)(1.0 / (double)(l.cosStart - l.cosEnd));
        l.softShadows = soft ? 1 : 0; l.photonOnly = pOnly ? 1 : 0; l.shadowFuzzy = ssfuzzy;
    } else if (type == "spherelight") {            // spherelight.cc:62-70,194-213
        v3 from(0.f); col3 color(1.f); float power = 1.0f, radius = 1.f; int samples = 4, object = 0;
        p.getPoint("from", from); p.getColor("color", color); p.get("power", power); p.get("radius", radius); p.get("samples", samples); p.get("object", object);
        l.type = LIGHT_SPHERE; l.samples = samples; l.position = from; l.radius = radius;
        l.color = color * power;
        l.squareRadius = radius * radius;
        l.squareRadiusEps = (float)((double)l.squareRadius * 1.000003815);
        l.area = (float)((double)l.squareRadius * 4.0 * YB_M_PI);
        l.invArea = 1.f / l.area;
        if (object != 0) { fail("spherelight '" + std::string(name) + "': binding to a sphere primitive ('object') is not supported (triangle meshes only)"); return 0; }
    } else {
        fail("Environment: light type '" + type + "' is not supported by the B200 path (supported: pointlight, sunlight, arealight, meshlight, directional, spotlight, spherelight)");
        return 0;
    }
    lights.push_back(l); lightObj.push_back(obj);
    lightNames[name] = (int)lights.size() - 1;
    return 0x4c000000u | (uintptr_t)lights.size();
}

uintptr_t Context::createCamera(const char *name)   // environment.cc:292-330 + perspectiveCamera.cc:28-71,181-232; camera.h:41-61
{
    std::string type;
    if (!params.get("type", type)) { fail(std::string("Environment: Type of camera not specified for ") + name); return 0; }
    if (cameras.count(name)) { warn(std::string("Environment: Sorry, camera already in table: ") + name); return 0; }
    const bool persp = type == "perspective", archi = type == "architect", ortho = type == "orthographic", angular = type == "angular";
    if (!persp && !archi && !ortho && !angular) { fail("Environment: camera type '" + type + "' is not supported by the B200 path (supported: perspective, architect, orthographic, angular)"); return 0; }
    const ParaMap &p = params;
    v3 from(0, 1, 0), to(0, 0, 0), up(0, 1, 1);
    int resx = 320, resy = 200;
    float aspect = 1, dfocal = 1, apt = 0, dofd = 0, bkhrot = 0, nearClip = 0.0f, farClip = -1.0f;
    double aspect_d = 1.0, scale_d = 1.0, angle_d = 90, max_angle_d = 90; bool circular = true, mirrored = false;
    std::string bkhtype = "disk1", bkhbias = "uniform";
    p.getPoint("from", from); p.getPoint("to", to); p.getPoint("up", up); p.get("resx", resx); p.get("resy", resy);
    p.get("nearClip", nearClip); p.get("farClip", farClip);
    if (persp || archi) {
        p.get("focal", dfocal); p.get("aperture", apt); p.get("dof_distance", dofd); p.get("bokeh_type", bkhtype); p.get("bokeh_bias", bkhbias); p.get("bokeh_rotation", bkhrot);
        p.get("aspect_ratio", aspect);
    } else {      // orthographicCamera.cc:83-101, angularCamera.cc:73-99: double-typed aspect / scale / angle parameters
        p.get("aspect_ratio", aspect_d); aspect = (float)aspect_d;
        p.get("scale", scale_d);
        p.get("angle", angle_d); max_angle_d = angle_d; p.get("max_angle", max_angle_d); p.get("circular", circular); p.get("mirrored", mirrored);
    }
    CameraDesc d;
    memset(&d, 0, sizeof(d));
    Camera &c = d.cam;
    c.position = from; c.resx = resx; c.resy = resy;
    const float aspect_ratio = aspect * (float)resy / (float)resx;
    v3 camY = up - from, camZ = to - from;
    v3 camX = cross(camZ, camY);
    camY = cross(camZ, camX);
    camX = normalize(camX); camY = normalize(camY); camZ = normalize(camZ);
    c.camZ = camZ;
    c.near_p = from + camZ * nearClip;
    c.far_p = from + camZ * farClip;
    c.nearClip = nearClip; c.farClip = farClip;
    c.aspectRatio = aspect_ratio;
    if (ortho) {                                            // orthoCam_t::setAxis (orthographicCamera.cc:36-49)
        const float scale = (float)scale_d;
        c.camType = CAM_ORTHO;
        c.vright = camX; c.vup = aspect_ratio * camY; c.vto = camZ;
        c.opos = from - (float)(0.5 * scale) * (c.vup + c.vright);
        c.vup *= scale / (float)resy; c.vright *= scale / (float)resx;
        cameras[name] = d;
        return 0x43000000u | (uintptr_t)cameras.size();
    }
    if (angular) {                                          // angularCam_t ctor / setAxis / factory (angularCamera.cc:27-48,97-99)
        c.camType = CAM_ANGULAR;
        c.horPhi = (float)((float)angle_d * YB_M_PI / 180.f);
        c.circular = circular ? 1 : 0;
        c.vright = camX; c.vup = camY; c.vto = camZ;
        if (mirrored) c.vright *= -1.0f;
        c.maxR = (float)(max_angle_d / angle_d);
        cameras[name] = d;
        return 0x43000000u | (uintptr_t)cameras.size();
    }
    c.camType = CAM_PERSPECTIVE;
    c.aperture = apt; c.dof_distance = dofd; c.sampleLens = apt != 0;
    // bokeh shape / bias (perspectiveCamera.cc:215-225): unknown strings fall back to disk1 / uniform like the reference
    c.bkhtype = bkhtype == "disk2" ? 1 : bkhtype == "triangle" ? 3 : bkhtype == "square" ? 4 : bkhtype == "pentagon" ? 5 : bkhtype == "hexagon" ? 6 : bkhtype == "ring" ? 7 : 0;
    c.bkhbias = bkhbias == "center" ? 1 : bkhbias == "edge" ? 2 : 0;
    if (c.bkhtype >= 3 && c.bkhtype <= 6) {                 // lens polygon corners (perspectiveCamera.cc:40-50)
        int ns = c.bkhtype;
        float w = (float)((double)bkhrot * 0.01745329251994329576922), wi = (float)(6.28318530717958647692 / (double)(float)ns);
        ns = (ns + 2) * 2;
        for (int i = 0; i < ns; i += 2) { c.LS[i] = fCos(w); c.LS[i + 1] = fSin(w); w += wi; }
    }
    c.dof_rt = apt * camX; c.dof_up = apt * camY;
    c.vright = camX;
    c.vup = archi ? aspect_ratio * v3(0.f, 0.f, -1.f) : aspect_ratio * camY;     // architectCam_t::setAxis (architectCamera.cc:53-67)
    c.vto = (camZ * dfocal) - 0.5f * (c.vup + c.vright);
    c.vup = v3(c.vup.x / (float)resy, c.vup.y / (float)resy, c.vup.z / (float)resy);
    c.vright = v3(c.vright.x / (float)resx, c.vright.y / (float)resx, c.vright.z / (float)resx);
    cameras[name] = d;
    return 0x43000000u | (uintptr_t)cameras.size();
}

uintptr_t Context::createBackground(const char *name)   // environment.cc:332-365 + textureback.cc:205-235
{
    std::string type;
    if (!params.get("type", type)) { fail(std::string("Environment: Type of background not specified for ") + name); return 0; }
    if (backgrounds.count(name)) { warn(std::string("Environment: Sorry, background already in table: ") + name); return 0; }
    if (type != "constant" && type != "gradientback") { fail("Environment: background type '" + type + "' is not supported by the B200 path (supported: constant, gradientback)"); return 0; }
    col3 col(0.f); float power = 1.0f; bool ibl = false;
    params.getColor("color", col); params.get("power", power); params.get("ibl", ibl);
    if (ibl) { fail("background '" + std::string(name) + "': ibl=true (background light) is not supported"); return 0; }
    BackgroundDesc b;
    memset((void *)&b, 0, sizeof(b));
    b.color = col * power;
    if (type == "gradientback") {                 // gradientback.cc:82-115
        col3 szenith(0.4f, 0.5f, 1.f), shoriz(1.f);
        params.getColor("horizon_color", shoriz); params.getColor("zenith_color", szenith);
        col3 gzenith = szenith, ghoriz = shoriz;
        params.getColor("horizon_ground_color", ghoriz); params.getColor("zenith_ground_color", gzenith);
        b.gradient = 1; b.gzenith = gzenith * power; b.ghoriz = ghoriz * power; b.szenith = szenith * power; b.shoriz = shoriz * power;
    }
    backgrounds[name] = b;
    return 0x42000000u | (uintptr_t)backgrounds.size();
}

uintptr_t Context::createIntegrator(const char *name)   // environment.cc:367-400 + directlight.cc:184-257, pathtracer.cc:335-409
{
    std::string type;
    if (!params.get("type", type)) { fail(std::string("Environment: Type of integrator not specified for ") + name); return 0; }
    if (integrators.count(name)) { warn(std::string("Environment: Sorry, integrator already in table: ") + name); return 0; }
    const ParaMap &p = params;
    IntegratorDesc d;
    memset(&d, 0, sizeof(d));
    Integrator &I = d.p;
    if (type == "none") { d.surface = false; d.volume_none = true; }          // EmptyVolumeIntegrator.cc:21-33
    else if (type == "directlighting" || type == "pathtracing") {
        d.surface = true;
        bool transpShad = false, bg_transp = true, bg_transp_refract = true, useSSS = false;
        int shadowDepth = 5, raydepth = 5;
        p.get("raydepth", raydepth); p.get("transpShad", transpShad); p.get("shadowDepth", shadowDepth);
        p.get("bg_transp", bg_transp); p.get("bg_transp_refract", bg_transp_refract); p.get("useSSS", useSSS);
        if (useSSS) { fail("integrator '" + std::string(name) + "': useSSS (subsurface photon maps) is not supported"); return 0; }
        I.trShad = transpShad; I.sDepth = shadowDepth; I.rDepth = raydepth;
        I.transpBackground = bg_transp; I.transpRefractedBackground = bg_transp_refract;
        if (type == "directlighting") {
            bool caustics = false, do_AO = false; int AO_samples = 32; double AO_dist = 1.0; col3 AO_col(1.f);
            p.get("caustics", caustics); p.get("do_AO", do_AO); p.get("AO_samples", AO_samples); p.get("AO_distance", AO_dist); p.getColor("AO_color", AO_col);
            // caustic photon map (directlight.cc:204-231 defaults and keys)
            int search = 100, photons = 500000, cDepth = 10; double cRad = 0.25;
            p.get("photons", photons); p.get("caustic_mix", search); p.get("caustic_depth", cDepth); p.get("caustic_radius", cRad);
            I.usePhotonCaustics = caustics ? 1 : 0; I.nCausPhotons = photons; I.nCausSearch = search; I.causDepth = cDepth; I.causRadius = (float)cRad;
            I.type = INTEG_DIRECT; I.useAO = do_AO; I.aoSamples = AO_samples; I.aoDist = (float)AO_dist; I.aoCol = AO_col;
        } else {
            bool noRec = false; int path_samples = 32, bounces = 3; std::string cMethod;
            p.get("path_samples", path_samples); p.get("bounces", bounces); p.get("no_recursive", noRec);
            if (p.get("caustic_type", cMethod)) {
                // pathtracer.cc:366-385,100-113: "photon" = map only, "both" = map + caustic paths, "none", anything else = paths (default)
                const bool usePhotons = cMethod == "photon" || cMethod == "both";
                I.traceCaustics = (cMethod == "none" || cMethod == "photon") ? 0 : 1;
                if (usePhotons) {
                    int search = 100, photons = 500000, cDepth = 10; double cRad = 0.25;
                    p.get("photons", photons); p.get("caustic_mix", search); p.get("caustic_depth", cDepth); p.get("caustic_radius", cRad);
                    I.usePhotonCaustics = 1; I.nCausPhotons = photons; I.nCausSearch = search; I.causDepth = cDepth; I.causRadius = (float)cRad;
                }
            } else I.traceCaustics = 1;
            I.type = INTEG_PATH; I.nPaths = path_samples; I.maxBounces = bounces; I.noRecursive = noRec;
            // opt-in extension, not in the reference (SURVEY.md discrepancy 1): unbiased Russian roulette on path throughput
            bool rr = false; float rrMin = 0.05f;
            p.get("russian_roulette", rr); p.get("rr_min_prob", rrMin);
            I.russianRoulette = rr; I.rrMinProb = rrMin;
        }
    } else {
        fail("Environment: integrator type '" + type + "' is not supported by the B200 path (supported: directlighting, pathtracing, none)");
        return 0;
    }
    integrators[name] = d;
    return 0x49000000u | (uintptr_t)integrators.size();
}

// ------------------------------------------------------------------------------------------------ film (imagefilm.cc:59-165)
static float filt_box(float, float) { return 1.f; }
static float filt_mitchell(float dx, float dy)
{
    float x = 2.f * fSqrt(dx * dx + dy * dy);
    if (x >= 2.f) return 0.f;
    if (x >= 1.f) return (float)(x * (x * (x * -0.38888889f + 2.0f) - 3.33333333f) + 1.77777778f);
    return (float)(x * x * (1.16666666f * x - 2.0f) + 0.88888889f);
}
static float filt_gauss(float dx, float dy)
{
    float r2 = dx * dx + dy * dy;
    return std::max(0.f, float(fExp(-6 * r2) - 0.00247875));
}
static float filt_lanczos(float dx, float dy)
{
    float x = fSqrt(dx * dx + dy * dy);
    if (x == 0.f) return 1.f;
    if (-2 < x && x < 2) {
        float a = (float)YB_M_PI * x, b = (float)YB_M_PI_2 * x;
        return ((fSin(a) * fSin(b)) / (a * b));
    }
    return 0.f;
}

bool Context::setupRender(int width, int height)   // environment.cc:596-694 + createImageFilm :481-531
{
    const ParaMap &p = params;
    std::string name;
    if (!p.get("camera_name", name)) return fail("Environment: Specify a Camera!!");
    std::map<std::string, CameraDesc>::iterator ci = cameras.find(name);
    if (ci == cameras.end()) return fail("Environment: Specify an _existing_ Camera!!");
    if (!p.get("integrator_name", name)) return fail("Environment: Specify an Integrator!!");
    std::map<std::string, IntegratorDesc>::iterator ii = integrators.find(name);
    if (ii == integrators.end()) return fail("Environment: Specify an _existing_ Integrator!!");
    if (!ii->second.surface) return fail("Environment: Integrator is no surface integrator!");
    if (!p.get("volintegrator_name", name)) return fail("Environment: Specify a Volume Integrator!");
    std::map<std::string, IntegratorDesc>::iterator vi = integrators.find(name);
    if (vi == integrators.end() || !vi->second.volume_none)
        return fail("volume integrator '" + name + "' must be of type 'none' (participating media are not supported)");
    sc.integ = ii->second.p;
    sc.integ.hasBackground = 0; sc.integ.bgColor = col3(0.f); sc.integ.bgGradient = 0;
    if (p.get("background_name", name)) {
        std::map<std::string, BackgroundDesc>::iterator bi = backgrounds.find(name);
        if (bi == backgrounds.end()) { if (verbosity >= 1) fprintf(stderr, "ERROR: yb200: please specify an _existing_ Background!!\n"); }
        else {
            sc.integ.hasBackground = 1; sc.integ.bgColor = bi->second.color; sc.integ.bgGradient = bi->second.gradient;
            sc.integ.bgGZenith = bi->second.gzenith; sc.integ.bgGHoriz = bi->second.ghoriz; sc.integ.bgSZenith = bi->second.szenith; sc.integ.bgSHoriz = bi->second.shoriz;
        }
    }
    sc.cam = ci->second.cam;
    int AA_passes = 1, AA_samples = 1, AA_inc = 1; double AA_threshold = 0.05; bool z_chan = false, norm_z_chan = true;
    p.get("AA_passes", AA_passes); p.get("AA_minsamples", AA_samples); AA_inc = AA_samples; p.get("AA_inc_samples", AA_inc);
    p.get("AA_threshold", AA_threshold); p.get("z_channel", z_chan); p.get("normalize_z_channel", norm_z_chan);
    // createImageFilm
    int w = 320, h = 240, xstart = 0, ystart = 0, tileSize = 32; float filt_sz = 1.5f, gamma = 1.f; bool clamp = false, premult = false;
    std::string ftype; bool haveF = p.get("filter_type", ftype);
    p.get("gamma", gamma); p.get("clamp_rgb", clamp); p.get("AA_pixelwidth", filt_sz); p.get("width", w); p.get("height", h);
    p.get("xstart", xstart); p.get("ystart", ystart); p.get("tile_size", tileSize); p.get("premult", premult);
    if (width > 0 && height > 0 && (w != width || h != height)) return fail("render: output buffer size differs from the 'width'/'height' render parameters");
    Film &f = sc.film;
    memset(&f, 0, sizeof(f));
    f.w = w; f.h = h; f.cx0 = xstart; f.cy0 = ystart; f.cx1 = xstart + w; f.cy1 = ystart + h;
    f.filterw = filt_sz * 0.5f;
    float (*ffunc)(float, float) = filt_box;
    if (haveF) {
        if (ftype == "mitchell") { ffunc = filt_mitchell; f.filterw *= 2.6f; }
        else if (ftype == "lanczos") ffunc = filt_lanczos;
        else if (ftype == "gauss") { ffunc = filt_gauss; f.filterw *= 2.f; }
    }
    f.filterw = std::min(std::max(0.501f, f.filterw), 0.5f * YB_MAX_FILTER_SIZE);
    const float scale = 1.f / (float)YB_FILTER_TABLE_SIZE;
    for (int y = 0; y < YB_FILTER_TABLE_SIZE; ++y)
        for (int x = 0; x < YB_FILTER_TABLE_SIZE; ++x) f.filterTable[y * YB_FILTER_TABLE_SIZE + x] = ffunc((x + .5f) * scale, (y + .5f) * scale);
    f.tableScale = (float)(0.9999 * YB_FILTER_TABLE_SIZE / f.filterw);
    f.clamp = clamp; f.premult = premult; f.gamma = gamma > 0 ? 1.f / gamma : 1.f;   // imagefilm.cc:631-635
    f.correctGamma = (gamma > 0 && fabsf(1.f - gamma) > 0.001f) ? 1 : 0;
    f.tileSize = tileSize > 0 ? tileSize : 32;
    f.AA_passes = AA_passes; f.n_samples = AA_samples > 0 ? AA_samples : 1; f.pass_offs = 0;
    f.AA_inc_samples = AA_inc; f.AA_threshold = (float)AA_threshold;
    f.zChan = z_chan ? 1 : 0; f.zNorm = norm_z_chan ? 1 : 0; f.minDepth = 1e38f; f.maxDepth = 0.f;
    return true;
}

// ------------------------------------------------------------------------------------------------ flatten + build
bool Context::ensureDevice()
{
#ifndef YB_HOSTEMU
    if (device < 0) {
        int n = 0;
        cudaError_t e = cudaGetDeviceCount(&n);
        if (e != cudaSuccess || n == 0) return fail(std::string("no CUDA device available (") + cudaGetErrorString(e) + "): the B200 path has no CPU fallback");
        int cur = 0; cudaGetDevice(&cur); device = cur;
    }
    if (cudaSetDevice(device) != cudaSuccess) return fail("cudaSetDevice failed");
    if (!stream) { if (cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking) != cudaSuccess) return fail("cudaStreamCreate failed"); streamOwned = true; }
#endif
    return true;
}

// Material and light tables -> device.  Mesh lights are bound to their emitter here, as meshLight_t::init does inside
// scene_t::update (scene.cc:828-835, meshlight.cc:49-76): triangles {a, b, c, normal} and the area cdf of pdf1D_t
// (sample_utils.h:85-118) go into one light-geometry buffer, and the device pointers are patched into the Light records.
bool Context::uploadTables()
{
    std::vector<f4> ltris; std::vector<float> lcdf;
    std::vector<size_t> triOff(lights.size(), 0), cdfOff(lights.size(), 0);
    for (size_t i = 0; i < lights.size(); ++i) {
        Light &l = lights[i];
        if (l.type != LIGHT_MESH) continue;
        const std::map<unsigned, MeshObj> &allMeshes = meshSrc ? *meshSrc : meshes;
        std::map<unsigned, MeshObj>::const_iterator it = allMeshes.find(lightObj[i]);
        if (it == allMeshes.end() || it->second.base) return fail("meshlight: object " + std::to_string(lightObj[i]) + " is not a triangle mesh of this scene");
        const MeshObj &o = it->second;
        const size_t nt = o.tri_mat.size();
        if (nt == 0) return fail("meshlight: object " + std::to_string(lightObj[i]) + " has no triangles");
        triOff[i] = ltris.size(); cdfOff[i] = lcdf.size();
        std::vector<float> areas(nt);
        double totalArea = 0.0;
        for (size_t t = 0; t < nt; ++t) {
            v3 P[3];
            for (int c = 0; c < 3; ++c) {
                const int pi = o.tri_idx[3 * t + c];
                if (pi < 0 || (size_t)pi >= o.points.size()) return fail("triangle references a vertex that does not exist");
                P[c] = o.points[pi];
            }
            const v3 cr = cross(P[1] - P[0], P[2] - P[0]);
            const v3 n = normalize(cr);                                  // triangle_t::recNormal (triangle.cc:117-124)
            areas[t] = (float)(0.5 * (double)length(cr));                // triangle_t::surfaceArea (triangle.cc:144-154)
            totalArea += areas[t];
            f4 r; r.x = P[0].x; r.y = P[0].y; r.z = P[0].z; r.w = n.x; ltris.push_back(r);
            r.x = P[1].x; r.y = P[1].y; r.z = P[1].z; r.w = n.y; ltris.push_back(r);
            r.x = P[2].x; r.y = P[2].y; r.z = P[2].z; r.w = n.z; ltris.push_back(r);
        }
        l.area = (float)totalArea; l.invArea = (float)(1.0 / totalArea); l.meshCount = (int32_t)nt;
        double c = 0.0; const double delta = 1.0 / (double)nt;          // CumulateStep1dDF
        std::vector<float> cdf(nt + 1);
        cdf[0] = 0.f;
        for (size_t t = 1; t < nt + 1; ++t) { c += (double)areas[t - 1] * delta; cdf[t] = (float)c; }
        const float integral = (float)c;
        for (size_t t = 1; t < nt + 1; ++t) cdf[t] /= integral;
        lcdf.insert(lcdf.end(), cdf.begin(), cdf.end());
    }
    try {
        d_light_tris.reserve(ltris.size() + 1); d_light_cdf.reserve(lcdf.size() + 1);
        rt_h2d(d_light_tris.p, ltris.data(), ltris.size() * sizeof(f4), stream);
        rt_h2d(d_light_cdf.p, lcdf.data(), lcdf.size() * sizeof(float), stream);
        for (size_t i = 0; i < lights.size(); ++i)
            if (lights[i].type == LIGHT_MESH) { lights[i].meshTris = d_light_tris.p + triOff[i]; lights[i].meshCdf = d_light_cdf.p + cdfOff[i]; }
        d_materials.reserve(materials.size() + 1); d_lights.reserve(lights.size() + 1);
        rt_h2d(d_materials.p, materials.data(), materials.size() * sizeof(Material), stream);
        rt_h2d(d_lights.p, lights.data(), lights.size() * sizeof(Light), stream);
        d_tex_pixels.reserve(texPixels.size() + 1);
        rt_h2d(d_tex_pixels.p, texPixels.data(), texPixels.size() * sizeof(f4), stream);
        for (size_t i = 0; i < textures.size(); ++i)
            textures[i].pixels = (i < texPixOffset.size() && texPixOffset[i] != (size_t)-1) ? d_tex_pixels.p + texPixOffset[i] : 0;
        d_textures.reserve(textures.size() + 1); d_shaderNodes.reserve(shaderNodes.size() + 1);
        rt_h2d(d_textures.p, textures.data(), textures.size() * sizeof(Texture), stream);
        rt_h2d(d_shaderNodes.p, shaderNodes.data(), shaderNodes.size() * sizeof(ShaderNode), stream);
        sc.textures = d_textures.p; sc.shaderNodes = d_shaderNodes.p;
        rt_sync(stream);    // the staging vectors above go out of scope
    } catch (const RtError &e) { return fail(std::string("scene tables upload failed: ") + e.what()); }
    return true;
}

// scene_t::update (scene.cc:748-850): gather the triangles of every visible, non-base TRIM mesh in ascending
// object-id order (this defines the global primitive index used by the parity hook), bake instance transforms
// (meshtypes.h:133-143), upload as SoA, build the BVH on the device.
bool Context::flattenAndBuild()
{
    if (!ensureDevice()) return false;
    auto t0 = std::chrono::steady_clock::now();
    size_t ntri = 0; bool anySmooth = false;
    for (std::map<unsigned, MeshObj>::iterator i = meshes.begin(); i != meshes.end(); ++i) {
        MeshObj &o = i->second;
        if (!o.visible || o.is_base) continue;
        const MeshObj &g = o.base ? *o.base : o;
        ntri += g.tri_mat.size();
        if (o.is_smooth || (o.base && o.normals_exported)) anySmooth = true;
    }
    if (ntri >= 0x7fffffffu / 3) return fail("scene has too many triangles for 32-bit indexing");
    // texture coordinates are only flattened when some material has shader nodes (nothing else reads sp.U/V or sp.orcoP)
    bool anyNodes = false, anyUV = false, anyOrco = false;
    for (size_t i = 0; i < materials.size(); ++i) if (materials[i].nNodes > 0) anyNodes = true;
    if (anyNodes)
        for (std::map<unsigned, MeshObj>::iterator i = meshes.begin(); i != meshes.end(); ++i) {
            MeshObj &o = i->second;
            if (!o.visible || o.is_base) continue;
            const MeshObj &g = o.base ? *o.base : o;
            if (g.has_uv && g.uv_offsets.size() == 3 * g.tri_mat.size()) anyUV = true;
            if (g.has_orco) anyOrco = true;
        }
    bool anyInst = false;
    for (std::map<unsigned, MeshObj>::iterator i = meshes.begin(); i != meshes.end(); ++i)
        if (i->second.visible && !i->second.is_base && i->second.base) anyInst = true;
    std::vector<float> h_uv; std::vector<v3> h_orco, h_ng;
    if (anyInst) h_ng.assign(ntri, v3(0.f));
    if (anyUV) h_uv.assign(6 * ntri, 0.f);
    if (anyOrco) h_orco.assign(3 * ntri, v3(0.f));
    h_in_v.reserve(3 * ntri + 1); if (anySmooth) h_in_n.reserve(3 * ntri + 1);
    f4 *hv = h_in_v.p; v3 *hn = h_in_n.p;
    // work list: (object, first global triangle) in scene order, cut into chunks that host threads fill in parallel
    struct Chunk { const MeshObj *o; size_t t0, t1, k0; };
    std::vector<Chunk> chunks;
    {
        size_t k = 0;
        for (std::map<unsigned, MeshObj>::iterator i = meshes.begin(); i != meshes.end(); ++i) {
            MeshObj &o = i->second;
            if (!o.visible || o.is_base) continue;
            const size_t nt = (o.base ? *o.base : o).tri_mat.size();
            for (size_t t0 = 0; t0 < nt; t0 += 32768) { Chunk c = {&o, t0, std::min(nt, t0 + 32768), k + t0}; chunks.push_back(c); }
            k += nt;
        }
    }
    std::atomic<size_t> nextChunk(0);
    std::atomic<int> bad(0);           // 1: vertex index out of range, 2: normal index out of range
    auto fill = [&]() {
        for (;;) {
            const size_t ci = nextChunk.fetch_add(1);
            if (ci >= chunks.size()) return;
            const Chunk &ch = chunks[ci];
            const MeshObj &o = *ch.o;
            const MeshObj &g = o.base ? *o.base : o;
            const bool inst = o.base != 0;
            const float *m = o.m;
            const bool smooth = inst ? (o.is_smooth || o.normals_exported) : o.is_smooth;   // triangle.cc:18 vs :188
            size_t k = ch.k0;
            for (size_t t = ch.t0; t < ch.t1; ++t, ++k) {
                uint32_t info = g.tri_mat[t] & TRI_MAT_MASK;
                if (smooth) info |= TRI_SMOOTH;
                for (int c = 0; c < 3; ++c) {
                    const int pi = g.tri_idx[3 * t + c];
                    if (pi < 0 || (size_t)pi >= g.points.size()) { bad = 1; return; }
                    v3 pt = g.points[pi];
                    if (inst) pt = v3(m[0] * pt.x + m[1] * pt.y + m[2] * pt.z + m[3], m[4] * pt.x + m[5] * pt.y + m[6] * pt.z + m[7],
                                      m[8] * pt.x + m[9] * pt.y + m[10] * pt.z + m[11]);      // matrix4.h:92-97
                    f4 &d = hv[3 * k + c];
                    d.x = pt.x; d.y = pt.y; d.z = pt.z; d.w = 0.f;
                    if (anySmooth) {
                        const int ni = g.tri_nidx[3 * t + c];
                        const bool has = inst ? (ni > 0) : (ni >= 0);                        // triangle.cc:20-22 vs :190-192
                        v3 nn(0.f);
                        if (smooth && has) {
                            if ((size_t)ni >= g.normals.size()) { bad = 2; return; }
                            nn = g.normals[ni];
                            if (inst) nn = v3(m[0] * nn.x + m[1] * nn.y + m[2] * nn.z, m[4] * nn.x + m[5] * nn.y + m[6] * nn.z, m[8] * nn.x + m[9] * nn.y + m[10] * nn.z);
                            info |= (TRI_HAS_NA << c);
                        }
                        hn[3 * k + c] = nn;
                    }
                }
                if (anyUV && g.has_uv && g.uv_offsets.size() == 3 * g.tri_mat.size()) {   // triangle.cc:50-58
                    bool ok = true;
                    for (int c = 0; c < 3; ++c) {
                        const int ui = g.uv_offsets[3 * t + c];
                        if (ui < 0 || 2 * (size_t)ui + 1 >= g.uv_values.size()) { ok = false; break; }
                        h_uv[6 * k + 2 * c] = g.uv_values[2 * (size_t)ui]; h_uv[6 * k + 2 * c + 1] = g.uv_values[2 * (size_t)ui + 1];
                    }
                    if (!ok) { bad = 3; return; }
                    info |= TRI_HAS_UV;
                }
                if (anyOrco && g.has_orco) {                                             // triangle.cc:28-41: orco of vertex p is points[p + 1]
                    for (int c = 0; c < 3; ++c) {
                        const size_t pi = (size_t)g.tri_idx[3 * t + c] + 1;
                        if (pi >= g.points.size()) { bad = 1; return; }
                        h_orco[3 * k + c] = g.points[pi];
                    }
                    info |= TRI_HAS_ORCO;
                }
                if (inst) {      // triangleInstance_t::getNormal (triangle_inline.h:183-186): objToWorld * (base triangle's recNormal), normalised
                    const v3 bn = tri_rec_normal(g, t);
                    h_ng[k] = normalize(v3(m[0] * bn.x + m[1] * bn.y + m[2] * bn.z, m[4] * bn.x + m[5] * bn.y + m[6] * bn.z, m[8] * bn.x + m[9] * bn.y + m[10] * bn.z));
                    info |= TRI_EXPLICIT_NG;
                }
                hv[3 * k].w = i2f((int32_t)k);
                hv[3 * k + 1].w = i2f((int32_t)info);
            }
        }
    };
    {
        unsigned nth = std::thread::hardware_concurrency();
        nth = std::max(1u, std::min(std::min(nth, 16u), (unsigned)chunks.size()));
        std::vector<std::thread> pool;
        for (unsigned i = 1; i < nth; ++i) pool.emplace_back(fill);
        fill();
        for (size_t i = 0; i < pool.size(); ++i) pool[i].join();
    }
    if (bad == 1) return fail("triangle references a vertex that does not exist");
    if (bad == 2) return fail("triangle references a normal that does not exist");
    if (bad == 3) return fail("triangle references a UV value that does not exist");
    nTris = (uint32_t)ntri; haveNormals = anySmooth;
    const auto t1 = std::chrono::steady_clock::now();
    // upload + build
    DevBuf<f4> &in_v = d_in_v; DevBuf<v3> &in_n = d_in_n;
    try {
        in_v.reserve(3 * ntri + 1);
        rt_h2d(in_v.p, hv, 3 * ntri * sizeof(f4), stream);
        if (anySmooth) { in_n.reserve(3 * ntri + 1); rt_h2d(in_n.p, hn, 3 * ntri * sizeof(v3), stream); }
        sc.prim_uv = 0; sc.prim_orco = 0; sc.prim_ng = 0;
        if (anyInst) { d_prim_ng.reserve(h_ng.size() + 1); rt_h2d(d_prim_ng.p, h_ng.data(), h_ng.size() * sizeof(v3), stream); sc.prim_ng = d_prim_ng.p; }
        if (anyUV) { d_prim_uv.reserve(h_uv.size() + 1); rt_h2d(d_prim_uv.p, h_uv.data(), h_uv.size() * sizeof(float), stream); sc.prim_uv = d_prim_uv.p; }
        if (anyOrco) { d_prim_orco.reserve(h_orco.size() + 1); rt_h2d(d_prim_orco.p, h_orco.data(), h_orco.size() * sizeof(v3), stream); sc.prim_orco = d_prim_orco.p; }
        rt_sync(stream);
        const auto t2 = std::chrono::steady_clock::now();
        bvh_build(in_v.p, anySmooth ? in_n.p : 0, nTris, d_nodes, d_tri_v, d_tri_n, nNodes, bvhDepth, boundLo, boundHi, bvhScratch, stream);
        if (!uploadTables()) return false;
        rt_sync(stream);
        if (getenv("YB200_PROFILE_BUILD")) {
            const auto t3 = std::chrono::steady_clock::now();
            auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
            fprintf(stderr, "yb200 build: flatten %.1f ms, upload %.1f ms, bvh %.1f ms (%u tris)\n", ms(t0, t1), ms(t1, t2), ms(t2, t3), nTris);
        }
    } catch (const RtError &e) { return fail(std::string("scene update failed: ") + e.what()); }
    sc.nodes = d_nodes.p; sc.tri_v = d_tri_v.p; sc.tri_n = anySmooth ? d_tri_n.p : 0;
    sc.materials = d_materials.p; sc.lights = d_lights.p;
    sc.nLights = (int32_t)lights.size(); sc.nMaterials = (int32_t)materials.size();
    sc.nTris = nTris; sc.nNodes = nNodes; sc.boundLo = boundLo; sc.boundHi = boundHi;
    geomChanged = false;
    stats.n_triangles = nTris; stats.n_bvh_nodes = nNodes;
    stats.bvh_bytes = (uint64_t)nNodes * sizeof(BvhNode);
    stats.tri_bytes = (uint64_t)nTris * (3 * sizeof(f4) + (anySmooth ? 3 * sizeof(v3) : 0));
    stats.build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (nTris == 0) warn("Scene: Scene is empty...");
    return true;
}

// Helper of a multi-GPU context: take over the leader's scene.  Host tables are copied (they carry device pointers that are
// patched per device in uploadTables); the big buffers - BVH nodes, triangles in leaf order, vertex normals, per-primitive
// texture coordinates - come over NVLink with cudaMemcpyPeerAsync instead of a second flatten + upload + build.
bool Context::adoptScene(Context &L)
{
#ifdef YB_HOSTEMU
    (void)L;
    return fail("several devices are not part of the host emulator");
#else
    if (!ensureDevice()) return false;
    try {
        // direct NVLink DMA between the two devices (without it cudaMemcpyPeer stages through host memory)
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, device, L.device) == cudaSuccess && can) {
            cudaError_t e = cudaDeviceEnablePeerAccess(L.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) warn(std::string("peer access to the leader device: ") + cudaGetErrorString(e));
            cudaGetLastError();
        }
        verbosity = L.verbosity;
        materials = L.materials; textures = L.textures; texPixels = L.texPixels; texPixOffset = L.texPixOffset;
        shaderNodes = L.shaderNodes; lights = L.lights; lightObj = L.lightObj; meshSrc = &L.meshes;
        nTris = L.nTris; nNodes = L.nNodes; bvhDepth = L.bvhDepth; haveNormals = L.haveNormals; boundLo = L.boundLo; boundHi = L.boundHi;
        sc = L.sc;                                  // camera, film, integrator, counts; every pointer is replaced below
        struct Copy { void *dst; const void *src; size_t bytes; };
        std::vector<Copy> copies;
        d_nodes.reserve((size_t)nNodes + 1); copies.push_back({d_nodes.p, L.d_nodes.p, (size_t)nNodes * sizeof(BvhNode)});
        d_tri_v.reserve(3 * (size_t)nTris + 1); copies.push_back({d_tri_v.p, L.d_tri_v.p, 3 * (size_t)nTris * sizeof(f4)});
        sc.nodes = d_nodes.p; sc.tri_v = d_tri_v.p; sc.tri_n = 0; sc.prim_uv = 0; sc.prim_orco = 0; sc.prim_ng = 0;
        if (L.sc.tri_n) { d_tri_n.reserve(3 * (size_t)nTris + 1); copies.push_back({d_tri_n.p, L.d_tri_n.p, 3 * (size_t)nTris * sizeof(v3)}); sc.tri_n = d_tri_n.p; }
        if (L.sc.prim_uv) { d_prim_uv.reserve(6 * (size_t)nTris + 1); copies.push_back({d_prim_uv.p, L.d_prim_uv.p, 6 * (size_t)nTris * sizeof(float)}); sc.prim_uv = d_prim_uv.p; }
        if (L.sc.prim_orco) { d_prim_orco.reserve(3 * (size_t)nTris + 1); copies.push_back({d_prim_orco.p, L.d_prim_orco.p, 3 * (size_t)nTris * sizeof(v3)}); sc.prim_orco = d_prim_orco.p; }
        if (L.sc.prim_ng) { d_prim_ng.reserve((size_t)nTris + 1); copies.push_back({d_prim_ng.p, L.d_prim_ng.p, (size_t)nTris * sizeof(v3)}); sc.prim_ng = d_prim_ng.p; }
        for (size_t i = 0; i < copies.size(); ++i)
            if (copies[i].bytes) YB_CUDA_CHECK(cudaMemcpyPeerAsync(copies[i].dst, device, copies[i].src, L.device, copies[i].bytes, stream));
        if (!uploadTables()) return false;          // materials, lights (+ mesh-light geometry), textures, shader nodes; syncs the stream
        sc.materials = d_materials.p; sc.lights = d_lights.p;
        sc.nLights = (int32_t)lights.size(); sc.nMaterials = (int32_t)materials.size();
        stats.n_triangles = nTris; stats.n_bvh_nodes = nNodes;
        sceneReady = true; geomChanged = false; sceneStarted = true;
    } catch (const RtError &e) { return fail(std::string("scene copy to device ") + std::to_string(device) + " failed: " + e.what()); }
    return true;
#endif
}

bool Context::updateScene()
{
    if (!sceneStarted) return fail("no scene: call startScene first");
    if (stateStack.back() != ST_READY) return fail("scene update inside startGeometry/startTriMesh");
    if (stream && !ensureDevice()) return false;     // the caller's thread may have another device current
    if (!setupRender(0, 0)) return false;
    if (geomChanged || !sceneReady) { if (!flattenAndBuild()) return false; }
    else {
        try {   // materials / lights may have changed
            if (!uploadTables()) return false;
            sc.materials = d_materials.p; sc.lights = d_lights.p;
            sc.nLights = (int32_t)lights.size(); sc.nMaterials = (int32_t)materials.size();
        } catch (const RtError &e) { return fail(e.what()); }
    }
    sceneReady = true;
    return true;
}

}  // namespace yb
