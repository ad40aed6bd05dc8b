// yb_types.h — plain-data records shared by the host flattener and the kernels.
// Layout notes (HBM): see DESIGN.md §"Data layout".  All records are PODs that are memcpy'd to the device.
#pragma once
#include "yb_math.cuh"

namespace yb {

// BSDF_t bits — include/core_api/material.h:32-45
enum : uint32_t {
    BSDF_NONE = 0x0000, BSDF_SPECULAR = 0x0001, BSDF_GLOSSY = 0x0002, BSDF_DIFFUSE = 0x0004,
    BSDF_DISPERSIVE = 0x0008, BSDF_REFLECT = 0x0010, BSDF_TRANSMIT = 0x0020, BSDF_FILTER = 0x0040,
    BSDF_EMIT = 0x0080, BSDF_VOLUMETRIC = 0x0100, BSDF_TRANSLUCENT = 0x0200,
    BSDF_ALL_SPECULAR = BSDF_SPECULAR | BSDF_REFLECT | BSDF_TRANSMIT,
    BSDF_ALL_GLOSSY = BSDF_GLOSSY | BSDF_REFLECT | BSDF_TRANSMIT,
    BSDF_ALL = BSDF_SPECULAR | BSDF_GLOSSY | BSDF_DIFFUSE | BSDF_DISPERSIVE | BSDF_REFLECT | BSDF_TRANSMIT | BSDF_FILTER
};

struct alignas(16) f4 { float x, y, z, w; };
struct alignas(16) u4 { uint32_t x, y, z, w; };

enum MatType : int32_t { MAT_SHINYDIFFUSE = 0, MAT_GLOSSY, MAT_COATED_GLOSSY, MAT_GLASS, MAT_MIRROR, MAT_NULL, MAT_LIGHT, MAT_ROUGH_GLASS, MAT_BLEND };

// One record per material; the union of the constant parameters of the supported material_t classes
// (shader-node inputs: the node table below + mat_resolve in yb_bsdf.cuh).  ~350 B, read through the read-only cache.
struct Material {
    int32_t type;
    uint32_t bsdfFlags;
    // --- shinyDiffuseMat_t (include/materials/shinydiff.h:63-108; config() shinydiffuse.cc:32-80)
    col3 diffuseColor, mirrorColor, emitColor;
    float mirrorStrength, transparencyStrength, translucencyStrength, diffuseStrength, transmitFilterStrength;
    float iorSquared, orenA, orenB;
    int32_t hasFresnel, useOrenNayar, isMirror, isTransparent, isTranslucent, isDiffuse;
    int32_t nBSDF;
    uint32_t cFlags[4];
    int32_t cIndex[4];
    // --- glossyMat_t / coatedGlossyMat_t (glossy.cc:31-80, coatedglossy.cc:38-101)
    col3 glossColor, diffColor;     // mirrorColor above doubles as the coat's mirror_color
    float exponent, expU, expV, reflectivity, mDiffuse, ior;
    int32_t asDiffuse, withDiffuse, anisotropic, orenNayar;
    // --- glassMat_t / mirrorMat_t (glass.cc:30-82,300-340)
    col3 filterCol, specRefCol;
    int32_t fakeShadow;
    float mirrorRefl;               // mirrorMat_t::ref
    // --- lightMat_t (simple.cc:36-70)
    col3 lightCol;
    int32_t doubleSided;
    // --- roughGlassMat_t (roughglass.cc:30-43,268-285): ior, filterCol, specRefCol, fakeShadow above + GGX alpha^2
    float rgAlpha2;
    // --- blendMat_t (blend.cc:32-41,395-431): the two blended materials as offsets RELATIVE to this record in the
    //     material table (sub-materials are leaf materials: a blend of blends is rejected at creation)
    int32_t blendOff1, blendOff2;
    float blendVal;
    // glass / rough_glass with absorption: beer_t volume handler of the inside (beer.cc:24-45), sigma_a = -log(absorption) / absorption_dist
    int32_t hasBeer;
    col3 beerSigma;
    // --- nodeMaterial_t (nodematerial.cc:77-243): the material's shader nodes in evaluation order (allSorted) live at
    //     [nodeBase, nodeBase + nNodes) of the scene's node table; slot[s] = index (relative to nodeBase) of the node bound to
    //     shader slot s, -1 = the constant above is used.  emitStrength: shinydiffuse.cc:251-257 scales the diffuse NODE colour
    int32_t nodeBase, nNodes;
    int32_t slot[8];
    float emitStrength;
    uint32_t bumpMask;               // nodes of the bump_shader tree (nodeMaterial_t::bumpNodes), bit i = node i of this material
};
enum ShaderSlot : int32_t { SLOT_DIFFUSE = 0, SLOT_MIRROR_COLOR, SLOT_MIRROR, SLOT_TRANSPARENCY, SLOT_TRANSLUCENCY, SLOT_GLOSSY, SLOT_GLOSSY_REFLECT, SLOT_BUMP, SLOT_COUNT,
                            SLOT_BLEND_MASK = SLOT_DIFFUSE /* blend_mat has no other slot */ };
#define YB_MAX_NODES 8               // nodes per material (creation fails beyond)

// texture_t implementations evaluated on the device (src/textures/basictex.cc): rgb_cube, blend, clouds, marble, wood with the
// noise generators newPerlin_t (default), blenderNoise_t, voronoi_t (real distance), cellNoise_t (noise.cc)
enum TexType : int32_t { TEX_RGB_CUBE = 0, TEX_BLEND, TEX_CLOUDS, TEX_MARBLE, TEX_WOOD, TEX_IMAGE, TEX_VORONOI, TEX_MUSGRAVE, TEX_DISTNOISE };
enum : int32_t { TCL_EXTEND = 0, TCL_CLIP, TCL_CLIPCUBE, TCL_REPEAT, TCL_CHECKER };   // imagetex.h:31-38
enum : int32_t { TXP_PLAIN = 0, TXP_CUBE, TXP_TUBE, TXP_SPHERE };                 // basicnodes.h texture_mapper projection
struct Texture {
    int32_t type;
    int32_t depth, hard, bias, wshape, rings;   // wshape: 0 sin, 1 saw, 2 tri
    int32_t noiseType;                          // newNoise (basictex.cc:26-50): 0 newperlin, 1 blender, 2 cellnoise, 3..8 voronoi f1 f2 f3 f4 f2f1 crackle, 9 stdperlin
    float size, turb, sharpness;
    col3 color1, color2;
    // textureVoronoi_t (basictex.cc:250-318): weights, |weights|, intensity scale, colour type, distance metric (0 real, 1 squared -
    // also what 'manhattan' selects, noise.cc:453-456 -, 2 chebychev, 3 minkovsky 1/2, 4 minkovsky 4, 5 minkovsky e) and exponent
    float vw[4], vaw[4], iscale, mkExp;
    int32_t coltype, distMetric;
    // textureMusgrave_t (basictex.cc:346-383; 0 fBm, 1 multifractal, 2 heteroterrain, 3 hybridmf, 4 ridgedmf) and
    // textureDistortedNoise_t (:409-441: second noise basis, distortion)
    int32_t musType, noiseType2;
    float musH, musLacu, musOcts, musOffs, musGain, distort;
    // textureImage_t (src/textures/imagetex.cc:27-240): pixels = resx*resy RGBA floats of the image handler, row 0 on top
    const f4 *pixels;
    int32_t resx, resy, intp, isHDR, normalmap;          // intp: 0 none, 1 bilinear, 2 bicubic
    int32_t xrepeat, yrepeat, rot90, cropx, cropy, clipmode, checkerOdd, checkerEven;
    float gamma, cropminx, cropmaxx, cropminy, cropmaxy, checkerDist;
};
// shaderNode_t implementations (src/textures/basicnodes.cc, layernode.cc)
enum NodeType : int32_t { NODE_TEXMAP = 0, NODE_VALUE, NODE_MIX, NODE_LAYER };
enum : int32_t { TXC_UV = 0, TXC_GLOB, TXC_ORCO, TXC_TRAN };                     // texture_mapper texco (window / normal / ... rejected)
enum : uint32_t { TXF_RGBTOINT = 1, TXF_STENCIL = 2, TXF_NEGATIVE = 4, TXF_ALPHAMIX = 8 };   // layernode.h:11
struct ShaderNode {
    int32_t type;
    // texture_mapper (basicnodes.cc:96-160,236-304)
    int32_t tex, texCoords, mapX, mapY, mapZ, doScalar, mapType;
    v3 scale, offset;
    float bumpStr, dU, dV;           // textureMapper_t::setup (basicnodes.cc:12-38): step 0.0002, bump_strength / |scale| / 100
    float mtx[12];                   // first three rows of the 'transform' matrix (texco = transformed)
    // value (col1, val1) and the mix family (basicnodes.cc:312-600): inputs are node indices relative to nodeBase or -1
    float col1[4], col2[4];
    float val1, val2, cfactor;
    int32_t in1, in2, inFactor, mode;
    // layer (layernode.cc:12-110,190-225): input / upper layer, flags and factors
    int32_t input, upper;
    uint32_t texflag;
    int32_t doColor, doScalarL, colorInput, useAlpha;
    float colfac, valfac, defVal;
    col3 defCol;
    float upperCol[4];
    float upperVal;
};

enum LightType : int32_t { LIGHT_POINT = 0, LIGHT_SUN, LIGHT_AREA, LIGHT_DIRECTIONAL, LIGHT_SPOT, LIGHT_SPHERE, LIGHT_MESH };
struct Light {
    int32_t type, samples;
    col3 color;                      // point: col*power; area: col*power*pi; sun: col*power
    v3 position;                     // point
    v3 direction, du, dv;            // sun (du/dv from the UN-normalised direction, sunlight.cc:57); area du/dv unused
    col3 colPdf;                     // sun
    float pdf, invpdf;               // sun
    double cosAngle;                 // sun (double member, sunlight.cc:44)
    v3 corner, c2, c3, c4, toX, toY, normal, fnormal;   // area (arealight.cc:33-49)
    float area, invArea;
    // directional (directional.cc:52-61): direction, position, radius, infinite.  spot (spotlight.cc:64-76): position,
    // direction = cone axis, ndir, du/dv, cosStart/cosEnd/icosDiff, softShadows, shadowFuzzy.  sphere (spherelight.cc:62-70)
    v3 ndir;
    float radius, cosStart, cosEnd, icosDiff, shadowFuzzy, squareRadius, squareRadiusEps;
    int32_t infinite, softShadows, photonOnly;
    // mesh light (meshlight.cc:32-66): the emitter's triangles as 3 x f4 {a, n.x | b, n.y | c, n.z} and the area cdf of
    // pdf1D_t (count + 1 floats), both in the scene's light-geometry buffer; area / invArea / color above
    const f4 *meshTris; const float *meshCdf;
    int32_t meshCount, doubleSided;
};

struct Camera {                       // perspectiveCam_t (perspectiveCamera.cc:28-71,127-149; camera.h:41-61)
    v3 position, vto, vup, vright, dof_rt, dof_up, camZ;
    v3 near_p, far_p;                // plane points (plane normal = camZ)
    float aperture, dof_distance;
    int32_t resx, resy, sampleLens;
    int32_t bkhtype, bkhbias;        // bokehType / bkhBiasType values of perspectiveCamera.h:33-34
    float LS[16];                    // polygon lens corners (perspectiveCamera.cc:40-50)
    // camType: CAM_PERSPECTIVE also covers architectCam_t (same shootRay, vup = aspect * (0,0,-1), architectCamera.cc:53-67);
    // CAM_ORTHO (orthographicCamera.cc:36-64): opos = position - 0.5 scale (vup + vright), direction vto = camZ;
    // CAM_ANGULAR (angularCamera.cc:38-71): fisheye, wt = 0 outside max_r when circular
    int32_t camType, circular;
    v3 opos;
    float horPhi, maxR, aspectRatio;
    float nearClip, farClip;         // camera_t::getNearClip / getFarClip (camera.h:58-59,77-78)
};
enum CamType : int32_t { CAM_PERSPECTIVE = 0, CAM_ORTHO = 1, CAM_ANGULAR = 2 };

enum IntegType : int32_t { INTEG_DIRECT = 0, INTEG_PATH = 1 };
struct Integrator {
    int32_t type;
    int32_t trShad, sDepth, rDepth;
    int32_t useAO, aoSamples;
    float aoDist;
    col3 aoCol;
    int32_t nPaths, maxBounces, noRecursive, traceCaustics;
    int32_t transpBackground, transpRefractedBackground;
    int32_t hasBackground;
    col3 bgColor;
    // gradientBackground_t (gradientback.cc:60-80): colour by ray direction z; bgGradient = 0: constant bgColor
    int32_t bgGradient;
    col3 bgGZenith, bgGHoriz, bgSZenith, bgSHoriz;
    int32_t russianRoulette;          // opt-in extension (north_star); OFF by default and in every parity test
    float rrMinProb;
    // caustic photon map (mcintegrator.h: usePhotonCaustics / nCausPhotons / nCausSearch / causDepth / causRadius)
    int32_t usePhotonCaustics, nCausPhotons, nCausSearch, causDepth;
    float causRadius;
};

// photon_t as the caustic map keeps it (photon.h:67-133 without _SMALL_PHOTONS): position, incoming direction, power
struct Photon { v3 pos; v3 dir; col3 c; };
// the caustic photon map on the device (replaces photonMap_t + pointKdTree, photon.h / pkdtree.h): photons sorted by the cell of a
// uniform grid whose cells are at least the gather radius wide; cellStart[c] .. cellStart[c + 1] are the photons of cell c
struct PhotonGrid {
    const Photon *photons; const uint32_t *cellStart;
    v3 lo; float invCell;
    int32_t nx, ny, nz;
    uint32_t nPhotons;                // 0: no map (causticMap.ready() == false)
    int32_t nPaths, nSearch;          // photonMap_t::nPaths(), nCausSearch
    float radius;                     // causRadius
};

#define YB_FILTER_TABLE_SIZE 16
#define YB_MAX_FILTER_SIZE 8
struct Film {                         // imageFilm_t (imagefilm.cc:119-165,453-511,383-443)
    int32_t w, h, cx0, cy0, cx1, cy1;
    float filterw;
    float tableScale;                 // float member in the reference (imagefilm.h:161)
    float filterTable[YB_FILTER_TABLE_SIZE * YB_FILTER_TABLE_SIZE];
    int32_t clamp, premult, correctGamma;
    float gamma;
    int32_t tileSize;
    // sampling plan of this render pass (integrator.cc:229-290)
    int32_t AA_passes, n_samples, pass_offs;
    int32_t AA_inc_samples; float AA_threshold;      // adaptive passes (integrator.cc:158-163, imagefilm.cc:226-262)
    // depth channel (z_channel / normalize_z_channel, environment.cc:661-684; integrator.cc:99-130,313-334)
    int32_t zChan, zNorm;
    float minDepth, maxDepth;                        // tiledIntegrator_t::minDepth and the INVERSE range it keeps in maxDepth
};

// ---- geometry in HBM ------------------------------------------------------------------------------
// tri_v : 3 x float4 per triangle, in BVH leaf order.  v0.w carries the global primitive index of
//         scene.cc:760-781 (as int bits), v1.w the material/flags word, v2.w is free.
// tri_n : 3 x v3 vertex normals per triangle (same order), only when some mesh is smooth.
enum : uint32_t {
    TRI_MAT_MASK = 0x000fffffu,
    TRI_SMOOTH = 1u << 20,            // mesh->is_smooth
    TRI_HAS_NA = 1u << 21, TRI_HAS_NB = 1u << 22, TRI_HAS_NC = 1u << 23,   // na/nb/nc >= 0 (triangle.cc:20-22)
    TRI_HAS_UV = 1u << 24, TRI_HAS_ORCO = 1u << 25,                        // mesh->has_uv / has_orco (triangle.cc:28-48)
    TRI_EXPLICIT_NG = 1u << 26,       // instance triangle: Ng = normalize(objToWorld * base normal) in prim_ng (triangle_inline.h:183-186)
};

// 8-wide BVH node with child boxes quantised to 8 bits on a per-node power-of-two grid (80 B = 5 x 16 B).
struct alignas(16) BvhNode {
    float px, py, pz;                 // grid origin (node box lower corner)
    uint8_t ex, ey, ez;               // grid scale exponents: cell = 2^(e-127) as float bits (e<<23)
    uint8_t imask;                    // bit s set: child slot s is an internal node
    uint32_t childBase;               // index of the first internal child (children are stored contiguously in slot order)
    uint32_t triBase;                 // index of the first triangle of this node's leaf children
    uint8_t meta[8];                  // 0: empty; internal child in slot s: 0x20 | (24 + s); leaf: (unary tri count 1|3|7) << 5 | offset
                                      // from triBase (0..23).  So (meta >> 5) << (meta & 31) is the child's bit field in a 32-bit hit mask:
                                      // bits 0..23 = triangles of the node, bits 24..31 = internal children by slot (Ylitie et al. 2017)
    uint8_t qlox[8], qloy[8], qloz[8], qhix[8], qhiy[8], qhiz[8];
};
static_assert(sizeof(BvhNode) == 80, "BvhNode must be 80 bytes");

struct SceneDev {
    const BvhNode *nodes;
    const f4 *tri_v;
    const v3 *tri_n;                  // may be null
    const Material *materials;
    const Light *lights;
    // textures / shader nodes; per-primitive texture coordinates in the ORIGINAL primitive order of scene.cc:760-781
    // (tri_v[3 t].w): prim_uv = 6 floats {u1,v1,u2,v2,u3,v3}, prim_orco = the three orco points; null when no mesh has them
    const Texture *textures;
    const ShaderNode *shaderNodes;
    const float *prim_uv;
    const v3 *prim_orco;
    const v3 *prim_ng;      // geometric normal of instance triangles (TRI_EXPLICIT_NG), by primitive index; null without instances
    int32_t nLights, nMaterials;
    uint32_t nTris, nNodes;
    v3 boundLo, boundHi;
    Camera cam;
    Integrator integ;
    Film film;
    PhotonGrid photonGrid;
};

// ---- trace records ----------------------------------------------------------------------------------
struct Ray { v3 from; float tmin; v3 dir; float tmax; };   // 32 B; tmax < 0 == infinite (scene.cc:856,888)
struct Hit { int32_t tri; float t, b1, b2; };              // tri = index into tri_v order, -1 = miss

// ---- wavefront records ------------------------------------------------------------------------------
enum : uint32_t { TASK_INTEGRATE = 0, TASK_PATH = 1 };
// One pending scene_t::intersect + shading step: either a whole integrate(state, ray) call (camera ray or a
// recursiveRaytrace child) or the next vertex of a path of pathIntegrator_t::integrate's bounce loop.
struct Task {
    Ray ray;                          // 32
    col3 thr;                         // product of the weights above this call (linear in the result)
    float aw;                         // weight with which this call's alpha enters the camera sample's alpha
    uint32_t sid;                     // camera sample slot in the batch
    uint32_t bits;                    // kind:1 | includeLights:1 | caustic:1 | raylevel:6 | depth:8 | ...
    uint32_t rayDivision, rayOffset;  // renderState_t trajectory splitting (scene.h:86-88)
    float dc1, dc2;
    uint32_t pathOffs;                // `offs` of the path this vertex belongs to (pathtracer.cc:202)
    uint32_t pad;
};
static_assert(sizeof(Task) == 80, "Task is 80 bytes");
YB_HD uint32_t task_kind(uint32_t b) { return b & 1u; }
YB_HD bool task_lights(uint32_t b) { return (b >> 1) & 1u; }
YB_HD uint32_t task_level(uint32_t b) { return (b >> 3) & 63u; }
YB_HD uint32_t task_depth(uint32_t b) { return (b >> 9) & 255u; }
YB_HD uint32_t task_bits(uint32_t kind, bool lights, uint32_t level, uint32_t depth)
{
    return kind | (lights ? 2u : 0u) | (level << 3) | (depth << 9);
}

// One pending scene_t::isShadowed: if unoccluded, `c` (times the transparent-shadow filter) is added to the sample.
struct ShadowTask { Ray ray; col3 c; uint32_t sid; };     // 48 B
static_assert(sizeof(ShadowTask) == 48, "ShadowTask is 48 bytes");

}  // namespace yb
