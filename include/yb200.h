/* yb200.h — C ABI of libyb200.so: the B200-native ray/integrator hot path behind the
 * yafrayInterface_t / renderEnvironment_t API of inferrna/Core (TheBounty 0.1.6).
 *
 * Every entry point replaces one reference interface (cited as file:line relative to the
 * reference tree).  Plain pointers and sizes only; no C++/torch types cross this boundary.
 * All functions returning int return 1 on success and 0 on failure (the reference's `bool`);
 * on failure yb200_last_error() holds a "Y_ERROR"-style message.  Unsupported features
 * (unknown materials / lights, volumes, universal scene mode, photon maps ...) FAIL — there is no
 * CPU fallback anywhere in this library.
 */
#ifndef YB200_H
#define YB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct yb200_ctx yb200_ctx;
/* borrowed token, like the material_t* / light_t* the reference hands back
 * (include/interface/yafrayinterface.h:103-112); 0 == NULL == creation failed */
typedef uintptr_t yb200_handle;

/* ---- lifetime: yafrayInterface_t ctor/dtor, getYafray() (src/interface/yafrayinterface.cc:11-29,386-392) */
yb200_ctx  *yb200_create(void);
void        yb200_destroy(yb200_ctx *c);
const char *yb200_last_error(yb200_ctx *c);
const char *yb200_version(void);                               /* yafrayinterface.cc:313-320 getVersion */
int         yb200_set_device(yb200_ctx *c, int cuda_device);   /* new: which GPU this context renders on */
/* new: several GPUs of this node behind ONE context, so that the reference-facing call - yb200_render, i.e.
 * yafrayInterface_t::render - scales without the caller doing anything else.  devices[0] is the leader (scene
 * update, BVH build, final film); the other devices get the BVH and triangles by peer copy over NVLink, render the tiles
 * i, i + n, ... (imageSpliter_t order) on their own streams from their own host threads, and their raw films are summed on
 * the leader before imageFilm_t::flush.  n == 1 goes back to one device.  Applies to yb200_render / yb200_get_depth; the
 * explicit rank/world forms below stay per device. */
int         yb200_set_devices(yb200_ctx *c, const int *devices, int n);
void        yb200_set_verbosity(yb200_ctx *c, int level);      /* yafrayinterface.cc:360-384 (0 mute .. 3 info) */
void        yb200_clear_all(yb200_ctx *c);                     /* yafrayinterface.cc:47-64 */

/* ---- scene state machine: scene_t (src/yafraycore/scene.cc:95-135,266-319,522-662,362-520,983-1008) */
int      yb200_start_scene(yb200_ctx *c, int type);            /* yafrayinterface.cc:66-73; type!=0 fails (universal mode) */
int      yb200_start_geometry(yb200_ctx *c);                   /* scene.cc:95-100 */
int      yb200_end_geometry(yb200_ctx *c);                     /* scene.cc:102-117 */
unsigned yb200_get_next_free_id(yb200_ctx *c);                 /* scene.cc:950-966 */
int      yb200_start_trimesh(yb200_ctx *c, unsigned id, int vertices, int triangles,
                             int has_orco, int has_uv, int type); /* scene.cc:266-294 */
int      yb200_end_trimesh(yb200_ctx *c);                      /* scene.cc:296-319 */
int      yb200_add_vertex(yb200_ctx *c, double x, double y, double z);            /* scene.cc:525-542 */
int      yb200_add_vertex_orco(yb200_ctx *c, double x, double y, double z,
                               double ox, double oy, double oz);                  /* scene.cc:544-566 */
void     yb200_add_normal(yb200_ctx *c, double nx, double ny, double nz);         /* scene.cc:568-583 */
int      yb200_add_triangle(yb200_ctx *c, int a, int b, int cc, yb200_handle mat);/* scene.cc:585-626 */
int      yb200_add_triangle_uv(yb200_ctx *c, int a, int b, int cc, int uv_a, int uv_b, int uv_c,
                               yb200_handle mat);                                  /* scene.cc:628-646 */
int      yb200_add_uv(yb200_ctx *c, float u, float v);                            /* scene.cc:648-662 */
int      yb200_smooth_mesh(yb200_ctx *c, unsigned id, double angle);              /* scene.cc:362-520 */
int      yb200_add_instance(yb200_ctx *c, unsigned base_id, const float m[16]);   /* scene.cc:983-1008; row-major 4x4 */
/* bulk SoA forms of the three calls above (same semantics as calling them n times) */
int      yb200_add_vertices(yb200_ctx *c, const float *xyz, const float *normals_or_null, int n);
int      yb200_add_triangles(yb200_ctx *c, const int *abc, int n, yb200_handle mat);

/* ---- parameter maps: paraMap_t (include/core_api/params.h:40-125; yafrayinterface.cc:146-222) */
void yb200_params_set_point (yb200_ctx *c, const char *name, double x, double y, double z);
void yb200_params_set_string(yb200_ctx *c, const char *name, const char *s);
void yb200_params_set_bool  (yb200_ctx *c, const char *name, int b);
void yb200_params_set_int   (yb200_ctx *c, const char *name, int i);
void yb200_params_set_float (yb200_ctx *c, const char *name, double f);
void yb200_params_set_color (yb200_ctx *c, const char *name, float r, float g, float b, float a);
void yb200_params_set_matrix(yb200_ctx *c, const char *name, const float m[16], int transpose);
void yb200_params_clear_all (yb200_ctx *c);
void yb200_params_start_list(yb200_ctx *c);
void yb200_params_push_list (yb200_ctx *c);
void yb200_params_end_list  (yb200_ctx *c);

/* ---- factories: renderEnvironment_t::create* (src/yafraycore/environment.cc:189-479) keyed by "type" */
yb200_handle yb200_create_light     (yb200_ctx *c, const char *name);  /* pointlight | sunlight | arealight | meshlight | directional | spotlight | spherelight */
yb200_handle yb200_create_material  (yb200_ctx *c, const char *name);  /* shinydiffusemat | glossy | coated_glossy | glass | rough_glass | blend_mat | mirror | null | light_mat; shader nodes = the parameter list */
yb200_handle yb200_create_camera    (yb200_ctx *c, const char *name);  /* perspective | architect | orthographic | angular */
yb200_handle yb200_create_background(yb200_ctx *c, const char *name);  /* constant, gradientback */
yb200_handle yb200_create_integrator(yb200_ctx *c, const char *name);  /* directlighting | pathtracing | none (volume) */
yb200_handle yb200_create_texture   (yb200_ctx *c, const char *name);  /* rgb_cube | blend | clouds | marble | wood | image (basictex.cc, imagetex.cc) */
/* Decoded pixels for an image texture, keyed by the `filename` the texture's parameters name: what imageHandler_t::getPixel
 * (include/core_api/imagehandler.h:44) returns for that file - width*height RGBA floats, row 0 on top, index x + width*y;
 * is_hdr = imageHandler_t::isHDR (no gamma).  The reference decodes files in its image-handler plugins; this library only
 * reads TGA (24/32-bit true colour, 8-bit grey) itself, every other format has to be registered before yb200_create_texture. */
int yb200_register_image(yb200_ctx *c, const char *file_name, const float *rgba, int width, int height, int is_hdr);
yb200_handle yb200_create_volume_region(yb200_ctx *c, const char *name); /* always fails */

/* ---- render: yafrayInterface_t::render + scene_t::render + imageFilm_t::flush
 *      (yafrayinterface.cc:336-342, scene.cc:931-945, imagefilm.cc:383-443).
 * The render parameters are the current paraMap (camera_name, integrator_name, volintegrator_name,
 * background_name, width, height, xstart, ystart, AA_*, filter_type, AA_pixelwidth, gamma, clamp_rgb,
 * premult, tile_size ...; environment.cc:493-505,607-664).
 * rgba: caller-owned HOST buffer, memoryIO_t layout (src/yafraycore/memoryIO.cc:27-43):
 *       width*height*4 floats, index (x + width*y)*4.  Callbacks mirror colorOutput_t/progressBar_t. */
typedef struct yb200_callbacks {
    void *user;
    int  (*should_abort)(void *user);                       /* scene_t::abort flag poll (scene.cc:71-85) */
    void (*progress)(void *user, int done, int total);      /* progressBar_t::update (monitor.h:24-36) */
} yb200_callbacks;
int yb200_render(yb200_ctx *c, float *rgba, int width, int height, const yb200_callbacks *cb);

/* Multi-GPU form: scene update + accumulate this rank's image tiles into a DEVICE film of
 * width*height*5 floats (r,g,b,a,weight raw sums; pixel_t, include/utilities/image_buffers.h:28-43)
 * laid out (x + width*y)*5.  Tiles (imageSpliter_t, imagesplitter.cc:29-52) are dealt round-robin:
 * tile i belongs to rank i % world.  The caller sums the films of all ranks (NCCL reduce) and calls
 * yb200_film_resolve on the sum.  dev_film must be zeroed by the caller. */
int yb200_render_film(yb200_ctx *c, float *dev_film, int rank, int world, const yb200_callbacks *cb);
/* One anti-aliasing pass of the above (tiledIntegrator_t::render's pass loop, integrator.cc:153-164), for adaptive
 * sampling (AA_passes > 1) on several ranks: pass 0 renders AA_minsamples on every owned pixel; pass i >= 1 renders
 * AA_inc_samples on the owned pixels that imageFilm_t::nextPass (imagefilm.cc:213-263) flags from dev_film_all = the raw
 * film sums of ALL ranks after pass i-1 (NULL: use dev_film, i.e. world == 1).  yb200_render_film runs all passes itself
 * when world == 1 and fails for world > 1 with AA_passes > 1. */
int yb200_render_film_pass(yb200_ctx *c, float *dev_film, const float *dev_film_all, int rank, int world, int pass,
                           const yb200_callbacks *cb);
/* normalise / clamp / gamma / premult exactly like imageFilm_t::flush; dev_film device, rgba HOST or DEVICE */
int yb200_film_resolve(yb200_ctx *c, const float *dev_film, float *rgba, int rgba_on_device);
/* Depth channel (render parameters z_channel / normalize_z_channel, environment.cc:661-685): what the reference hands to
 * colorOutput_t::putPixel(x, y, c, alpha, depth = true, z) (imagefilm.cc:360-362,425-427) - the filtered per-pixel depth of
 * imageFilm_t::addDepthSample (imagefilm.cc:513-564), normalised to [0, 1] by tiledIntegrator_t::precalcDepths
 * (integrator.cc:99-130) unless normalize_z_channel is false.  z: HOST buffer of width*height floats, index x + width*y.
 * yb200_get_depth reads the depth film of the last yb200_render / yb200_render_film of this context; with several ranks,
 * sum the raw depth films (yb200_depth_film: DEVICE pointer, width*height*2 floats = sum(depth*w), sum(w)) like the colour
 * film and resolve the sum with yb200_depth_resolve. */
int    yb200_get_depth(yb200_ctx *c, float *z);
float *yb200_depth_film(yb200_ctx *c);
int    yb200_depth_resolve(yb200_ctx *c, const float *dev_depth_film, float *z);
void yb200_abort(yb200_ctx *c);                              /* yafrayinterface.cc:278 */

/* ---- scene update only (scene_t::update, scene.cc:748-850): flatten + GPU BVH build */
int yb200_update_scene(yb200_ctx *c);
/* mark the geometry as changed (what every geometry call does through state.changes |= C_GEOM, scene.cc:288):
 * the next render/update re-flattens, re-uploads and rebuilds the BVH.  Used to time the full host-to-host path. */
void yb200_touch_geometry(yb200_ctx *c);

/* ---- parity / benchmark hook for triKdTree_t::Intersect / IntersectS / IntersectTS
 *      through scene_t::intersect / isShadowed (scene.cc:852-928, kdtree.cc:675-1108).
 * rays: n records of 8 floats {from.xyz, tmin, dir.xyz, tmax}; tmax<0 = infinite (scene.cc:856,888).
 * mode YB200_TRACE_CLOSEST: out = n records {int32 prim, float t, float b1, float b2} (prim=-1: miss);
 *      prim indexes the global triangle order of scene.cc:760-781.
 * mode YB200_TRACE_SHADOW:  out = n int32 (1 shadowed / 0 not)           — isShadowed(state,ray)
 * mode YB200_TRACE_TSHADOW: out = n records {int32 shadowed, float r,g,b} — isShadowed(state,ray,depth,filt)
 * on_device: rays/out are device pointers (no copies, asynchronous on the context stream). */
enum { YB200_TRACE_CLOSEST = 0, YB200_TRACE_SHADOW = 1, YB200_TRACE_TSHADOW = 2 };
int yb200_trace_batch(yb200_ctx *c, const float *rays, size_t n, int mode, int shadow_depth,
                      void *out, int on_device);

/* Allocation-free form of yb200_trace_batch for callers that must not allocate (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same rays / out records / modes as yb200_trace_batch, except YB200_TRACE_TSHADOW (rejected).  The entry
 * point calls no cudaMalloc / cudaFree or allocating library routine: the ray-fetch counter of the traversal kernel and,
 * for HOST pointers, the device staging of rays and results live in one file-scope device array of
 * YB200_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.
 * Limits (a larger n fails with a message in yb200_last_error, nothing is traced):
 *   on_device == 0: n <= YB200_STATIC_MAX_RAYS_HOST (the staging area); copies in, traces, copies out, waits;
 *   on_device != 0: n <= YB200_STATIC_MAX_RAYS_DEVICE (one launch of the 32-bit counter); asynchronous on the context stream.
 * Needs a scene already on the device (yb200_update_scene / a render - those do allocate).  The array is shared by every
 * call on that GPU and keeps the last call's contents; the counter is cleared on the stream before each launch.  Calls that
 * overlap on the GPU (other streams, contexts or threads) race on it: serialise them. */
#define YB200_STATIC_MAX_RAYS_HOST   65536u
#define YB200_STATIC_MAX_RAYS_DEVICE (1u << 30)
#define YB200_STATIC_SCRATCH_BYTES   (256u + YB200_STATIC_MAX_RAYS_HOST * (32u + 16u))
int yb200_trace_batch_static(yb200_ctx *c, const float *rays, size_t n, int mode, void *out, int on_device);

/* ---- device known-answer hook (TEST INFRASTRUCTURE, no counterpart in the reference's interface): one GPU thread
 * per record runs the device functions the path is built from, so tests can compare them with vectors generated by the compiled
 * reference.  in / out are HOST buffers.
 *  YB200_KAT_QMC : in n x {uint32 kind, a, b, c}, out n x double.  kind 0 scrHalton(dim a, n b) (scr_halton.h:47-69); 1 RI_vdC(a, b);
 *                  2 RI_S(a, b); 3 RI_LP(a, b) (mcqmc.h:100-122); 4 fnv_32a_buf(a) (mcqmc.h:155-168); 5 Halton(base a in {2,3,5})
 *                  .setStart(b), value of call c + 1 of getNext() (mcqmc.h:29-94)
 *  YB200_KAT_MATH: in n x {uint32 fn, float x, float y, pad}, out n x float.  fn 0 fSin, 1 fCos, 2 fPow(x, y), 3 fExp, 4 fLog, 5 fSqrt
 *                  (mathOptimizations.h:100-296)
 *  YB200_KAT_BSDF: needs an updated scene.  in n x 16 words {uint32 material index (creation order), float N.xyz, Ng.xyz, wo.xyz,
 *                  wi.xyz, s1, s2, uint32 bsdf flags}; out n x 16 floats {eval.rgb, pdf, sample(): wi.xyz, colour.rgb, W, s.pdf,
 *                  uint32 sampledFlags, uint32 flags of initBSDF, 0, 0}  (material_t::initBSDF / eval / pdf / sample, material.h:91-192) */
enum { YB200_KAT_QMC = 0, YB200_KAT_MATH = 1, YB200_KAT_BSDF = 2 };
int yb200_selftest(yb200_ctx *c, int what, const void *in, size_t n, void *out);

/* Dependent-load latency of this GPU (measurement hook for the latency roofline of SURVEY §8d; no reference counterpart): one
 * thread walks a random cycle through bytes / 128 cache lines (ld.global.cg) for `steps` loads.  bytes well above the L2 size
 * gives the L2-miss (HBM) latency t_L2miss, a few MB the L2-hit latency.  Allocates and frees `bytes` of device memory. */
int yb200_probe_latency(yb200_ctx *c, size_t bytes, int steps, double *ns_per_load);

/* ---- counters (new: the reference has none) */
typedef struct yb200_stats {
    uint64_t n_triangles, n_bvh_nodes;
    uint64_t rays_closest, rays_shadow;        /* scene_t::intersect / isShadowed calls of the last render */
    uint64_t camera_samples, kernel_launches;
    double   build_ms, render_ms, trace_closest_ms, trace_shadow_ms, shade_ms, film_ms;
    uint64_t bvh_bytes, tri_bytes;
} yb200_stats;
int yb200_get_stats(yb200_ctx *c, yb200_stats *out);
void *yb200_cuda_stream(yb200_ctx *c);          /* cudaStream_t the context launches on */

#ifdef __cplusplus
}
#endif
#endif /* YB200_H */
