/* srt.h — C ABI of the B200-native path-tracing device ("libsrt_b200.so").
 *
 * This is the drop-in boundary for ONE path of swishersnaaake/sexy-raytracer: the per-pixel
 * path-tracing loop (reference main.cpp:200-227 -> camera.h:40-50 -> main.cpp:33-52 -> bvh.h:97-105
 * -> aabb.h:11-27 / model.h:104-181 / sphere.h:54-83 -> material.h:91-245 -> texture.h:26-148
 * -> color.h:25-41).
 *
 * The reference has no FFI; the only device boundary it ever sketched is the per-frame one of
 * gl.h:28-31
 *     bool glDevice::init(int w, int h, std::vector<hittableIndexed>&);
 *     bool glDevice::rtFrame(void* frameData, int w, int h, std::vector<hittableIndexed>&);
 *     void glDevice::terminate();
 * fed by hittableVector::build (hittablevector.h:27-31, main.cpp:148-149) and wrapped around the pixel
 * loop (main.cpp:190-199,228-232).  The entry points below have that shape (hand a flattened scene
 * over once, then ask for frames), with the 160-byte hittableIndexed record (hittableindexed.h:24-38)
 * replaced by SoA arrays + 32-byte BVH nodes in the same pre-order (bvh.h:112-148).
 *
 * Conventions: plain C structs, plain pointers and sizes, int status codes (0 = OK), no exceptions
 * cross the boundary.  The caller owns every host pointer; the library owns all device memory.
 * Calls block and are not re-entrant per srt_device (the reference is single threaded).
 * Pointers named *_dev are device pointers on the srt_device's GPU; everything else is host memory.
 * There is no CPU fallback: every compute entry point fails with SRT_ERR_CUDA when no GPU is present.
 */
#ifndef SRT_H_
#define SRT_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SRT_ABI_VERSION 1

typedef int srt_status;
enum {
  SRT_OK = 0,
  SRT_ERR_INVALID = 1,      /* bad argument / malformed scene description            */
  SRT_ERR_CUDA = 2,         /* CUDA runtime error (srt_last_error() has the text)     */
  SRT_ERR_UNSUPPORTED = 3,  /* valid but outside what the device implements           */
  SRT_ERR_NOMEM = 4
};

/* ---- primitives ------------------------------------------------------------------------------
 * Primitive ids are global: triangles first [0, n_tri), then spheres [n_tri, n_tri + n_sph)
 * (main.cpp:82-86 adds the mesh triangles, main.cpp:90,127,140,144 the spheres).               */

/* One BVH node, 32 bytes, array in the pre-order of bvhNode::populateVector (bvh.h:112-148):
 * a node's index is assigned before its children, left subtree first.
 * child >= 0 : index of an inner node;  child < 0 : leaf, primitive id = ~child.
 * right == SRT_BVH_NONE : the reference's span-1 node (left == right, bvh.h:67-69) - its leaf is
 * visited once (the second visit of the reference cannot change the record).                  */
typedef struct srt_bvh_node {
  float   bmin[3];
  float   bmax[3];
  int32_t left;
  int32_t right;
} srt_bvh_node;
#define SRT_BVH_NONE INT32_MIN

/* ---- textures (texture.h) ------------------------------------------------------------------- */
enum {
  SRT_TEX_SOLID = 0,    /* solidColor (texture.h:18-32): value = color                          */
  SRT_TEX_CHECKER = 1,  /* checker (texture.h:34-52): sinf(10x)sinf(10y)sinf(10z) < 0 ? odd : even, x255 */
  SRT_TEX_IMAGE = 2     /* imagePNG / image3bpp (texture.h:54-154): nearest texel, 0..255 floats */
};
typedef struct srt_texture {
  int32_t        kind;
  float          color[3];     /* SOLID                                                         */
  int32_t        even, odd;    /* CHECKER: texture ids (must be SOLID or IMAGE)                 */
  int32_t        width, height, bpp;  /* IMAGE: bpp = bytes per texel as passed to imagePNG     */
  const uint8_t* data;         /* IMAGE: height*width*bpp bytes, row 0 first; NULL = the file
                                  was missing -> value() = (1,0,1) (texture.h:130-131)          */
} srt_texture;

/* ---- materials (material.h) ----------------------------------------------------------------- */
enum {
  SRT_MAT_PBR = 0,         /* pbrMetallicRoughness (material.h:23-85,156-245)                   */
  SRT_MAT_METAL = 1,       /* metal (material.h:87-102)                                         */
  SRT_MAT_DIELECTRIC = 2,  /* dielectric (material.h:104-137)                                   */
  SRT_MAT_LIGHT = 3        /* diffuseLight (material.h:139-154)                                 */
};
typedef struct srt_material {
  int32_t kind;
  int32_t albedo_tex;     /* PBR albedoMap / LIGHT emit texture; -1 = none                      */
  int32_t normal_tex;     /* PBR normalMap;    -1 = none                                        */
  int32_t metallic_tex;   /* PBR metallicMap;  -1 = none                                        */
  int32_t roughness_tex;  /* PBR roughnessMap; -1 = none (metallicRoughnessMap is never sampled
                             by the reference, material.h:78, so it has no slot)                */
  float   albedo[4];      /* PBR albedo factor (vec4f) / METAL albedo (rgb)                     */
  float   metalness;      /* PBR scalar used when metallic_tex < 0                              */
  float   roughness;      /* PBR scalar used when roughness_tex < 0                             */
  float   fuzz;           /* METAL (already clamped to <= 1, material.h:89)                     */
  float   ir;             /* DIELECTRIC index of refraction                                     */
  int32_t flags;          /* SRT_MATF_*; 0 = the reference's behaviour                          */
} srt_material;
/* PBR, NOT the reference's behaviour (off unless the loader is told to): metallic_tex is a glTF 2.0
   metallicRoughnessTexture and IS sampled - metalness = blue/255 * metalness, roughness = green/255 * roughness;
   roughness_tex is ignored.  (The reference stores this map and never samples it, material.h:78.) */
#define SRT_MATF_PACKED_MR 1

/* ---- scene ---------------------------------------------------------------------------------- */
typedef struct srt_scene_desc {
  /* triangles: SoA, 3 floats per vertex position, 2 floats per vertex uv (model.h:71-72 resolved
     through the uint16 indices of model.h:46)                                                  */
  int64_t        n_tri;
  const float*   tri_v0;   /* [n_tri*3] */
  const float*   tri_v1;
  const float*   tri_v2;
  const float*   tri_uv0;  /* [n_tri*2] */
  const float*   tri_uv1;
  const float*   tri_uv2;
  const int32_t* tri_mat;  /* [n_tri]   */
  /* spheres (sphere.h:40-44) */
  int64_t        n_sph;
  const float*   sph_c0;   /* [n_sph*3] center0 */
  const float*   sph_c1;   /* [n_sph*3] center1 */
  const float*   sph_t0;   /* [n_sph]   */
  const float*   sph_t1;   /* [n_sph]   */
  const float*   sph_r;    /* [n_sph]   */
  const int32_t* sph_mat;  /* [n_sph]   */
  /* BVH: root is node 0 */
  int64_t             n_nodes;
  const srt_bvh_node* nodes;
  /* tables */
  int32_t             n_mat;
  const srt_material* materials;
  int32_t             n_tex;
  const srt_texture*  textures;
  /* optional per-vertex normals, [n_tri*3] each, all three or none (NULL = flat face normals, which is what the reference
     does: glTF NORMAL is never read, model.h:335-365).  When given, the shading normal of a triangle hit is
     unit(r0 n0 + r1 n1 + r2 n2) with the reference's own inverse-distance weights r_k (model.h:158-166), NOT the
     reference's behaviour; frontFace and the tangent basis stay those of the face. */
  const float*   tri_n0;
  const float*   tri_n1;
  const float*   tri_n2;
  /* optional per-vertex tangents (glTF TANGENT: xyz + handedness w = +-1), [n_tri*4] each, all three or none (NULL = the
     reference's per-face basis from the UV deltas, calcTangentBasis, model.h:214-235: glTF TANGENT is never read, model.h:335-365).
     When given, a triangle whose vertices carry a non-zero w is shaded with tangent = unit(r0 t0 + r1 t1 + r2 t2) (the weights of
     model.h:158-166) and bitangent = w0 * unit(shading normal x tangent); triangles whose three w are 0, and hits where that
     basis is not finite, keep the per-face basis.  NOT the reference's behaviour. */
  const float*   tri_tan0;
  const float*   tri_tan1;
  const float*   tri_tan2;
} srt_scene_desc;

/* camera: the ten derived members of camera.h:52-62 (private there; the host class exposes them) */
typedef struct srt_camera {
  float origin[3];
  float lleft[3];
  float horizontal[3];
  float vertical[3];
  float w[3];
  float hor[3];
  float vert[3];
  float lens_radius;
  float t0, t1;
} srt_camera;

/* ray.h:6-23 */
typedef struct srt_ray {
  float o[3];
  float dir[3];
  float time;
} srt_ray;

/* hitRecord (hittable.h:9-22) + the primitive id the reference does not keep */
typedef struct srt_hit {
  int32_t prim;        /* -1 = miss                                                             */
  int32_t mat;
  int32_t front_face;
  float   t;
  float   p[3];
  float   normal[3];   /* face-forwarded (hittable.h:18-21)                                     */
  float   tangent[3];
  float   bitangent[3];
  float   uv[2];
} srt_hit;

typedef struct srt_render_params {
  int32_t  width, height;
  int32_t  spp;            /* numSamples of the whole frame (main.cpp:178): the divisor of resolve */
  int32_t  sample_begin;   /* this call renders global samples [sample_begin, sample_begin+sample_count) */
  int32_t  sample_count;   /* of every pixel (spp split across GPUs); 0 = all spp      */
  int32_t  max_bounce;     /* main.cpp:180; 0 = every path is black and no segment is traced (main.cpp:36-37) */
  float    background[3];  /* main.cpp:170                                                       */
  uint64_t seed;           /* Philox key                                                         */
  int32_t  rr_begin;       /* Russian roulette from this bounce on; 0 = off (the reference has none) */
  /* variant 2 adds every finished path to its pixel with red.global.add.f32: the order of a pixel's samples depends on the
     schedule, so two runs of the same frame agree to fp32 summation order (~1e-6 relative, RGBA8 within 1), not bit for bit;
     variants 1 and 3 sum a pixel's samples in index order and are bit-reproducible.                                      */
  int32_t  variant;        /* 0 = auto (wavefront for frames of >= 4 Mi paths over trees of >= 256 nodes, else megakernel),
                              1 = flat megakernel, 2 = wavefront, 3 = phased megakernel         */
  int32_t  sampler;        /* where in the pixel a sample lies: 0 = random jitter (main.cpp:210-211, the live code),
                              1 = the 4x rotated grid main.cpp:205-207,214-215 keeps commented out: sample s sits at
                              R_z(15 deg + 90 deg * s) * (0, 0.8) from the pixel's corner (four positions, evaluated
                              in the reference's fp32 arithmetic); lens, shutter time and scatter draws are unchanged */
} srt_render_params;

typedef struct srt_stats {
  uint64_t paths;
  uint64_t segments;       /* world.hit calls, main.cpp:39                                       */
  uint64_t node_visits;    /* aabb::hit calls (only when counters are enabled)                   */
  uint64_t leaf_tests;     /* triangle::hit + sphere::hit calls (only when counters are enabled) */
  float    kernel_ms;      /* device time of the render kernel(s), CUDA events on the launch stream */
  int32_t  launches;       /* kernels launched by the call                                       */
  /* per-kernel device time (CUDA events on the launch stream around each launch):
     traversal = k_wf_extend launches (wavefront) or the whole megakernel; shade = k_wf_shade      */
  float    extend_ms;
  float    shade_ms;
  int32_t  extend_launches;
  int32_t  variant;        /* the variant that ran: 1 flat megakernel, 2 wavefront, 3 phased megakernel */
  int32_t  extend_kernel;  /* variant 2: 0 = k_wf_extend (exact ordered walk), 16 / 32 = k_wf_extend2 with 16- / 32-byte nodes */
  int32_t  fused;          /* variant 2: 1 = the frame ran as ONE launch of k_wf_path (extend + shade in one persistent kernel) */
} srt_stats;

typedef struct srt_device srt_device;
typedef struct srt_scene  srt_scene;

/* ---- life cycle: replaces glDevice::init / terminate (gl.h:28,31) ----------------------------- */
srt_status  srt_device_create(int gpu_id, srt_device** out);
void        srt_device_destroy(srt_device* dev);
/* copies the flattened scene to the GPU (glDevice::init's SSBO upload, gl.h:252-256)            */
srt_status  srt_scene_create(srt_device* dev, const srt_scene_desc* desc, srt_scene** out);
void        srt_scene_destroy(srt_scene* scene);
int64_t     srt_scene_device_bytes(const srt_scene* scene);

/* ---- parity gate: hittableList::hit on explicit rays (hittablelist.h:33-47 -> bvh.h:97-105) -- */
srt_status  srt_trace_rays(srt_scene* scene, const srt_ray* rays, int64_t n, float t_min, float t_max,
                           srt_hit* out_hits);

/* the same rays through the PRODUCTION extend kernel of the wavefront renderer (variant 2) in the launch shape, tree layout and
 * ray ordering it uses for this scene - one wave over n pool slots; returns the primitive id (-1 = miss) and t bits per ray  */
srt_status  srt_trace_rays_wavefront(srt_scene* scene, const srt_ray* rays, int64_t n, int32_t* out_prim, float* out_t);

/* ---- the same wave for callers that may not allocate -------------------------------------------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc (code
 * under CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same kernels, launch shape and results as
 * srt_trace_rays_wavefront, but:
 *   - rays_dev, out_prim_dev and out_t_dev are DEVICE pointers on the scene's GPU (n srt_ray in, n int32 + n float out); the call
 *     makes no host copy and calls no cudaMalloc / cudaFree / allocating library routine: its scratch (ray slots, hit records,
 *     cursors) is a static device array of SRT_STATIC_MAX_RAYS * 40 + 256 bytes (2.5 MiB) that every GPU which loads the library
 *     holds until the process exits;
 *   - n > SRT_STATIC_MAX_RAYS returns SRT_ERR_UNSUPPORTED (split the batch, or use srt_trace_rays_wavefront);
 *   - every call on a GPU shares that array (it keeps the last call's contents; what must start at zero is cleared by the call):
 *     calls that overlap on one GPU race - one call at a time per GPU;
 *   - rays are never re-ordered before the wave (trees beyond the L2 sort them in srt_trace_rays_wavefront); results do not
 *     depend on the order.
 * The call returns after the scene's device stream has been synchronised.                                                      */
#define SRT_STATIC_MAX_RAYS 65536
int64_t     srt_static_max_rays(void);
srt_status  srt_trace_rays_wavefront_static(srt_scene* scene, const srt_ray* rays_dev, int64_t n, int32_t* out_prim_dev,
                                            float* out_t_dev);

/* ---- one frame: replaces glDevice::rtFrame + the loop main.cpp:200-227 ------------------------
 * out_sum_rgb  (optional) [H*W*3] float: sum over this call's samples of rayColor, row 0 = top
 * out_rgba8    (optional) [H*W*4] writeColorTarget output (color.h:25-41), only meaningful when this
 *              call rendered all spp samples                                                    */
srt_status  srt_render(srt_scene* scene, const srt_camera* cam, const srt_render_params* params,
                       float* out_sum_rgb, uint8_t* out_rgba8, srt_stats* stats);

/* device-resident variants for the multi-GPU path (one process per GPU, NCCL reduce outside):
 * accumulate ADDS this call's samples into accum_dev [H*W*3] float (caller zeroes it);
 * resolve applies color.h:25-41 to a summed buffer.                                            */
srt_status  srt_render_accumulate(srt_scene* scene, const srt_camera* cam, const srt_render_params* params,
                                  float* accum_dev, srt_stats* stats);
srt_status  srt_resolve(srt_device* dev, const float* sum_dev, int32_t width, int32_t height, int32_t spp,
                        uint8_t* rgba8_dev);

/* ---- several GPUs behind one handle: glDevice::init / rtFrame (gl.h:28-31, main.cpp:190-199) for a whole box ---------------
 * A device group owns one srt_device per GPU.  srt_scene_group_create uploads a replica of the flattened scene to each of them
 * (in parallel); srt_render_group splits the samples of every pixel across the GPUs (the remainder to the first ones - the
 * Philox stream is keyed by the global sample index, so the frame does not depend on the split, up to fp32 summation order),
 * renders on all of them at once (one host thread per GPU), sums the per-GPU fp32 radiance buffers on the first GPU (peer
 * copies over NVLink + one add per peer: the frame's only exchange) and resolves there.  Outputs and statistics are those of
 * srt_render; kernel_ms is the slowest GPU's.  gpu_ids = NULL with n_gpus = 0 takes every visible GPU; an id may repeat (two
 * devices sharing one GPU - of use for tests on a single-GPU box).  Multi-process callers (one process per GPU, NCCL) use
 * srt_render_accumulate + their own reduce + srt_resolve instead.                                                        */
typedef struct srt_device_group srt_device_group;
typedef struct srt_scene_group  srt_scene_group;
srt_status  srt_device_group_create(const int* gpu_ids, int n_gpus, srt_device_group** out);
void        srt_device_group_destroy(srt_device_group* group);
int32_t     srt_device_group_size(const srt_device_group* group);
srt_device* srt_device_group_device(srt_device_group* group, int32_t k);      /* borrowed: counters, streams */
srt_status  srt_scene_group_create(srt_device_group* group, const srt_scene_desc* desc, srt_scene_group** out);
void        srt_scene_group_destroy(srt_scene_group* scenes);
srt_scene*  srt_scene_group_scene(srt_scene_group* scenes, int32_t k);         /* borrowed */
srt_status  srt_sample_split(int32_t begin, int32_t count, int32_t n, int32_t k, int32_t* out_begin, int32_t* out_count);
srt_status  srt_render_group(srt_scene_group* scenes, const srt_camera* cam, const srt_render_params* params,
                             float* out_sum_rgb, uint8_t* out_rgba8, srt_stats* stats);

/* ---- progressive accumulation: glDevice::rtFrame as a per-frame loop (main.cpp:195-199, gl.h:269-320), headless ----
 * A film is a persistent fp32 sum buffer on the device + the number of samples per pixel it holds.  add_samples
 * renders the next `count` GLOBAL sample indices of every pixel into it (params->spp / sample_begin / sample_count are
 * ignored): the Philox stream is keyed by the global index, so K calls of n samples give the frame of one call of
 * K*n samples up to fp32 summation order.  resolve applies writeColorTarget (color.h:25-41) with the samples so far;
 * load restores a saved sum buffer (checkpoint / resume, see srt_checkpoint_* in srt_host.h).                     */
typedef struct srt_film srt_film;
srt_status  srt_film_create(srt_device* dev, int32_t width, int32_t height, srt_film** out);
void        srt_film_destroy(srt_film* film);
srt_status  srt_film_add_samples(srt_film* film, srt_scene* scene, const srt_camera* cam, const srt_render_params* params,
                                 int32_t count, srt_stats* stats);
int32_t     srt_film_samples(const srt_film* film);
srt_status  srt_film_resolve(srt_film* film, uint8_t* out_rgba8 /*optional, W*H*4*/, float* out_sum_rgb /*optional, W*H*3*/);
srt_status  srt_film_load(srt_film* film, const float* sum_rgb /*W*H*3*/, int32_t samples_done);

/* ---- per-function known-answer entry points (random numbers injected) ------------------------ */
/* material::scatter + emitted for n (ray, hit) pairs; xi[n*4] are the uniform [0,1) numbers the
 * sampler consumes ("random numbers").  out_scattered[i] = 0/1 return of scatter. */
srt_status  srt_shade_kat(srt_scene* scene, const srt_ray* rays_in, const srt_hit* hits, const float* xi,
                          int64_t n, float* out_attenuation /*n*3*/, float* out_emitted /*n*3*/,
                          srt_ray* out_rays, int32_t* out_scattered);
/* texture::value for n lookups: uvp[n*5] = (u, v, p.x, p.y, p.z)                                */
srt_status  srt_texture_kat(srt_scene* scene, int32_t tex, const float* uvp, int64_t n, float* out_rgb);
/* camera::getRay (camera.h:40-50) with its random inputs supplied: in[n*5] = (s, t, disk.x, disk.y,
 * xi_time), disk = the randomInUnitDisk point, xi_time = the uniform number behind randomFloat(t0,t1) */
srt_status  srt_getray_kat(srt_device* dev, const srt_camera* cam, const float* in, int64_t n, srt_ray* out_rays);
/* the primary ray the renderer generates for explicit (pixel, sample) pairs: Philox blocks 0 and 1,
 * the pixel jitter of main.cpp:210-211, then camera::getRay                                      */
srt_status  srt_primary_rays(srt_device* dev, const srt_camera* cam, const srt_render_params* params,
                             const int32_t* pixel_xy /*n*2*/, const int32_t* sample /*n*/, int64_t n,
                             srt_ray* out_rays);

/* aabb::hit (aabb.h:11-27) through the device's box predicates, one (ray, box, tMin, tMax) row per decision:
 * boxes[n*6] = min3 max3, tminmax[n*2].  predicate 0 = filtered predicate of the exact walks (with its fallback), 1 = 16-bit snapped
 * box filter of the compact tree (with its fallback): out_decision must equal aabb::hit; out_path (optional) = 1 where the
 * reference's own arithmetic ran.  predicate 2 / 3 = the conservative cull of k_wf_extend2 on 32- / 16-byte nodes (tMin is the
 * walk's constant 0.001): 1 = kept, 0 = culled - never 0 where aabb::hit is true; out_path = 1 where the ray takes the exact walk */
srt_status  srt_aabb_kat(srt_device* dev, const srt_ray* rays, const float* boxes, const float* tminmax, int64_t n,
                         int32_t predicate, int32_t* out_decision, int32_t* out_path);

/* ---- misc ----------------------------------------------------------------------------------- */
int         srt_abi_version(void);
const char* srt_status_string(srt_status s);
const char* srt_last_error(void);
/* 1 = count node visits / leaf tests in srt_stats (slower kernels); default 0                   */
srt_status  srt_set_counters(srt_device* dev, int enabled);
/* counters of the last trace/render call: node visits, leaf tests, exact (unfiltered) box tests,
 * sphere leaf tests, segments that ended on a triangle                                          */
srt_status  srt_last_counters(srt_device* dev, uint64_t* out5);
/* the CUDA stream the device launches on (cudaStream_t as void*), for event timing outside      */
void*       srt_device_stream(srt_device* dev);
/* launch on the caller's stream instead (e.g. the framework's current stream, so that its events
 * and NCCL calls order with the render); NULL restores the device's own stream                  */
srt_status  srt_device_set_stream(srt_device* dev, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* SRT_H_ */
