/* srt_host.h — C ABI of the HOST side helpers that sit immediately before the path ("libsrt_host.so", plain C++,
 * no CUDA): the parts of scene construction whose results the device consumes and that must equal the
 * reference's bit for bit.
 *
 *   srt_bvh_build    : bvhNode::bvhNode(objects, 0, n, time0, time1)   (reference bvh.h:55-95), in place
 *   srt_camera_make  : camera::camera(...)                              (reference camera.h:10-38)
 *   srt_prim_boxes   : triangle::boundingBox / sphere::boundingBox      (reference model.h:183-212, sphere.h:85-94)
 *
 * Why a builder of our own: the reference constructor copies the whole object vector at EVERY node (bvh.h:57), i.e.
 * O(N^2) - 92 ms for 3 046 primitives, unusable for BASELINE config 5 (10.1 M primitives); and triangle holds uint16
 * indices (model.h:46).  This builder sorts each node's span in place (children only ever look at their own slice, so
 * the result is identical), draws the split axes from the same std::mt19937 stream in the same pre-order, uses the same
 * comparator (box minimum at time 0, bvh.h:32-40) with the same std::sort, and builds sub-trees on several threads
 * (the tree SHAPE depends on n alone, so every node's pre-order index - hence its random draw - is known up front).
 */
#ifndef SRT_HOST_H_
#define SRT_HOST_H_

#include "srt.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Boxes of the primitives of `prims` (nodes / tables ignored): out[i*6..] = min.xyz, max.xyz over [time0, time1],
 * primitive order = triangles then spheres.                                                                  */
srt_status srt_prim_boxes(const srt_scene_desc* prims, float time0, float time1, float* out_boxes);

/* Builds the tree over all primitives of `prims`.
 *  rng_skip   : outputs of the default-seeded std::mt19937 (seed 5489) already consumed by randomFloat() before the
 *               bvhNode constructor runs (0 for main.cpp's default scene; each randomFloat() = one output)
 *  threads    : worker threads (<= 0: hardware concurrency)
 *  out_nodes  : malloc'ed array in srt_bvh_node form (pre-order, root 0), free with srt_bvh_free
 *  out_draws  : optional, number of random outputs the build consumed (= number of nodes)                      */
srt_status srt_bvh_build(const srt_scene_desc* prims, float time0, float time1, uint64_t rng_skip, int threads,
                         srt_bvh_node** out_nodes, int64_t* out_n_nodes, uint64_t* out_draws);
void       srt_bvh_free(srt_bvh_node* nodes);
/* As srt_bvh_build, for scenes whose objects were NOT inserted triangles-first: order[i] = primitive id (in the
 * triangles-then-spheres numbering of srt.h) of the i-th object of the reference's `objects` vector.  std::sort is not
 * stable, so the initial order of the vector is part of the contract.  order == NULL means identity.            */
srt_status srt_bvh_build_ordered(const srt_scene_desc* prims, const int32_t* order, float time0, float time1, uint64_t rng_skip,
                                 int threads, srt_bvh_node** out_nodes, int64_t* out_n_nodes, uint64_t* out_draws);

/* camera.h:10-38: the ten derived members from (eye, lookAt, up, vFOV degrees, aspect, aperture, focusDist, t0, t1) */
srt_status srt_camera_make(const float eye[3], const float look_at[3], const float up[3], float vfov, float aspect,
                           float aperture, float focus_dist, float time0, float time1, srt_camera* out);

/* ---- images: stbi_load as imagePNG uses it (texture.h:112-123) and stbi_write_png (main.cpp:235-237) ---------------
 * read : 8-bit samples converted to req_comp (1..4) channels by stb_image's rules, row 0 = top; *out_pixels is malloc'ed
 *        with 4 spare zero bytes at the end, free with srt_host_free.
 * write: 8-bit, comp = 1..4 channels, stride_bytes between rows (0 = w*comp).                                       */
srt_status srt_png_read(const char* path, int32_t req_comp, int32_t* out_w, int32_t* out_h, uint8_t** out_pixels);
srt_status srt_png_read_memory(const uint8_t* file_bytes, int64_t len, int32_t req_comp, int32_t* out_w, int32_t* out_h,
                               uint8_t** out_pixels);
srt_status srt_png_write(const char* path, int32_t w, int32_t h, int32_t comp, const uint8_t* pixels, int32_t stride_bytes);
void       srt_host_free(void* p);

/* ---- raw radiance dump for parity tooling: Portable Float Map, colour ("PF"), little endian.
 * `rgb` = width*height*3 floats, row 0 = TOP (the layout of srt_render's out_sum_rgb); `scale` multiplies every value on the
 * way out (1 / spp turns the sums into mean radiance).  PFM stores rows bottom-up: the writer flips, the reader flips back. */
srt_status srt_pfm_write(const char* path, int32_t width, int32_t height, const float* rgb, float scale);
srt_status srt_pfm_read(const char* path, int32_t* out_width, int32_t* out_height, float** out_rgb);   /* free with srt_host_free */

/* ---- checkpoint of a progressive render: the fp32 sum buffer + how many samples per pixel it holds.
 * File: "SRTF" u32 version (1; write2: 2), i32 width, height, samples_done, u64 seed, then width*height*3 little-endian floats.   */
srt_status srt_checkpoint_write(const char* path, int32_t width, int32_t height, int32_t samples_done, uint64_t seed,
                                const float* sum_rgb);
/* *out_sum_rgb is malloc'ed (free with srt_host_free) */
srt_status srt_checkpoint_read(const char* path, int32_t* width, int32_t* height, int32_t* samples_done, uint64_t* seed,
                               float** out_sum_rgb);
/* version 2 of the file: i32 sampler, i32 max_bounce follow the seed, so that a resume with other parameters (which would mix
 * two sample streams under continuing global sample indices) can be refused.  read2 also reads version 1 (-1 = not recorded);
 * out_sampler / out_max_bounce may be NULL.                                                                            */
srt_status srt_checkpoint_write2(const char* path, int32_t width, int32_t height, int32_t samples_done, uint64_t seed,
                                 int32_t sampler, int32_t max_bounce, const float* sum_rgb);
srt_status srt_checkpoint_read2(const char* path, int32_t* width, int32_t* height, int32_t* samples_done, uint64_t* seed,
                                int32_t* out_sampler, int32_t* out_max_bounce, float** out_sum_rgb);

/* ---- whole scenes through the C++ host classes of srt_world.hpp (for callers that are not C++: tests, bench) ----------
 * which: 1 = the Shirley sphere block of main.cpp:92-122 re-enabled (+ main.cpp:124-131,143-144)   [BASELINE config 1]
 *        2 = cube.gltf + square.gltf + ground + light                                              [BASELINE config 2]
 *        3 = randomScene()'s default branch, main.cpp:72-76,82-90,126-146                          [BASELINE configs 3, 4]
 *        4 = as 3, the globe's four maps pointed at files that exist in this checkout
 * data_dir holds the reference's assets (its data/ directory).  The process-wide generator is reset first, as at the
 * start of main().  The scene owns everything srt_host_scene_desc points at.                                      */
typedef struct srt_host_scene srt_host_scene;
srt_status srt_host_scene_reference(int which, const char* data_dir, srt_host_scene** out);
/* options: bit 0 = gltfOptions::referenceQuirks (1 = load exactly like model.h:301-460), bit 1 = smoothNormals,
 * bit 2 = sampleMetallicRoughness, bit 3 = vertexTangents (include/srt_world.hpp; the last three are NOT reference behaviour) */
#define SRT_GLTF_REFERENCE_QUIRKS 1
#define SRT_GLTF_SMOOTH_NORMALS 2
#define SRT_GLTF_SAMPLE_MR 4
#define SRT_GLTF_VERTEX_TANGENTS 8
srt_status srt_host_scene_gltf(const char* gltf_path, const char* texture_dir, int options, srt_host_scene** out);
const srt_scene_desc* srt_host_scene_desc(const srt_host_scene* s);
uint64_t   srt_host_scene_draws(const srt_host_scene* s);      /* generator outputs consumed by the set-up */
void       srt_host_scene_free(srt_host_scene* s);

const char* srt_host_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* SRT_HOST_H_ */
