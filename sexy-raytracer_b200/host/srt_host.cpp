// srt_host.cpp — host-side helpers of include/srt_host.h (plain C++17, no CUDA).  Built into libsrt_host.so with
//   g++ -O2 -ffp-contract=off -fno-fast-math   (no FMA contraction: results must equal the reference's, which is built
//   without -march / fast-math, CMakeLists.txt:16).
// Everything here restates reference code paths whose RESULTS the device consumes: bvh.h:32-95, aabb.h:33-43,
// model.h:183-212, sphere.h:47-52,85-94, camera.h:10-38, globals.h:26-43, vec3.h:29-31,54-60.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "../../include/srt_host.h"

namespace srt_host_detail {
thread_local std::string g_err;
srt_status fail(srt_status s, const std::string& m) { g_err = m; return s; }
}  // namespace srt_host_detail
using srt_host_detail::fail;
using srt_host_detail::g_err;

namespace {

struct Box { float mn[3], mx[3]; };

// aabb.h:33-43
inline Box surrounding(const Box& a, const Box& b) {
  Box r;
  for (int k = 0; k < 3; k++) { r.mn[k] = fminf(a.mn[k], b.mn[k]); r.mx[k] = fmaxf(a.mx[k], b.mx[k]); }
  return r;
}

// triangle::boundingBox, model.h:183-212 (time independent)
inline Box tri_box(const srt_scene_desc* d, int64_t i) {
  const float inf = std::numeric_limits<float>::infinity();
  Box b;
  const float* v[3] = {d->tri_v0 + 3 * i, d->tri_v1 + 3 * i, d->tri_v2 + 3 * i};
  for (int k = 0; k < 3; k++) { b.mn[k] = inf; b.mx[k] = -inf; }
  for (int q = 0; q < 3; q++)
    for (int k = 0; k < 3; k++) { b.mn[k] = std::min(b.mn[k], v[q][k]); b.mx[k] = std::max(b.mx[k], v[q][k]); }
  for (int k = 0; k < 3; k++)
    if (b.mn[k] == b.mx[k]) { b.mn[k] -= 0.0001f; b.mx[k] += 0.0001f; }
  return surrounding(b, b);
}

// sphere::center, sphere.h:47-52
inline void sph_center(const srt_scene_desc* d, int64_t j, float time, float c[3]) {
  const float *c0 = d->sph_c0 + 3 * j, *c1 = d->sph_c1 + 3 * j;
  if (c0[0] != c1[0] || c0[1] != c1[1] || c0[2] != c1[2]) {
    const float f = (time - d->sph_t0[j]) / (d->sph_t1[j] - d->sph_t0[j]);
    for (int k = 0; k < 3; k++) c[k] = c0[k] + f * (c1[k] - c0[k]);
  } else {
    for (int k = 0; k < 3; k++) c[k] = c0[k];
  }
}

// sphere::boundingBox, sphere.h:85-94
inline Box sph_box(const srt_scene_desc* d, int64_t j, float time0, float time1) {
  float a[3], b[3];
  const float r = d->sph_r[j];
  sph_center(d, j, time0, a);
  sph_center(d, j, time1, b);
  Box b0, b1;
  for (int k = 0; k < 3; k++) { b0.mn[k] = a[k] - r; b0.mx[k] = a[k] + r; b1.mn[k] = b[k] - r; b1.mx[k] = b[k] + r; }
  return surrounding(b0, b1);
}

inline Box prim_box(const srt_scene_desc* d, int64_t p, float t0, float t1) {
  return p < d->n_tri ? tri_box(d, p) : sph_box(d, p - d->n_tri, t0, t1);
}

// number of bvhNode objects the constructor creates for a span of s primitives (bvh.h:66-86)
int64_t node_count(int64_t s) {
  if (s <= 2) return s > 0 ? 1 : 0;
  return 1 + node_count(s / 2) + node_count(s - s / 2);
}

struct Builder {
  const srt_scene_desc* d;
  std::vector<float> key[3];          // box minimum at time 0 per axis: the comparator's key (bvh.h:32-40 calls boundingBox(0, 0))
  std::vector<Box> pbox;              // primitive boxes over [time0, time1]
  std::vector<int32_t> ids;           // the object vector, sorted in place span by span
  std::vector<uint8_t> axis;          // per node, in pre-order: randomInt(0, 2) (bvh.h:60)
  srt_bvh_node* nodes;
  int max_depth;

  struct KeyId { float k; int32_t id; };

  Box build(int64_t start, int64_t end, int64_t idx, int depth) {
    const int ax = axis[idx];
    const float* K = key[ax].data();
    const int64_t span = end - start;
    srt_bvh_node& nd = nodes[idx];
    Box box;
    if (span == 1) {
      nd.left = ~ids[start]; nd.right = SRT_BVH_NONE;
      box = surrounding(pbox[ids[start]], pbox[ids[start]]);
    } else if (span == 2) {
      int32_t a = ids[start], b = ids[start + 1];
      if (!(K[a] < K[b])) std::swap(a, b);
      nd.left = ~a; nd.right = ~b;
      box = surrounding(pbox[a], pbox[b]);
    } else {
      // std::sort with the reference's comparator.  Sorting (key, id) records performs the same comparisons and moves
      // as sorting the ids through an indirect comparator, hence yields the same permutation - without the cache misses.
      if (span >= 64) {
        std::vector<KeyId> tmp((size_t)span);
        for (int64_t q = 0; q < span; q++) tmp[q] = KeyId{K[ids[start + q]], ids[start + q]};
        std::sort(tmp.begin(), tmp.end(), [](const KeyId& x, const KeyId& y) { return x.k < y.k; });
        for (int64_t q = 0; q < span; q++) ids[start + q] = tmp[q].id;
      } else {
        std::sort(ids.begin() + start, ids.begin() + end, [K](int32_t x, int32_t y) { return K[x] < K[y]; });
      }
      const int64_t mid = start + span / 2;
      const int64_t li = idx + 1, ri = idx + 1 + node_count(mid - start);
      Box bl, br;
      if (depth < max_depth && span >= 8192) {
        std::thread t([&] { bl = build(start, mid, li, depth + 1); });
        br = build(mid, end, ri, depth + 1);
        t.join();
      } else {
        bl = build(start, mid, li, depth + 1);
        br = build(mid, end, ri, depth + 1);
      }
      nd.left = (int32_t)li; nd.right = (int32_t)ri;
      box = surrounding(bl, br);
    }
    for (int k = 0; k < 3; k++) { nd.bmin[k] = box.mn[k]; nd.bmax[k] = box.mx[k]; }
    return box;
  }
};

inline float length3(const float v[3]) { return sqrtf((v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]); }   // vec3.h:29-35
inline void unit3(const float v[3], float o[3]) {                                                     // vec3.h:54-60
  const float len = length3(v);
  if (len != 0) { o[0] = v[0] / len; o[1] = v[1] / len; o[2] = v[2] / len; }
  else { o[0] = v[0]; o[1] = v[1]; o[2] = v[2]; }
}
inline void cross3(const float a[3], const float b[3], float o[3]) {                                  // Eigen OrthoMethods.h:47-49
  o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}

}  // namespace

extern "C" {

const char* srt_host_last_error(void) { return g_err.c_str(); }

srt_status srt_prim_boxes(const srt_scene_desc* d, float time0, float time1, float* out) try {
  if (!d || !out) return fail(SRT_ERR_INVALID, "NULL argument");
  const int64_t n = d->n_tri + d->n_sph;
  for (int64_t p = 0; p < n; p++) {
    const Box b = prim_box(d, p, time0, time1);
    for (int k = 0; k < 3; k++) { out[6 * p + k] = b.mn[k]; out[6 * p + 3 + k] = b.mx[k]; }
  }
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

void srt_bvh_free(srt_bvh_node* nodes) { free(nodes); }

srt_status srt_bvh_build(const srt_scene_desc* d, float time0, float time1, uint64_t rng_skip, int threads,
                         srt_bvh_node** out_nodes, int64_t* out_n, uint64_t* out_draws) try {
  return srt_bvh_build_ordered(d, nullptr, time0, time1, rng_skip, threads, out_nodes, out_n, out_draws);
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_bvh_build_ordered(const srt_scene_desc* d, const int32_t* order, float time0, float time1, uint64_t rng_skip,
                                 int threads, srt_bvh_node** out_nodes, int64_t* out_n, uint64_t* out_draws) try {
  if (!d || !out_nodes || !out_n) return fail(SRT_ERR_INVALID, "NULL argument");
  *out_nodes = nullptr; *out_n = 0;
  if (d->n_tri < 0 || d->n_sph < 0) return fail(SRT_ERR_INVALID, "negative count");
  const int64_t n = d->n_tri + d->n_sph;
  if (n > (int64_t)INT32_MAX / 2) return fail(SRT_ERR_UNSUPPORTED, "too many primitives");
  if (out_draws) *out_draws = 0;
  if (n == 0) return SRT_OK;
  const int64_t n_nodes = node_count(n);

  Builder B;
  B.d = d;
  for (int k = 0; k < 3; k++) B.key[k].resize((size_t)n);
  B.pbox.resize((size_t)n);
  B.ids.resize((size_t)n);
  unsigned hw = threads > 0 ? (unsigned)threads : std::max(1u, std::thread::hardware_concurrency());
  {  // primitive boxes, in parallel
    std::vector<std::thread> pool;
    for (unsigned w = 0; w < hw; w++)
      pool.emplace_back([&, w] {
        for (int64_t p = n * w / hw; p < n * (w + 1) / hw; p++) {
          B.pbox[p] = prim_box(d, p, time0, time1);
          const Box b0 = prim_box(d, p, 0.0f, 0.0f);             // boxCompare: boundingBox(0, 0, ...), bvh.h:36
          for (int k = 0; k < 3; k++) B.key[k][p] = b0.mn[k];
          B.ids[p] = order ? order[p] : (int32_t)p;
        }
      });
    for (auto& t : pool) t.join();
  }
  if (order) {                                                     // must be a permutation of [0, n)
    std::vector<uint8_t> seen((size_t)n, 0);
    for (int64_t p = 0; p < n; p++) {
      if (order[p] < 0 || order[p] >= n || seen[order[p]]) return fail(SRT_ERR_INVALID, "order is not a permutation of the primitive ids");
      seen[order[p]] = 1;
    }
  }
  // one randomInt(0, 2) per node, drawn in construction (= pre-order) order from the reference's generator:
  // static std::uniform_real_distribution<float>(0, 1) over a default-seeded std::mt19937 (globals.h:30-43)
  B.axis.resize((size_t)n_nodes);
  {
    std::mt19937 gen;
    gen.discard(rng_skip);
    std::uniform_real_distribution<float> dist(0.0f, 1.0f);
    for (int64_t k = 0; k < n_nodes; k++) {
      const float mn = 0, mx = 2 + 1;
      B.axis[k] = (uint8_t)static_cast<int>(mn + (mx - mn) * dist(gen));
      if (B.axis[k] > 2) B.axis[k] = 2;     // cannot happen with libstdc++'s generate_canonical fix-up; guards the index
    }
  }
  B.nodes = (srt_bvh_node*)malloc((size_t)n_nodes * sizeof(srt_bvh_node));
  if (!B.nodes) return fail(SRT_ERR_NOMEM, "node array");
  B.max_depth = 0;
  while ((1u << B.max_depth) < hw) ++B.max_depth;
  B.build(0, n, 0, 0);
  *out_nodes = B.nodes; *out_n = n_nodes;
  if (out_draws) *out_draws = (uint64_t)n_nodes;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_camera_make(const float eye[3], const float look_at[3], const float up[3], float vfov, float aspect,
                           float aperture, float focus_dist, float time0, float time1, srt_camera* c) try {
  if (!eye || !look_at || !up || !c) return fail(SRT_ERR_INVALID, "NULL argument");
  const float pi = 3.1415926535897932385f;                      // globals.h:15
  const float theta = vfov * pi / 180.0f;                       // deg2rad, globals.h:26-28
  // `auto h = tan(theta / 2.0f)` (camera.h:20) resolves to ::tan(double): h, vpHeight and vpWidth are doubles, and
  // `focusDist * vpWidth * hor` rounds the double product to float once (Eigen converts the scalar factor to float)
  const double h = tan((double)(theta / 2.0f));
  const double vp_h = 2.0f * h, vp_w = aspect * vp_h;
  float d[3] = {eye[0] - look_at[0], eye[1] - look_at[1], eye[2] - look_at[2]}, t[3];
  unit3(d, c->w);
  cross3(up, c->w, t); unit3(t, c->hor);
  cross3(c->w, c->hor, t); unit3(t, c->vert);
  const float sw = (float)(focus_dist * vp_w), sh = (float)(focus_dist * vp_h);
  for (int k = 0; k < 3; k++) {
    c->origin[k] = eye[k];
    c->horizontal[k] = sw * c->hor[k];
    c->vertical[k] = sh * c->vert[k];
    c->lleft[k] = ((c->origin[k] - c->horizontal[k] / 2.0f) - c->vertical[k] / 2.0f) - focus_dist * c->w[k];
  }
  c->lens_radius = aperture / 2.0f;
  c->t0 = time0; c->t1 = time1;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

}  // extern "C"
