// srt_kernels.cuh — the sm_100a kernels.
//   k_setup_tris      : N and dd of every triangle, with the reference's arithmetic (model.h:276-283,:125)
//   k_trace_rays      : hittableList::hit for explicit rays -> full hit records (parity gate)
//   k_render_mega     : persistent megakernel: ray generation + ordered traversal + shading + accumulation
//   k_resolve         : writeColorTarget (color.h:25-41)
//   k_shade_kat / k_texture_kat / k_getray_kat : per-function known-answer entry points
#pragma once
#include "srt_shade.cuh"

namespace srt {

__global__ void k_setup_tris(Tri* tris, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float4 A = tris[i].a, B = tris[i].b, C = tris[i].c;
  V3 v0 = mk(A.x, A.y, A.z), v1 = mk(B.x, B.y, B.z), v2 = mk(C.x, C.y, C.z);
  V3 N = cross(v1 - v0, v2 - v0);          // getNormal, model.h:276-283
  float dd = -dot(N, v0);                  // model.h:125
  tris[i].a.w = N.x; tris[i].b.w = N.y; tris[i].c.w = N.z; tris[i].d.x = dd;
}

// image texels: texel (i, j) = (px[0], px[1], px[2], 255) for px = data + j*bpp*w + i*bpp, bytes past the end of the image
// read 0 (a 1- or 2-byte-per-texel map read with the 3-byte lookup of texture.h:129-148 sees its neighbours)
__global__ void k_expand_texels(const unsigned char* __restrict__ data, size_t n, int w, int h, int bpp, uchar4* __restrict__ out) {
  const size_t k = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (k >= (size_t)w * (size_t)h) return;
  const size_t j = k / (size_t)w, i = k % (size_t)w;
  const size_t off = j * ((size_t)bpp * (size_t)w) + i * (size_t)bpp;
  uchar4 p;
  p.x = off < n ? data[off] : 0; p.y = off + 1 < n ? data[off + 1] : 0; p.z = off + 2 < n ? data[off + 2] : 0; p.w = 255;
  out[k] = p;
}

// leaf-ordered copies of the primitives for the compact tree (srt_core.cuh: LPrim); runs after k_setup_tris
__global__ void k_build_lprims(Scene sc, const int32_t* __restrict__ src, int n, LPrim* __restrict__ out) {
  int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= n) return;
  const int p = src[u];
  LPrim r;
  if (p < sc.n_tri) {
    const Tri t = sc.tris[p];
    r.a = t.a; r.b = t.b; r.c = t.c; r.d = make_float4(t.d.x, t.d.y, __int_as_float(p), __int_as_float(0));
  } else {
    const Sph q = sc.sphs[p - sc.n_tri];
    r.a = q.a; r.b = q.b; r.c = q.c; r.d = make_float4(0.0f, q.c.y, __int_as_float(p), __int_as_float(1));
  }
  out[u] = r;
}

__device__ __forceinline__ Ray load_ray(const srt_ray* q) {
  Ray r;
  r.o = mk(q->o[0], q->o[1], q->o[2]); r.d = mk(q->dir[0], q->dir[1], q->dir[2]); r.time = q->time;
  return r;
}
__device__ __forceinline__ void store_ray(srt_ray* q, const Ray& r) {
  q->o[0] = r.o.x; q->o[1] = r.o.y; q->o[2] = r.o.z; q->dir[0] = r.d.x; q->dir[1] = r.d.y; q->dir[2] = r.d.z; q->time = r.time;
}

// cooperative copy of the node array into shared memory (whole tree resident next to the SM)
__device__ __forceinline__ const Node* stage_nodes(const Scene& sc, bool in_smem, float4* smem) {
  if (!in_smem) return sc.nodes;
  const float4* src = reinterpret_cast<const float4*>(sc.nodes);
  for (int k = threadIdx.x; k < (sc.n_nodes + 1) * 2; k += blockDim.x) smem[k] = __ldg(&src[k]);   // + the sentinel
  __syncthreads();
  return reinterpret_cast<const Node*>(smem);
}

template <bool COUNT>
__global__ void __launch_bounds__(256) k_trace_rays(Scene sc, int nodes_in_smem, const srt_ray* __restrict__ rays,
                                                    long long n, float tMin, float tMax, srt_hit* __restrict__ out,
                                                    unsigned long long* counters) {
  extern __shared__ float4 smem_nodes[];
  const Node* nodes = stage_nodes(sc, nodes_in_smem != 0, smem_nodes);
  Counters cnt = {0, 0, 0, 0, 0};
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    Ray r = load_ray(&rays[i]);
    int prim; float t;
    if (sc.cnodes && !nodes_in_smem) traverse_compact<COUNT>(sc, r, tMin, tMax, prim, t, &cnt);   // big trees: the compact walk
    else traverse<COUNT>(sc, nodes, r, tMin, tMax, prim, t, &cnt);
    srt_hit h;
    if (prim < 0) {
      h.prim = -1; h.mat = -1; h.front_face = 0; h.t = 0;
      for (int k = 0; k < 3; k++) h.p[k] = h.normal[k] = h.tangent[k] = h.bitangent[k] = 0;
      h.uv[0] = h.uv[1] = 0;
    } else {
      Rec rec;
      finish_hit<false>(sc, prim, r, t, true, true, rec);
      h.prim = rec.prim; h.mat = rec.mat; h.front_face = rec.front_face; h.t = rec.t;
      h.p[0] = rec.p.x; h.p[1] = rec.p.y; h.p[2] = rec.p.z;
      h.normal[0] = rec.normal.x; h.normal[1] = rec.normal.y; h.normal[2] = rec.normal.z;
      h.tangent[0] = rec.tangent.x; h.tangent[1] = rec.tangent.y; h.tangent[2] = rec.tangent.z;
      h.bitangent[0] = rec.bitangent.x; h.bitangent[1] = rec.bitangent.y; h.bitangent[2] = rec.bitangent.z;
      h.uv[0] = rec.u; h.uv[1] = rec.v;
    }
    out[i] = h;
  }
  if (COUNT) {
    atomicAdd(&counters[2], cnt.nodes); atomicAdd(&counters[3], cnt.leaves); atomicAdd(&counters[4], cnt.exact);
    atomicAdd(&counters[6], cnt.sph_leaves);
  }
}

// ---- the frame ----------------------------------------------------------------------------------
struct RenderArgs {
  srt_camera cam;
  int W, H;
  int sample_begin, sample_count;   // global sample indices [begin, begin+count) of every pixel
  int max_bounce, rr_begin;
  float bg[3];
  unsigned long long seed;
  int tiles_x, n_tiles;             // 8x4-pixel tiles
  int chunk, n_chunks;              // samples per work unit
  long long n_units;
  int sampler;                      // srt_render_params::sampler
  const float2* grid; int grid_begin;   // sampler 1: in-pixel position of global sample s at grid[s - grid_begin]
};

// main.cpp:181,205-207: samplePos = (0, 0.8, 0) turned about z by deg2rad(15) + deg2rad(90 s) - Eigen's AngleAxisf matrix
// has m01 = -sin, m11 = cos for the z axis, and the products with samplePos's zeros are exact.  Evaluated on the host
// (glibc sinf / cosf, as the reference would) for every sample of the call: the angle is not reduced mod 360, so the
// four positions drift by a few ulp as s grows, and the table keeps that.
inline void srt_rotated_grid(int s, float* x, float* y) {
  const float pi = 3.1415926535897932385f;                                   // globals.h:15
  const float ang = 15.0f * pi / 180.0f + (90.0f * (float)s) * pi / 180.0f;  // globals.h:26-28
  *x = -sinf(ang) * 0.8f;
  *y = cosf(ang) * 0.8f;
}

// One sample of one pixel: the two ray-generation Philox blocks, main.cpp:210-211, camera.h:40-50
template <bool SMALL = false>
__device__ __forceinline__ Ray generate_ray(const RenderArgs& A, int x, int y, uint32_t pixel, uint32_t sample) {
  uint4 a = philox_sel<SMALL>(A.seed, pixel, sample, 0), c = philox_sel<SMALL>(A.seed, pixel, sample, 1);
  float ju = u01(a.x), jv = u01(a.y);
  if (A.sampler == 1) { const float2 g = __ldg(A.grid + ((int)sample - A.grid_begin)); ju = g.x; jv = g.y; }   // main.cpp:214-215
  float u = ((float)x + ju) / (float)(A.W - 1);
  float v = ((float)(A.H - y) + jv) / (float)(A.H - 1);
  float dx, dy;
  sample_disk(u01(a.z), u01(a.w), dx, dy);
  return camera_get_ray(A.cam, u, v, dx, dy, u01(c.x));
}

// One finished segment of a path: main.cpp:39-51 after world.hit returned (prim, t).  Updates the path's
// radiance L and throughput thr (rayColor's recursion L = e0 + a0*(e1 + a1*(...)) unrolled), replaces r by
// the scattered ray and advances bounce.  Returns true when the path ends here.
template <bool COUNT, bool SMALL = false>
__device__ __forceinline__ bool shade_segment(const Scene& sc, const RenderArgs& A, uint32_t pixel, uint32_t sample, int& bounce,
                                              Ray& r, int prim, float t, V3& L, V3& thr, Counters& cnt) {
  if (prim < 0) {                                                         // miss -> background, main.cpp:39-40
    L = mk(L.x + thr.x * A.bg[0], L.y + thr.y * A.bg[1], L.z + thr.z * A.bg[2]);
    return true;
  }
  if (COUNT && prim < sc.n_tri) ++cnt.tri_hits;
  const int mat_id = prim < sc.n_tri ? __float_as_int(__ldg(&sc.tris[prim].d.y))
                                     : __float_as_int(__ldg(&sc.sphs[prim - sc.n_tri].c.y));
  const Mat m = sc.mats[mat_id];
  Rec rec;
  finish_hit<true>(sc, prim, r, t, (m.flags & MAT_NEEDS_UV) != 0, (m.flags & MAT_NEEDS_TBN) != 0, rec);
  const uint4 x4 = philox_sel<SMALL>(A.seed, pixel, sample, 2u + (uint32_t)bounce);
  V3 rvec = mk(0, 0, 0);
  if (m.kind == SRT_MAT_PBR) rvec = sample_unit_vector(u01(x4.x), u01(x4.y));
  else if (m.kind == SRT_MAT_METAL) rvec = sample_unit_ball(u01(x4.x), u01(x4.y), u01(x4.z));
  V3 e = material_emitted(sc, m, rec);
  L = mk(L.x + thr.x * e.x, L.y + thr.y * e.y, L.z + thr.z * e.z);
  V3 att; Ray next;
  if (!material_scatter<true>(sc, m, r, rec, rvec, u01(x4.z), att, next)) return true;
  thr = mk(thr.x * att.x, thr.y * att.y, thr.z * att.z);
  if (A.rr_begin > 0 && bounce + 1 >= A.rr_begin) {     // Russian roulette: not in the reference, off by default
    float q = fminf(fmaxf(fmaxf(thr.x, fmaxf(thr.y, thr.z)), 0.05f), 1.0f);
    if (!(u01(x4.w) < q)) return true;
    float inv = 1.0f / q; thr = mk(thr.x * inv, thr.y * inv, thr.z * inv);
  }
  r = next;
  return ++bounce >= A.max_bounce;                       // maxBounce <= 0 -> (0,0,0), main.cpp:36-37
}

// Persistent megakernel.  A warp owns one work unit at a time = (8x4 pixel tile, chunk of samples);
// lane l owns pixel l of the tile.  The body is ONE flat loop over path segments: a lane whose path
// ended regenerates the next sample of its pixel in the same iteration, so the warp stays full until
// the unit runs dry (no queues, no atomics on the hot path).  rayColor's recursion (main.cpp:33-52)
//   L = e0 + a0*(e1 + a1*(e2 + ...))   becomes   L += thr*e_k ; thr *= a_k.
// counters: [0] paths, [1] segments, [2] node visits, [3] leaf tests, [4] exact box tests, [5] unit cursor,
//           [6] sphere leaf tests, [7] segments that ended on a triangle
template <bool COUNT>
__global__ void __launch_bounds__(512, 1) k_render_mega(Scene sc, int nodes_in_smem, RenderArgs A, float* __restrict__ accum,
                                                        unsigned long long* counters) {
  extern __shared__ float4 smem_nodes[];
  const Node* nodes = stage_nodes(sc, nodes_in_smem != 0, smem_nodes);
  const int lane = threadIdx.x & 31;
  unsigned long long n_paths = 0, n_segments = 0;
  Counters cnt = {0, 0, 0, 0, 0};

  for (;;) {
    long long unit = 0;
    if (lane == 0) unit = (long long)atomicAdd(&counters[5], 1ULL);
    unit = __shfl_sync(0xffffffffu, unit, 0);
    if (unit >= A.n_units) break;
    const int tile = (int)(unit % A.n_tiles), chunk = (int)(unit / A.n_tiles);
    const int x = (tile % A.tiles_x) * 8 + (lane & 7), y = (tile / A.tiles_x) * 4 + (lane >> 3);
    const bool valid = x < A.W && y < A.H;
    const uint32_t pixel = (uint32_t)(y * A.W + x);
    int s = A.sample_begin + chunk * A.chunk;
    const int s_end = valid ? min(s + A.chunk, A.sample_begin + A.sample_count) : s;

    V3 sum = mk(0, 0, 0), L = mk(0, 0, 0), thr = mk(1, 1, 1);
    Ray r; r.o = r.d = mk(0, 0, 0); r.time = 0;
    int bounce = -1;                       // -1: no path in flight
    uint32_t sample = 0;
    for (;;) {
      if (bounce < 0) {
        if (s >= s_end) break;
        sample = (uint32_t)s++;
        r = generate_ray(A, x, y, pixel, sample);
        L = mk(0, 0, 0); thr = mk(1, 1, 1); bounce = 0;
        ++n_paths;
      }
      // one path segment: world.hit (main.cpp:39) + emitted/scatter (main.cpp:44-47)
      ++n_segments;
      int prim; float t;
      traverse<COUNT>(sc, nodes, r, 0.001f, __int_as_float(0x7f800000), prim, t, &cnt);
      const bool done = shade_segment<COUNT>(sc, A, pixel, sample, bounce, r, prim, t, L, thr, cnt);
      if (done) { sum = sum + L; bounce = -1; }
    }
    if (valid) {
      float* px = accum + (size_t)pixel * 3;
      if (A.n_chunks == 1) { px[0] += sum.x; px[1] += sum.y; px[2] += sum.z; }      // sole owner of the pixel
      else { atomicAdd(px, sum.x); atomicAdd(px + 1, sum.y); atomicAdd(px + 2, sum.z); }
    }
  }
  // statistics: one atomic per warp
  for (int o = 16; o > 0; o >>= 1) {
    n_paths += __shfl_down_sync(0xffffffffu, n_paths, o);
    n_segments += __shfl_down_sync(0xffffffffu, n_segments, o);
    if (COUNT) {
      cnt.nodes += __shfl_down_sync(0xffffffffu, cnt.nodes, o);
      cnt.leaves += __shfl_down_sync(0xffffffffu, cnt.leaves, o);
      cnt.exact += __shfl_down_sync(0xffffffffu, cnt.exact, o);
      cnt.sph_leaves += __shfl_down_sync(0xffffffffu, cnt.sph_leaves, o);
      cnt.tri_hits += __shfl_down_sync(0xffffffffu, cnt.tri_hits, o);
    }
  }
  if (lane == 0) {
    atomicAdd(&counters[0], n_paths); atomicAdd(&counters[1], n_segments);
    if (COUNT) {
      atomicAdd(&counters[2], cnt.nodes); atomicAdd(&counters[3], cnt.leaves); atomicAdd(&counters[4], cnt.exact);
      atomicAdd(&counters[6], cnt.sph_leaves); atomicAdd(&counters[7], cnt.tri_hits);
    }
  }
}

// writeColorTarget, color.h:25-41
__global__ void k_resolve(const float* __restrict__ sum, int n_pixels, int spp, uchar4* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_pixels) return;
  float scale = 1.0f / spp;
  float r = sqrtf(sum[3 * i] * scale), g = sqrtf(sum[3 * i + 1] * scale), b = sqrtf(sum[3 * i + 2] * scale);
  uchar4 px;
  // static_cast<uint8_t>(float): x86 converts through int32 and keeps the low byte; NaN -> 0
  px.x = (unsigned char)(int)(256 * clampf(r, 0.0f, 0.999f));
  px.y = (unsigned char)(int)(256 * clampf(g, 0.0f, 0.999f));
  px.z = (unsigned char)(int)(256 * clampf(b, 0.0f, 0.999f));
  px.w = 255;
  out[i] = px;
}

// ---- known-answer kernels -------------------------------------------------------------------------
__global__ void k_shade_kat(Scene sc, const srt_ray* __restrict__ rays, const srt_hit* __restrict__ hits,
                            const float* __restrict__ xi, long long n, float* att, float* emit, srt_ray* out_rays,
                            int* scattered) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  Ray r = load_ray(&rays[i]);
  const srt_hit& h = hits[i];
  Rec rec;
  rec.p = mk(h.p[0], h.p[1], h.p[2]); rec.normal = mk(h.normal[0], h.normal[1], h.normal[2]);
  rec.tangent = mk(h.tangent[0], h.tangent[1], h.tangent[2]); rec.bitangent = mk(h.bitangent[0], h.bitangent[1], h.bitangent[2]);
  rec.u = h.uv[0]; rec.v = h.uv[1]; rec.t = h.t; rec.front_face = h.front_face; rec.mat = h.mat; rec.prim = h.prim;
  const Mat m = sc.mats[rec.mat];
  V3 e = material_emitted(sc, m, rec), a = mk(0, 0, 0);
  Ray o; o.o = o.d = mk(0, 0, 0); o.time = 0;
  bool ok = material_scatter<false>(sc, m, r, rec, mk(xi[4 * i], xi[4 * i + 1], xi[4 * i + 2]), xi[4 * i + 3], a, o);
  if (!ok && m.kind == SRT_MAT_LIGHT) { o.o = o.d = mk(0, 0, 0); o.time = 0; }
  att[3 * i] = a.x; att[3 * i + 1] = a.y; att[3 * i + 2] = a.z;
  emit[3 * i] = e.x; emit[3 * i + 1] = e.y; emit[3 * i + 2] = e.z;
  store_ray(&out_rays[i], o);
  scattered[i] = ok ? 1 : 0;
}

__global__ void k_texture_kat(Scene sc, int tex, const float* __restrict__ uvp, long long n, float* out) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  V3 c = texture_value(sc, tex, uvp[5 * i], uvp[5 * i + 1], mk(uvp[5 * i + 2], uvp[5 * i + 3], uvp[5 * i + 4]));
  out[3 * i] = c.x; out[3 * i + 1] = c.y; out[3 * i + 2] = c.z;
}

__global__ void k_getray_kat(srt_camera cam, const float* __restrict__ in, long long n, srt_ray* out) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  Ray r = camera_get_ray(cam, in[5 * i], in[5 * i + 1], in[5 * i + 2], in[5 * i + 3], in[5 * i + 4]);
  store_ray(&out[i], r);
}

// (pixel, sample) -> the primary ray the renderer generates for it (Philox blocks 0 and 1)
__global__ void k_primary_rays(RenderArgs A, const int* __restrict__ pixel_xy, const int* __restrict__ sample, long long n,
                               srt_ray* out) {
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  int x = pixel_xy[2 * i], y = pixel_xy[2 * i + 1];
  Ray r = generate_ray(A, x, y, (uint32_t)(y * A.W + x), (uint32_t)sample[i]);
  store_ray(&out[i], r);
}

}  // namespace srt
