// srt_extend2.cuh — k_wf_extend2: the EXTEND kernel of the wavefront for trees that fit shared memory (configs 1-4).
//
// What the reference's result depends on (hittablelist.h:33-47 -> bvh.h:97-105 -> aabb.h:11-27):
// the LEAF nodes of the tree in pre-order, each one's box tested against [tMin, tCur] with aabb::hit's own arithmetic, its
// primitives tested when that box is hit, every accept overwriting (prim, tCur).  The inner nodes only cull: in a tree whose
// boxes are nested (a node's box contains its children's - bvh.h:88-94 builds them as unions) a leaf is reached by the
// recursion exactly when its OWN box test passes at that moment:
//   * aabb::hit is monotone under nesting for rays with finite non-zero direction components: fl(c - o) and fl(x / d) are
//     monotone in c, so the slab interval of a contained box lies inside the container's at every early-out stage - a
//     contained box can only be hit if the container is (same tMin, tMax);
//   * if an ancestor's test failed, nothing was accepted between that test and the leaf, so tCur is still the value the
//     ancestor was tested with, and the leaf's own test fails with it as well (induction over the leaves in pre-order; tCur
//     is NOT monotone - triangle::hit ignores tMax, model.h:128 - which is why the argument goes through "nothing accepted").
// So inner-node decisions need not be the reference's: any test that never rejects a box the reference would accept walks
// through a superset of the reference's nodes and reaches the same leaves with the same tCur.  This kernel uses that:
//   walk  : a CONSERVATIVE slab test on outward-rounded boxes, no exact re-evaluation, no undecided branch - 6 FFMA, 12
//           FMNMX, one subtract, one compare, one select per node; both links of a node (next if kept / next if culled) are
//           stored with it, so the step has no control logic.  Nodes are staged per CTA either as 32 bytes (fp32 boxes
//           inflated by 2^-20 |c|) or as 16 bytes (boxes relative to the scene's centre, rounded outward to fp16: one
//           LDS.128 per step instead of two - the loop is bound by shared-memory wavefronts, profiles/r2_extend2.md).
//   leaf  : a lane that reaches a leaf PARKS the ray (16 bytes: slot, node, tCur, prim) in its warp's shared-memory list and
//           takes another ray at once - a parked one whose leaf is done, or the next ray of the wave.  When 32 rays are
//           parked the whole warp tests 32 leaves together, one per lane, with the reference's arithmetic: aabb::hit on the
//           leaf's true box (box_hit_exact), then triangle::hit / sphere::hit in order.  Walking lanes never wait for a leaf
//           and leaf tests run with full warps (k_wf_extend: 2-8 lanes, 25 % of its warp instructions).
// Rays the conservative test cannot serve (zero / non-finite / extreme direction components) are traced by the exact
// ordered walk (traverse(), srt_core.cuh) by the lane that fetched them.  Scenes whose tree is not nested, has box-less
// primitives under inner nodes, or coordinates beyond 2^40 keep k_wf_extend.
//
// Error budget of the conservative test (u = 2^-24): with id = rcp.approx(d) (<= 2u), o' = fl(o - centre), oid = fl(o' id),
// the slab value fma(c', id, -oid) differs from the reference's fl(fl(c - o) / d) by at most 5.03u |c'| |id| + 7.05u |o' id|
// when c' is exact.  The first term is covered by inflating every coordinate by 2^-20 |c'| (16u) before it is rounded
// outward, the second by culling only when h - l < -(2^-19 M + 1e-30), M = max_k |oid_k| (32u against 2 x 7.05u).
#pragma once
#include <cuda_fp16.h>
#include "srt_wavefront.cuh"

namespace srt {

#define X2_DONE (-1)                 /* lane state: the walk is finished                                     */
/* lane state <= -2: stopped on a leaf, code = -2 - state = leaf node | (leaf holds a sphere) << X2_KIND_SHIFT;
   >= 0: walking (FMT 32: byte offset of the node, FMT 16: node index) */
#define X2_KIND_SHIFT(FMT) ((FMT) == 32 ? 24 : 14)
#define X2_MAX_NODES(FMT) ((FMT) == 32 ? (1 << 24) - 1 : 16382)   /* FMT 16: -2 - (16382 | 1 << 14) = -32768 still fits 16 bits */
#define X2_SPH_PARK 64               /* k_wf_path: entries of a warp's sphere-leaf list (it never holds more than 63)  */

struct Aux2 { float idx, idy, idz, oidx, oidy, oidz, nMs; };

__device__ __forceinline__ float rcp_approx(float v) { float y; asm("rcp.approx.f32 %0, %1;" : "=f"(y) : "f"(v)); return y; }

// -> false: the ray must take the exact walk
__device__ __forceinline__ bool make_aux2(const Ray& r, const float* centre, Aux2& a) {
  a.idx = rcp_approx(r.d.x); a.idy = rcp_approx(r.d.y); a.idz = rcp_approx(r.d.z);
  a.oidx = (r.o.x - centre[0]) * a.idx; a.oidy = (r.o.y - centre[1]) * a.idy; a.oidz = (r.o.z - centre[2]) * a.idz;
  const float ix = fabsf(a.idx), iy = fabsf(a.idy), iz = fabsf(a.idz);
  const float ox = fabsf(a.oidx), oy = fabsf(a.oidy), oz = fabsf(a.oidz);
  const float M = fmaxf(ox, fmaxf(oy, oz));
  a.nMs = -fmaf(1.9073486e-6f, M, 1e-30f);
  // every comparison is false for NaN: 2^-100 <= |1/d| <= 2^40, |o'/d| <= 2^60
  const float LO = 7.8886091e-31f, HI = 1.0995116e12f, OM = 1.1529215e18f;
  return ix >= LO && ix <= HI && iy >= LO && iy <= HI && iz >= LO && iz <= HI && ox <= OM && oy <= OM && oz <= OM && r.time == r.time;
}

// the conservative decision on one box given as (lo.xyz, hi.xyz) in the walk's coordinates: true = certainly missed
__device__ __forceinline__ bool cull_certain(float lx, float ly, float lz, float hx, float hy, float hz, const Aux2& x, float tCur) {
  const float ax = fmaf(lx, x.idx, -x.oidx), bx = fmaf(hx, x.idx, -x.oidx);
  const float ay = fmaf(ly, x.idy, -x.oidy), by = fmaf(hy, x.idy, -x.oidy);
  const float az = fmaf(lz, x.idz, -x.oidz), bz = fmaf(hz, x.idz, -x.oidz);
  const float l = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.001f));
  const float h = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), tCur));
  return (h - l) < x.nMs;
}

// outward rounding of one box coordinate into the walk's coordinates (relative to the centre, inflated by 2^-20 |c'|)
__device__ __forceinline__ float x2_lo(float c, float centre) { const float v = __fsub_rd(c, centre); return __fmaf_rd(-9.5367432e-7f, fabsf(v), v); }
__device__ __forceinline__ float x2_hi(float c, float centre) { const float v = __fsub_ru(c, centre); return __fmaf_ru(9.5367432e-7f, fabsf(v), v); }

__device__ __forceinline__ unsigned pack_h2(__half a, __half b) { return (unsigned)__half_as_ushort(a) | ((unsigned)__half_as_ushort(b) << 16); }

// One node of the walk tree (Scene::x2_nodes: the caller's threaded tree with low-value inner nodes left out) -> the walk's record.  Links: >= 0 next node, X2_DONE, or
// (-2 - k) = stop on leaf k.
template <int FMT>
__device__ __forceinline__ void x2_encode(const Scene& sc, int k, uint4* s_nodes, bool two_lists = true, int hi_index = -1) {
  const int n = sc.x2_n;
  const float4 lo = __ldg(&sc.x2_nodes[k].lo), hi = __ldg(&sc.x2_nodes[k].hi);
  const int w6 = __float_as_int(hi.z), w7 = __float_as_int(hi.w);
  const bool inner = w7 == NODE_INNER;
  const int scale = FMT == 32 ? 16 : 1;
  const int nxt = k + 1 < n ? (k + 1) * scale : X2_DONE;
  // a leaf's stop code carries whether it holds a sphere (its parked ray then waits in the sphere list)
  const bool has_sph = two_lists && !inner && (w6 >= sc.n_tri || w7 >= sc.n_tri);
  const int link_hit = inner ? nxt : -2 - (k | (has_sph ? 1 << X2_KIND_SHIFT(FMT) : 0));
  const int link_miss = inner ? (w6 < n ? w6 * scale : X2_DONE) : nxt;
  const float* c = sc.x2_centre;
  const float lx = x2_lo(lo.x, c[0]), ly = x2_lo(lo.y, c[1]), lz = x2_lo(lo.z, c[2]);
  const float hx = x2_hi(lo.w, c[0]), hy = x2_hi(hi.x, c[1]), hz = x2_hi(hi.y, c[2]);
  if (FMT == 32) {
    s_nodes[k] = make_uint4(__float_as_uint(lx), __float_as_uint(ly), __float_as_uint(lz), __float_as_uint(hx));
    s_nodes[(hi_index < 0 ? n : hi_index) + k] = make_uint4(__float_as_uint(hy), __float_as_uint(hz), (unsigned)link_miss, (unsigned)link_hit);
  } else {
    s_nodes[k] = make_uint4(pack_h2(__float2half_rd(lx), __float2half_rd(ly)), pack_h2(__float2half_rd(lz), __float2half_ru(hx)),
                            pack_h2(__float2half_ru(hy), __float2half_ru(hz)), ((unsigned)link_miss << 16) | ((unsigned)link_hit & 0xffffu));
  }
}

__device__ __forceinline__ uint4 lds_u4(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
  return v;
}

template <bool COUNT, int FMT>
__device__ __forceinline__ void walk2_step(uint32_t nodes_s, uint32_t hi_off, const Aux2& x, float tCur, int& i, Counters& cnt) {
  if (COUNT) ++cnt.nodes;
  if (FMT == 32) {
    const uint4 a = lds_u4(nodes_s + (uint32_t)i), b = lds_u4(nodes_s + hi_off + (uint32_t)i);
    const bool miss = cull_certain(__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(a.z), __uint_as_float(a.w),
                                   __uint_as_float(b.x), __uint_as_float(b.y), x, tCur);
    i = miss ? (int)b.z : (int)b.w;
  } else {
    const uint4 q = lds_u4(nodes_s + (uint32_t)i * 16u);
    const float2 p0 = __half22float2(*reinterpret_cast<const __half2*>(&q.x));
    const float2 p1 = __half22float2(*reinterpret_cast<const __half2*>(&q.y));
    const float2 p2 = __half22float2(*reinterpret_cast<const __half2*>(&q.z));
    const bool miss = cull_certain(p0.x, p0.y, p1.x, p1.y, p2.x, p2.y, x, tCur);
    const int w = (int)q.w;
    i = miss ? (w >> 16) : (int)(short)(w & 0xffff);
  }
}

#ifndef SRT_X2_NB
#define SRT_X2_NB 1     /* leaf tests without early returns (A/B builds: 0 = the early-return forms of srt_core.cuh) */
#endif
// aabb::hit / triangle::hit as in srt_core.cuh (same operations, same order, same NaN behaviour) without the early returns:
// in a leaf phase some lane nearly always needs the last stage, so the returns only cost branches.
__device__ __forceinline__ bool box_hit_exact_nb(const float4& lo, const float4& hi, const Ray& r, float tMin, float tMax) {
  float a, b;
  a = (lo.x - r.o.x) / r.d.x; b = (lo.w - r.o.x) / r.d.x;
  tMin = fmaxf(fminf(a, b), tMin); tMax = fminf(fmaxf(a, b), tMax);
  bool miss = tMax <= tMin;
  a = (lo.y - r.o.y) / r.d.y; b = (hi.x - r.o.y) / r.d.y;
  tMin = fmaxf(fminf(a, b), tMin); tMax = fminf(fmaxf(a, b), tMax);
  miss = miss || tMax <= tMin;
  a = (lo.z - r.o.z) / r.d.z; b = (hi.y - r.o.z) / r.d.z;
  tMin = fmaxf(fminf(a, b), tMin); tMax = fminf(fmaxf(a, b), tMax);
  return !(miss || tMax <= tMin);
}
__device__ __forceinline__ bool tri_hit_nb(const Tri* __restrict__ tris, int i, const Ray& r, float tMin, float& tOut) {
  const float4 A = __ldg(&tris[i].a), B = __ldg(&tris[i].b), C = __ldg(&tris[i].c);
  const float dd = __ldg(&tris[i].d.x);
  const V3 N = mk(A.w, B.w, C.w);
  const float NdotDir = dot(N, r.d);
  // model.h:119,122,128: the plane stage still returns - half of the candidates are back faces
  if (fabsf(NdotDir) < FLT_EPSILON || NdotDir > 0) return false;
  const float t = -(dot(N, r.o) + dd) / NdotDir;
  if (t < tMin) return false;
  const V3 v0 = mk(A.x, A.y, A.z), v1 = mk(B.x, B.y, B.z), v2 = mk(C.x, C.y, C.z);
  const V3 p = mk(r.o.x + t * r.d.x, r.o.y + t * r.d.y, r.o.z + t * r.d.z);
  const float e0 = dot(N, cross(v1 - v0, p - v0)), e1 = dot(N, cross(v2 - v1, p - v1)), e2 = dot(N, cross(v0 - v2, p - v2));
  tOut = t;
  return !(e0 < 0) && !(e1 < 0) && !(e2 < 0);           // model.h:136-154
}

// the leaf a parked ray waits on, with the reference's arithmetic: aabb::hit on the true box, then its primitives in order.
// SPH = false: the leaf holds triangles only (the warp's triangle list) - no sphere code in that phase at all.
template <bool COUNT, bool SPH>
__device__ __forceinline__ void x2_leaf(const Scene& sc, int node, const Ray& r, float& tCur, int& prim, Counters& cnt) {
  float4 lo, hi;
  load_node_soa<false>(sc.x2_nodes, 0u, 0u, node, lo, hi);
  if (COUNT) ++cnt.exact;
#if SRT_X2_NB
  if (!box_hit_exact_nb(lo, hi, r, 0.001f, tCur)) return;
#else
  if (!box_hit_exact(lo, hi, r, 0.001f, tCur)) return;
#endif
  const int pa = __float_as_int(hi.z), pb = __float_as_int(hi.w);
#pragma unroll 1
  for (int k = 0; k < 2; k++) {
    const int p = k == 0 ? pa : pb;
    if (p < 0) break;
    if (COUNT) { ++cnt.leaves; if (p >= sc.n_tri) ++cnt.sph_leaves; }
    float t;
    bool acc;
#if SRT_X2_NB
    if (SPH) acc = p < sc.n_tri ? tri_hit_nb(sc.tris, p, r, 0.001f, t) : sph_hit(sc.sphs, p - sc.n_tri, r, 0.001f, tCur, t);
    else acc = tri_hit_nb(sc.tris, p, r, 0.001f, t);
#else
    if (SPH) acc = p < sc.n_tri ? tri_hit(sc.tris, p, r, 0.001f, t) : sph_hit(sc.sphs, p - sc.n_tri, r, 0.001f, tCur, t);
    else acc = tri_hit(sc.tris, p, r, 0.001f, t);
#endif
    if (acc) { tCur = t; prim = p; }
  }
}

// One leaf phase of a warp: the top `take` <= 32 entries of the pending list, one per lane; rays whose walk goes on are
// pushed on the ready stack (ready_top[-1 - k]); a ray whose leaf was the last node has its hit stored at once.
template <bool COUNT>
__device__ __forceinline__ void x2_leaf_phase(const Scene& sc, const WfPool& pool, uint4* list, int& n_list, uint4* ready_top, int& r,
                                              int lane, int n, int step_scale, Counters& cnt) {
  const unsigned FULL = 0xffffffffu, below = (1u << lane) - 1u;
  const int take = n_list < 32 ? n_list : 32;
  bool push = false;
  uint4 res = make_uint4(0, 0, 0, 0);
  if (lane < take) {
    const uint4 q = list[n_list - take + lane];
    const float4 d = __ldg(&pool.ray_d[q.x]), o = __ldg(&pool.ray_o[q.x]);
    Ray ry; ry.o = mk(o.x, o.y, o.z); ry.time = o.w; ry.d = mk(d.x, d.y, d.z);
    float tc = __uint_as_float(q.z);
    int pr = (int)q.w;
    x2_leaf<COUNT, true>(sc, (int)q.y, ry, tc, pr, cnt);
    const int nxt = (int)q.y + 1;
    if (nxt >= n) pool.hit[q.x] = make_int2(pr, __float_as_int(tc));          // the leaf was the last node: the walk is over
    else { push = true; res = make_uint4(q.x, (unsigned)(nxt * step_scale), __float_as_uint(tc), (unsigned)pr); }
  }
  n_list -= take;
  const unsigned mp = __ballot_sync(FULL, push);      // (also orders the reads above before the writes below)
  if (push) ready_top[-1 - r - __popc(mp & below)] = res;
  r += __popc(mp);
  __syncwarp();
}

#ifndef SRT_X2_REPS
#define SRT_X2_REPS 2             /* box tests per vote (B200, config 4, th_walk 8: 2 -> 1.385 ms per launch, 4 -> 1.403) */
#endif
#ifndef SRT_X2_DEFAULT_FMT
#define SRT_X2_DEFAULT_FMT 32     /* node bytes plan_render asks for first: 16, 32, or 0 = keep k_wf_extend */
#endif
#ifndef SRT_X2_DEFAULT_TH
#define SRT_X2_DEFAULT_TH 8       /* lanes that must still be walking for the walk loop to go on (B200 sweep: 8 < 12 < 16 < 24) */
#endif
#define SRT_X2_PARK 64            /* park entries per warp */
#ifndef SRT_X2_PREFETCH
#define SRT_X2_PREFETCH 0
#endif

// Shared memory: the staged nodes (FMT 32: lo[n] | hi[n], FMT 16: q[n]) followed by `park` uint4 entries per warp:
// pending entries (slot, leaf node, tCur, prim) grow from index 0, ready entries (slot, next state, tCur, prim) from the top.
// A warp never holds more than 63 parked rays (fresh rays are fetched only when no ready one is left and fewer than 32 are
// pending), so park = 64 always suffices: 32 KB per CTA, which keeps config 4's 129 KB tree inside the 164 KB shared-memory
// step of the SM - i.e. leaves 64 KB of L1 for the rays and primitives the leaf phases re-read.  Measured alternatives
// (profiles/r2_extend2_ab.md): separate lists for leaves with / without a sphere make the phases 1.4 % cheaper but need 160
// entries per warp: with 16-byte entries (209 KB, 19 KB of L1) 1.50 ms per launch against 1.385; with 8-byte entries and the
// ray's (prim, tCur) kept in pool.hit (170 KB, the 196 KB step) 1.43-1.45.
// HIOFF (FMT 32 only, A/B builds): byte offset of the hi[] half of the nodes as a compile-time constant - the second LDS.128
// of a step then takes an immediate offset instead of an add per step.  Measured with 65536 on config 4: 1.381 vs 1.345 ms
// per launch (lo[k] and hi[k] then sit in the same banks); the host always uses 0 = hi[] follows lo[] at n * 16.
template <bool COUNT, int FMT, int THREADS, int HIOFF = 0>
__global__ void __launch_bounds__(THREADS, 1) k_wf_extend2(Scene sc, WfPool pool, WfCursors* cur, int th_walk, int park,
                                                          unsigned long long* counters) {
  extern __shared__ uint4 smem_x2[];
  const int n = sc.x2_n;
  for (int k = threadIdx.x; k < n; k += THREADS) x2_encode<FMT>(sc, k, smem_x2, false, HIOFF ? HIOFF / 16 : -1);
  const int node_u4 = FMT == 32 ? (HIOFF ? HIOFF / 16 + n : 2 * n) : n;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint4* s_park = smem_x2 + node_u4 + warp * park;
  uint4* s_top = s_park + park;                     // ready entries: s_top[-1 - k]
  const uint32_t nodes_s = (uint32_t)__cvta_generic_to_shared(smem_x2);
  const uint32_t hi_off = HIOFF ? (uint32_t)HIOFF : (uint32_t)n * 16u;
  __syncthreads();
  const unsigned FULL = 0xffffffffu, below = (1u << lane) - 1u;
  const float INF = __int_as_float(0x7f800000);
  const int root = n > 0 ? 0 : X2_DONE;
  const int step_scale = FMT == 32 ? 16 : 1;
  Counters cnt = {0, 0, 0, 0, 0};
  if (blockIdx.x == 0 && threadIdx.x == 0) cur->alive = 0;   // set anew by the next shade pass

  Aux2 x = {0, 0, 0, 0, 0, 0, 0};
  int i = X2_DONE, slot = -1, prim = -1;
  float tCur = INF;
  int p = 0, r = 0;                 // warp-uniform: pending entries [0, p), ready entries [park - r, park)
  bool pool_dry = false;            // warp-uniform: the wave's rays are all handed out

  for (;;) {
    // ---- A. lanes whose walk ended store the hit; lanes stopped on a leaf park their ray
    const bool has = slot >= 0;
    const bool stp = has && i <= -2;
    if (has && i == X2_DONE) { pool.hit[slot] = make_int2(prim, __float_as_int(tCur)); slot = -1; }
    const unsigned ms = __ballot_sync(FULL, stp);
    if (stp) {
      s_park[p + __popc(ms & below)] = make_uint4((unsigned)slot, (unsigned)(-2 - i), __float_as_uint(tCur), (unsigned)prim);
#if SRT_X2_PREFETCH
      asm volatile("prefetch.global.L1 [%0];" :: "l"(sc.x2_nodes + (-2 - i)));      // the leaf's true box + primitive ids, for the phase
#endif
      slot = -1; i = X2_DONE;
    }
    p += __popc(ms);
    __syncwarp();
    // ---- B. leaf phase: 32 parked rays, one per lane.
    // The end of a wave (no ray left to fetch, none ready): a phase for every idle lane would test one or two leaves at a
    // time, each behind three dependent global loads (64 such phases per warp measured).  Instead the walkers run on until
    // every lane has stopped, and one phase serves them all.
    if (p >= 32 || (p > 0 && pool_dry && r == 0 && __ballot_sync(FULL, slot < 0) == FULL))
      x2_leaf_phase<COUNT>(sc, pool, s_park, p, s_top, r, lane, n, step_scale, cnt);
    // ---- C. empty lanes take a ready ray, else the next ray of the wave (one atomic per warp)
    const unsigned m_empty = __ballot_sync(FULL, slot < 0);
    if (m_empty) {
      const int n_empty = __popc(m_empty), q = __popc(m_empty & below);
      const int take = r < n_empty ? r : n_empty;
      const int fresh = pool_dry ? 0 : n_empty - take;
      unsigned long long base = 0;
      if (fresh) {
        const int leader = __ffs(m_empty) - 1;
        if (lane == leader) base = atomicAdd(&cur->ray, (unsigned long long)fresh);
        base = __shfl_sync(FULL, base, leader);
        if (base + fresh >= (unsigned long long)pool.n_slots) pool_dry = true;
      }
      if (slot < 0) {
        long long mine = -1;
        int link = root;
        bool is_fresh = false;
        if (q < take) {
          const uint4 e = s_top[-1 - (r - take + q)];
          mine = (long long)e.x; link = (int)e.y; tCur = __uint_as_float(e.z); prim = (int)e.w;
        } else if (q - take < fresh) {
          const unsigned long long idx = base + (unsigned long long)(q - take);
          if (idx < (unsigned long long)pool.n_slots) { mine = pool.order ? (long long)pool.order[idx] : (long long)idx; tCur = INF; prim = -1; is_fresh = true; }
        }
        if (mine >= 0) {
          const float4 d = __ldg(&pool.ray_d[mine]), o = __ldg(&pool.ray_o[mine]);
          if (!is_fresh || __float_as_int(d.w) == 1) {
            Ray ry; ry.o = mk(o.x, o.y, o.z); ry.time = o.w; ry.d = mk(d.x, d.y, d.z);
            if (make_aux2(ry, sc.x2_centre, x)) { slot = (int)mine; i = link; }
            else {                                   // only a fresh ray can land here: the exact ordered walk, start to end
              int pr; float t;
              traverse<COUNT>(sc, sc.nodes, ry, 0.001f, INF, pr, t, &cnt);
              pool.hit[mine] = make_int2(pr, __float_as_int(t));
            }
          }
        }
      }
      r -= take;
    }
    __syncwarp();
    // ---- D. walk while enough lanes are walking
    const unsigned mW = __ballot_sync(FULL, i >= 0);
    if (!mW) {
      if (__ballot_sync(FULL, slot >= 0)) continue;          // finished / stopped lanes only: serve them
      if (p == 0 && r == 0 && pool_dry) break;
      continue;
    }
    const int th = (pool_dry && r == 0) ? 1 : (__popc(mW) < th_walk ? __popc(mW) : th_walk);   // nothing to refill with: walk until all have stopped
    do {
#pragma unroll
      for (int rep = 0; rep < SRT_X2_REPS; rep++)
        if (i >= 0) walk2_step<COUNT, FMT>(nodes_s, hi_off, x, tCur, i, cnt);
    } while (__popc(__ballot_sync(FULL, i >= 0)) >= th);
  }

  if (COUNT) {
    for (int o = 16; o > 0; o >>= 1) {
      cnt.nodes += __shfl_down_sync(FULL, cnt.nodes, o);
      cnt.leaves += __shfl_down_sync(FULL, cnt.leaves, o);
      cnt.exact += __shfl_down_sync(FULL, cnt.exact, o);
      cnt.sph_leaves += __shfl_down_sync(FULL, cnt.sph_leaves, o);
    }
    if (lane == 0) {
      atomicAdd(&counters[2], cnt.nodes); atomicAdd(&counters[3], cnt.leaves); atomicAdd(&counters[4], cnt.exact);
      atomicAdd(&counters[6], cnt.sph_leaves);
    }
  }
}

// ---- known-answer kernel for the box predicates (srt_aabb_kat) ---------------------------------------------------------
// One (ray, box, tMin, tMax) per thread.  predicate 0: box_hit<> = the filtered predicate of the exact walks with its
// fallback; 1: cbox_filter on the box snapped outward to the 16-bit grid `g`, fallback on the full box; 2 / 3: the
// conservative cull of k_wf_extend2 on 32-byte / 16-byte nodes (decision 1 = kept, 0 = culled).
// path: 0 = the cheap evaluation decided, 1 = the reference's arithmetic ran (0/1), or the ray is not served by the
// conservative walk (2/3).
struct KatGrid { float org[3], cell[3], centre[3]; };
__global__ void k_aabb_kat(const srt_ray* __restrict__ rays, const float* __restrict__ boxes, const float* __restrict__ tmm, long long n,
                           int predicate, KatGrid g, int* __restrict__ decision, int* __restrict__ path) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  Ray r; r.o = mk(rays[i].o[0], rays[i].o[1], rays[i].o[2]); r.d = mk(rays[i].dir[0], rays[i].dir[1], rays[i].dir[2]); r.time = rays[i].time;
  const float* b = boxes + 6 * i;
  const float4 lo = make_float4(b[0], b[1], b[2], b[3]), hi = make_float4(b[4], b[5], 0.0f, 0.0f);
  const float tMin = tmm[2 * i], tMax = tmm[2 * i + 1];
  int dec = 0, pth = 0;
  if (predicate == 0) {
    const RayAux aux = make_aux(r, 1);
    unsigned long long ex = 0;
    dec = box_hit<true>(lo, hi, r, aux, tMin, tMax, &ex) ? 1 : 0;
    pth = (int)ex;
  } else if (predicate == 1) {
    Scene sc; sc.q_org[0] = g.org[0]; sc.q_org[1] = g.org[1]; sc.q_org[2] = g.org[2];
    sc.q_cell[0] = g.cell[0]; sc.q_cell[1] = g.cell[1]; sc.q_cell[2] = g.cell[2];
    const RayAuxC aux = make_aux_c(sc, r);
    unsigned q6[6];
    for (int k = 0; k < 3; k++) {      // outward snap, as srt_scene_create does on the host (floor / ceil in double)
      const double l = floor(((double)b[k] - (double)g.org[k]) / (double)g.cell[k]), h = ceil(((double)b[3 + k] - (double)g.org[k]) / (double)g.cell[k]);
      q6[k] = (unsigned)fmin(fmax(l, 0.0), 65535.0); q6[3 + k] = (unsigned)fmin(fmax(h, 0.0), 65535.0);
    }
    const uint4 q = make_uint4(q6[0] | (q6[1] << 16), q6[2] | (q6[3] << 16), q6[4] | (q6[5] << 16), 0u);
    const int f = cbox_filter(q, aux, tMin, tMax);
    if (f == 0) { pth = 1; dec = box_hit_exact(lo, hi, r, tMin, tMax) ? 1 : 0; }
    else dec = f > 0 ? 1 : 0;
  } else {
    Aux2 x;
    const bool ok = make_aux2(r, g.centre, x);
    float c[6] = {x2_lo(b[0], g.centre[0]), x2_lo(b[1], g.centre[1]), x2_lo(b[2], g.centre[2]),
                  x2_hi(b[3], g.centre[0]), x2_hi(b[4], g.centre[1]), x2_hi(b[5], g.centre[2])};
    if (predicate == 3) {
      for (int k = 0; k < 3; k++) { c[k] = __half2float(__float2half_rd(c[k])); c[3 + k] = __half2float(__float2half_ru(c[3 + k])); }
    }
    // the walk's tMin is the constant 0.001 (main.cpp:39): the KAT feeds rows with that tMin only
    dec = ok ? (cull_certain(c[0], c[1], c[2], c[3], c[4], c[5], x, tMax) ? 0 : 1) : 1;
    pth = ok ? 0 : 1;
  }
  decision[i] = dec;
  if (path) path[i] = pth;
}

// explicit rays -> pool slots (srt_trace_rays_wavefront)
__global__ void k_wf_load_rays(const srt_ray* __restrict__ rays, long long n, WfPool pool) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  pool.ray_o[i] = make_float4(rays[i].o[0], rays[i].o[1], rays[i].o[2], rays[i].time);
  pool.ray_d[i] = make_float4(rays[i].dir[0], rays[i].dir[1], rays[i].dir[2], __int_as_float(1));
  pool.hit[i] = make_int2(-2, 0);
}

}  // namespace srt
