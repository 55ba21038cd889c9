// srt_wavefront.cuh — variant 2: the frame as alternating EXTEND / SHADE kernels over a pool of path slots.
//
// Why (profiles/r1_v1_megakernel_summary.md, profiles/r1_v3_phased_summary.md): in the megakernels a lane
// that finished its traversal must wait for shading code (~450 instructions + ray generation) that the
// rest of its warp does not need, and leaf tests run with 2-3 lanes.  Here
//   k_wf_extend : ONLY hittableList::hit (bvh.h:97-105 + leaf tests).  Persistent warps; a lane whose ray
//                 is finished stores (prim, t) and fetches the next ray of the pool (warp-aggregated
//                 atomic), so the warp never waits for a long ray; leaf tests are batched by a warp vote.
//   k_wf_shade  : main.cpp:44-51 for every slot with all lanes active, then path regeneration
//                 (main.cpp:204-217): a slot whose path ended adds its radiance to the pixel
//                 (red.global.add.f32) and takes the next (pixel, sample) of the frame from a global cursor.
// A path's radiance is not carried: every non-zero term thr * emitted / thr * background is added to the pixel
// (red.global.add.f32) where it arises - a path of configs 1-4 has one such term, so the pixel sum sees the same addends.
// Path state lives in HBM between the two (SoA, ~150 B of traffic per segment — two orders of magnitude
// below what 6.5 TB/s allows at this path's segment rates).  Traversal order per ray, arithmetic and the
// Philox keying (pixel, sample, bounce) are the megakernel's: frames agree up to fp32 summation order
// of a pixel's samples.
#pragma once
#include "srt_phased.cuh"

// build-time tuning switches (A/B builds: tools/b.sh -DSRT_...=...)
#ifndef SRT_FMNMX3
#define SRT_FMNMX3 0
#endif
#ifndef SRT_WALK_REPS
#define SRT_WALK_REPS 2
#endif

namespace srt {

struct WfPool {
  float4* ray_o;      // (o.xyz, time)
  float4* ray_d;      // (d.xyz, active flag as int bits: 1 = a segment is waiting to be traced)
  int2*   hit;        // (prim, t bits) written by extend
  uint2*  id;         // (pixel, sample)
  float4* thr;        // (throughput.xyz, bounce as int bits; bounce < 0 = slot has no path)
  int     n_slots;
  const unsigned* order;   // extend hands out slots order[0], order[1], ... (nullptr: 0, 1, ...) - see k_wf_keys
};

// cursors (device memory, zeroed per frame): [0] next path id, [1] extend ray cursor (zeroed per wave),
// [2] slots still alive after the last shade, [3] unused
struct WfCursors { unsigned long long path, ray, alive, pad; };

// One box test of the ordered traversal, written without data-dependent branches (the exact
// re-evaluation of an undecided filtered test is the one rare exception): the loop is issue bound, so
// control is a handful of predicate operations and two selects.
// The node array carries a SENTINEL at index n (NODE_NO_BOX leaf without primitives): every ray ends by
// "hitting" it - no end-of-array compare per step.  A lane that hits a leaf's box stops walking with i
// still on that leaf; the service phase reads its primitives from the node again.
// Lane state is ONE integer: i >= 0 = walking, next node i;  i < 0 = stopped on node ~i.
// `nodes_s` is the 32-bit shared-window address of the staged tree (SMEM) - LDS needs no generic->shared
// conversion per step that way.
//
// Shared-memory layout (SMEM): the two 16-byte halves of the nodes are staged as two arrays lo[n+1], hi[n+1]
// instead of 32-byte records.  The loop is bound by shared-memory wavefronts (every lane reads a different
// node): with 32-byte records all `lo` halves fall into 4 of the 8 bank groups, so the 8 lanes of a quarter
// warp collide ~3.4 deep; with 16-byte stride they spread over all 8 (~2.4 deep).
template <bool SMEM>
__device__ __forceinline__ void load_node_soa(const Node* nodes, uint32_t nodes_s, uint32_t hi_off, int i, float4& lo, float4& hi) {
  if (SMEM) {
    const uint32_t a = nodes_s + (uint32_t)i * 16u;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w) : "r"(a));
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w) : "r"(a + hi_off));
  } else {     // one 256-bit load: a node is one 32-byte sector
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w) : "l"(nodes + i));
  }
}

// cooperative staging of the (n + 1)-node array as lo[] | hi[]
__device__ __forceinline__ void stage_nodes_soa(const Scene& sc, float4* smem) {
  const float4* src = reinterpret_cast<const float4*>(sc.nodes);
  const int n1 = sc.n_nodes + 1;
  for (int k = threadIdx.x; k < n1 * 2; k += blockDim.x) smem[(k & 1) * n1 + (k >> 1)] = __ldg(&src[k]);
  __syncthreads();
}

template <bool COUNT, bool SMEM>
__device__ __forceinline__ void walk_step(const Node* nodes, uint32_t nodes_s, uint32_t hi_off, int n, const Ray& r, const RayAux& x,
                                          int& i, float tCur, Counters& cnt) {
  float4 lo, hi;
  load_node_soa<SMEM>(nodes, nodes_s, hi_off, i, lo, hi);
  const int w6 = __float_as_int(hi.z), w7 = __float_as_int(hi.w);
  if (COUNT && i < n) ++cnt.nodes;
  const float ax = fmaf(lo.x, x.idx, -x.oidx), bx = fmaf(lo.w, x.idx, -x.oidx);
  const float ay = fmaf(lo.y, x.idy, -x.oidy), by = fmaf(hi.x, x.idy, -x.oidy);
  const float az = fmaf(lo.z, x.idz, -x.oidz), bz = fmaf(hi.y, x.idz, -x.oidz);
#if SRT_FMNMX3
  // three-input min/max (FMNMX3 on sm_100a; NaN -> the other operands, like fminf/fmaxf): 10 instead of 12 on the ALU pipe
  float l, h;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(l) : "f"(fminf(ax, bx)), "f"(fminf(ay, by)), "f"(fmaxf(fminf(az, bz), 0.001f)));
  asm("min.f32 %0, %1, %2, %3;" : "=f"(h) : "f"(fmaxf(ax, bx)), "f"(fmaxf(ay, by)), "f"(fminf(fmaxf(az, bz), tCur)));
#else
  const float l = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.001f));
  const float h = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), tCur));
#endif
  const float slack = fmaf(9.5367432e-7f, x.M + fabsf(l) + fabsf(h), 1e-30f);   // x.M = +inf when the filter is off
  const float diff = h - l;
  bool hit = diff > 0.0f;
  const bool nobox = w7 == NODE_NO_BOX;
  if (!(fabsf(diff) > slack) && !nobox) {          // undecided (or NaN, or filter off): the reference's own arithmetic
    if (COUNT) ++cnt.exact;
    hit = box_hit_exact(lo, hi, r, 0.001f, tCur);
  }
  hit = hit || nobox;
  const bool inner = w7 == NODE_INNER;
  const int next = (inner && !hit) ? w6 : i + 1;
  i = (hit && !inner) ? ~i : next;
}

// The same step over the compact tree (srt_core.cuh): one 16-byte load, the snapped-box filter, and the reference's
// arithmetic on the full node only for the undecided shell.
// An undecided test does NOT run here: a dependent HBM round trip for one or two lanes would stall the whole warp on
// ~40 % of its steps (profiles/r1_c5_extend_summary.md).  The lane stops on the node with `pend` set, the full node is
// prefetched, and the service phase evaluates aabb::hit together with the other stopped lanes.
template <bool COUNT>
__device__ __forceinline__ void walk_step_c(const Scene& sc, int n, const RayAuxC& x, int& i, bool& pend, float tCur, Counters& cnt) {
  const uint4 q = __ldg(&sc.cnodes[i]);
  if (COUNT && i < n) ++cnt.nodes;
  const bool inner = (q.w & CNODE_INNER) != 0;
  const bool nobox = !inner && (q.w & CNODE_NO_BOX);
  const int f = cbox_filter(q, x, 0.001f, tCur);
  const bool und = f == 0 && !nobox;
  const bool hit = f > 0 || nobox;
  if (und) { pend = true; asm volatile("prefetch.global.L2 [%0];" :: "l"(sc.nodes + i)); }
  const int next = (inner && !hit) ? (int)(q.w & 0x7fffffffu) : i + 1;
  i = (und || (hit && !inner)) ? ~i : next;
}

// ---- EXTEND ------------------------------------------------------------------------------------------
// MODE 0: tree staged in shared memory; 1: 32-byte nodes in global memory; 2: compact tree in global memory
template <bool COUNT, int MODE, int THREADS>
__global__ void __launch_bounds__(THREADS, (MODE != 0 && THREADS <= 512) ? 1536 / THREADS : (MODE != 0 && THREADS == 768) ? 2 : 1)
k_wf_extend(Scene sc, WfPool pool, WfCursors* cur, int th_walk,
                                                          unsigned long long* counters) {
  constexpr bool SMEM = MODE == 0;
  constexpr bool CPT = MODE == 2;
  extern __shared__ float4 smem_nodes[];
  const Node* nodes = sc.nodes;
  if (SMEM) stage_nodes_soa(sc, smem_nodes);
  const int lane = threadIdx.x & 31;
  const unsigned FULL = 0xffffffffu;
  const int n = sc.n_nodes;
  const uint32_t hi_off = (uint32_t)(n + 1) * 16u;
  const float INF = __int_as_float(0x7f800000);
  Counters cnt = {0, 0, 0, 0, 0};
  if (blockIdx.x == 0 && threadIdx.x == 0) cur->alive = 0;   // set anew by the next shade pass

  Ray r; r.o = r.d = mk(0, 0, 0); r.time = 0;
  RayAux aux = make_aux(r, 0);
  RayAuxC auxc = {0, 0, 0, 0, 0, 0, 0, 0};
  const uint32_t nodes_s = SMEM ? (uint32_t)__cvta_generic_to_shared(smem_nodes) : 0u;
  int slot = -1;                   // -1: this lane holds no ray
  int i = -1, prim = -1;           // i >= 0: walking, next node i;  i < 0 with a slot: stopped on node ~i (hit leaf, or sentinel = done)
  float tCur = INF;
  bool pend = false;               // compact tree: stopped on node ~i whose snapped box could not decide - aabb::hit is due
  bool pool_dry = false;           // warp-uniform: the wave's rays are all handed out

  for (;;) {
    // ---- service phase 1: lanes stopped on a leaf test its primitives in reference order; on the sentinel, store the hit
    if (slot >= 0 && i < 0) {
      const int at = ~i;
      if (at >= n) {
        pool.hit[slot] = make_int2(prim, __float_as_int(tCur));
        slot = -1;
      } else if (CPT) {
        const unsigned w = __ldg(&sc.cnodes[at].w);
        bool go = true;                           // the leaf's box is hit: test its primitives
        if (pend) {                               // the reference's own arithmetic on the full node
          pend = false;
          float4 lo, hi;
          load_node_soa<false>(nodes, 0u, 0u, at, lo, hi);
          if (COUNT) ++cnt.exact;
          go = box_hit_exact(lo, hi, r, 0.001f, tCur);
          if (w & CNODE_INNER) { i = go ? at + 1 : (int)(w & 0x7fffffffu); go = false; }
          else if (!go) i = at + 1;
        }
        if (go) {
          const unsigned u = w & CNODE_NO_PRIM;
          const int np = u == CNODE_NO_PRIM ? 0 : ((w & CNODE_TWO) ? 2 : 1);
#pragma unroll 1
          for (int k = 0; k < np; k++) {
            float t; int orig; bool sph;
            const bool acc = lprim_hit(sc.lprims, u + k, r, 0.001f, tCur, t, orig, sph);
            if (COUNT) { ++cnt.leaves; if (sph) ++cnt.sph_leaves; }
            if (acc) { tCur = t; prim = orig; }
          }
          i = at + 1;
        }
      } else {
        float4 lo, hi;
        load_node_soa<SMEM>(nodes, nodes_s, hi_off, at, lo, hi);
        const int pa = __float_as_int(hi.z), pb = __float_as_int(hi.w);
#pragma unroll 1
        for (int k = 0; k < 2; k++) {
          const int p = k == 0 ? pa : pb;
          if (p < 0) break;
          if (COUNT) { ++cnt.leaves; if (p >= sc.n_tri) ++cnt.sph_leaves; }
          float t;
          const bool acc = p < sc.n_tri ? tri_hit(sc.tris, p, r, 0.001f, t) : sph_hit(sc.sphs, p - sc.n_tri, r, 0.001f, tCur, t);
          if (acc) { tCur = t; prim = p; }
        }
        i = at + 1;
      }
    }
    // ---- service phase 2: idle lanes fetch the next rays of the wave (one atomic per warp)
    if (!pool_dry) {
      const unsigned need = __ballot_sync(FULL, slot < 0);
      if (need) {
        unsigned long long base = 0;
        const int leader = __ffs(need) - 1;
        if (lane == leader) base = atomicAdd(&cur->ray, (unsigned long long)__popc(need));
        base = __shfl_sync(FULL, base, leader);
        if (base + __popc(need) >= (unsigned long long)pool.n_slots) pool_dry = true;
        if (slot < 0) {
          unsigned long long mine = base + __popc(need & ((1u << lane) - 1));
          if (mine < (unsigned long long)pool.n_slots) {
            if (pool.order) mine = __ldg(&pool.order[mine]);
            const float4 d = __ldg(&pool.ray_d[mine]), o = __ldg(&pool.ray_o[mine]);   // both in flight together
            if (__float_as_int(d.w) == 1) {
              slot = (int)mine;
              r.o = mk(o.x, o.y, o.z); r.time = o.w; r.d = mk(d.x, d.y, d.z);
              if (CPT) auxc = make_aux_c(sc, r);
              else {
                aux = make_aux(r, sc.filter_ok);
                if (!aux.ok) aux.M = INF;        // walk_step then always takes the exact test
              }
              i = 0; tCur = INF; prim = -1; pend = false;
            }
          }
        }
      }
    }
    // ---- walk phase: box tests while enough lanes are walking
    const unsigned mW = __ballot_sync(FULL, i >= 0);
    if (!mW) {
      if (__ballot_sync(FULL, slot >= 0)) continue;    // only stopped lanes left: serve them
      if (pool_dry) break;
      continue;                                         // fetched nothing but dead slots: fetch again
    }
    do {
#pragma unroll
      for (int rep = 0; rep < SRT_WALK_REPS; rep++)     // steps per vote
        if (i >= 0) {
          if (CPT) walk_step_c<COUNT>(sc, n, auxc, i, pend, tCur, cnt);
          else walk_step<COUNT, SMEM>(nodes, nodes_s, hi_off, n, r, aux, i, tCur, cnt);
        }
    } while (__popc(__ballot_sync(FULL, i >= 0)) >= th_walk);
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

#ifndef SRT_SHADE_PREFETCH
#define SRT_SHADE_PREFETCH 1
#endif
#ifndef SRT_SHADE_MIN_BLOCKS
#define SRT_SHADE_MIN_BLOCKS 4   /* 64 registers, 52 B of spills: 4 CTAs per SM measured +3-4 % per frame over 3 (80 registers) */
#endif
// ---- SHADE + REGENERATE ----------------------------------------------------------------------------
// One thread per slot, persistent blocks striding over 256-slot chunks.  The frame's paths are numbered
// 0 .. n_tiles*32*sample_count-1: path id -> (sample offset, 8x4 tile, pixel of the tile), so that slots
// regenerated together get neighbouring pixels.  Path ids are allocated with ONE global atomic per chunk
// (a per-warp atomic on the single cursor address serialised the whole kernel: profiles/r1_wavefront.md);
// statistics stay in registers until the block ends.
// counters: [0] paths, [1] segments, [7] segments that ended on a triangle (COUNT only)
template <bool COUNT>
__global__ void __launch_bounds__(256, SRT_SHADE_MIN_BLOCKS) k_wf_shade(Scene sc, RenderArgs A, WfPool pool, WfCursors* cur, float* __restrict__ accum,
                                                  unsigned long long* counters) {
  __shared__ unsigned s_cnt[8];
  __shared__ unsigned long long s_base;
  __shared__ unsigned long long s_stat[3];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned FULL = 0xffffffffu;
  if (blockIdx.x == 0 && threadIdx.x == 0) cur->ray = 0;    // the next extend wave starts at slot 0 (no extend kernel runs now)
  if (threadIdx.x < 3) s_stat[threadIdx.x] = 0;
  const unsigned long long per_sample = (unsigned long long)A.n_tiles * 32ull;
  const unsigned long long total = per_sample * (unsigned long long)A.sample_count;
  Counters cnt = {0, 0, 0, 0, 0};
  unsigned n_new = 0, n_seg = 0;

  for (int chunk = blockIdx.x; chunk * 256 < pool.n_slots; chunk += gridDim.x) {
    const int k = chunk * 256 + threadIdx.x;
    const bool in_pool = k < pool.n_slots;
    int bounce = -1;
    uint32_t pixel = 0, sample = 0;
    V3 L = mk(0, 0, 0), thr = mk(1, 1, 1);
    Ray r; r.o = r.d = mk(0, 0, 0); r.time = 0;
    bool wrote_ray = false;
    if (in_pool) {
      // all six loads are issued before the first use (a slot without a path reads stale but valid memory)
      const float4 t4 = pool.thr[k], o = pool.ray_o[k], d = pool.ray_d[k];
      const int2 h = pool.hit[k];
      const uint2 id = pool.id[k];
      bounce = __float_as_int(t4.w);
      if (bounce >= 0) {
        pixel = id.x; sample = id.y;
        r.o = mk(o.x, o.y, o.z); r.time = o.w; r.d = mk(d.x, d.y, d.z);
        thr = mk(t4.x, t4.y, t4.z); L = mk(0, 0, 0);
        const bool done = shade_segment<COUNT, true>(sc, A, pixel, sample, bounce, r, h.x, __int_as_float(h.y), L, thr, cnt);
        if (L.x != 0.0f || L.y != 0.0f || L.z != 0.0f) {                           // this segment's term (NaN included)
          float* px = accum + (size_t)pixel * 3;
          atomicAdd(px, L.x); atomicAdd(px + 1, L.y); atomicAdd(px + 2, L.z);     // result unused -> red.global.add.f32
        }
        if (done) bounce = -1; else wrote_ray = true;
      }
    }
    // regeneration: slots without a path take the next path ids of the frame
    const bool want = in_pool && bounce < 0;
    const unsigned need = __ballot_sync(FULL, want);
    if (lane == 0) s_cnt[warp] = __popc(need);
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned run = 0;
#pragma unroll
      for (int w = 0; w < 8; w++) { const unsigned c = s_cnt[w]; s_cnt[w] = run; run += c; }
      s_base = run ? atomicAdd(&cur->path, (unsigned long long)run) : total;
    }
    __syncthreads();
    bool new_path = false;
    if (want) {
      const unsigned long long mine = s_base + s_cnt[warp] + __popc(need & ((1u << lane) - 1));
      if (mine < total) {
        const unsigned q = (unsigned)(mine % per_sample);
        const int s_off = (int)(mine / per_sample);
        const int tile = (int)(q >> 5), l = (int)(q & 31);
        const int x = (tile % A.tiles_x) * 8 + (l & 7), y = (tile / A.tiles_x) * 4 + (l >> 3);
        if (x < A.W && y < A.H) {
          pixel = (uint32_t)(y * A.W + x);
          sample = (uint32_t)(A.sample_begin + s_off);
          r = generate_ray<true>(A, x, y, pixel, sample);
          L = mk(0, 0, 0); thr = mk(1, 1, 1); bounce = 0;
          new_path = wrote_ray = true;
        }
      }
    }
    if (in_pool) {
      if (wrote_ray) {
        pool.ray_o[k] = make_float4(r.o.x, r.o.y, r.o.z, r.time);
        pool.ray_d[k] = make_float4(r.d.x, r.d.y, r.d.z, __int_as_float(1));
        pool.thr[k] = make_float4(thr.x, thr.y, thr.z, __int_as_float(bounce));
        if (new_path) pool.id[k] = make_uint2(pixel, sample);
      } else {                                    // no path: extend skips the slot, the next shade pass retries
        pool.ray_d[k] = make_float4(0, 0, 0, __int_as_float(0));
        pool.thr[k] = make_float4(0, 0, 0, __int_as_float(-1));
      }
    }
    n_new += new_path ? 1u : 0u;
    n_seg += wrote_ray ? 1u : 0u;
    __syncthreads();                              // s_cnt / s_base are reused by the next chunk
  }
  // statistics + liveness: one atomic per counter per block
  unsigned long long v0 = n_new, v1 = n_seg, v2 = cnt.tri_hits;
  for (int o = 16; o > 0; o >>= 1) {
    v0 += __shfl_down_sync(FULL, v0, o); v1 += __shfl_down_sync(FULL, v1, o);
    if (COUNT) v2 += __shfl_down_sync(FULL, v2, o);
  }
  if (lane == 0) { atomicAdd(&s_stat[0], v0); atomicAdd(&s_stat[1], v1); if (COUNT) atomicAdd(&s_stat[2], v2); }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (s_stat[0]) atomicAdd(&counters[0], s_stat[0]);
    if (s_stat[1]) { atomicAdd(&counters[1], s_stat[1]); cur->alive = 1; }     // flag, not a count
    if (COUNT && s_stat[2]) atomicAdd(&counters[7], s_stat[2]);
  }
}

// ---- SHADE + REGENERATE, sorted phases --------------------------------------------------------------------------
// k_wf_shade runs one thread per slot: after a wave ~45 % of the slots hold a miss (a dozen instructions), ~55 % a hit
// (emitted + scatter + textures, ~450) and more than half of them then regenerate (two Philox blocks + camera::getRay) -
// 16 of 32 lanes active on average (profiles/r1_wavefront_summary.md).  Here a block takes 512 slots at a time and runs the
// three jobs as separate passes over COMPACTED index lists in shared memory, so every pass runs with full warps:
//   pass 0  (thread = slot)      : classify; misses add the background and retire at once; hits are listed
//   pass 1  (thread = list item) : shade_segment for the listed hits; paths that end retire
//   pass 2  (thread = list item) : retired / empty slots take the next path ids of the frame and generate their ray
// Same arithmetic, same Philox keys; only which thread serves which slot and the order paths are numbered change.
// The block barriers between the passes are what keeps this kernel fast: a barrier-free version (one warp per 64 slots,
// path ids drawn 1024 at a time per warp) ran 2.4x slower - with the warps of an SM spread over 4.7 k instructions of
// shading code the instruction cache misses all the time (ncu: 20 warps stalled on no_instruction per issue against 2.5
// here, profiles/r2_extend2_ab.md).  The same effect sank the fused kernel (srt_path.cuh).
template <bool COUNT, int TS>
__global__ void __launch_bounds__(TS, SRT_SHADE_MIN_BLOCKS * 256 / TS) k_wf_shade2(Scene sc, RenderArgs A, WfPool pool, WfCursors* cur, float* __restrict__ accum,
                                                   unsigned long long* counters) {
  constexpr int CH = 2 * TS;      // TS = 256 (8 warps meet at the barriers) or 128 (4 warps, twice the CTAs)
  // lists and their counts are double buffered by chunk parity: three barriers per chunk instead of five
  __shared__ unsigned short s_hit[2][CH], s_free[2][CH];
  __shared__ unsigned s_nhit[2], s_nfree[2];
  __shared__ unsigned long long s_base;
  __shared__ unsigned long long s_stat[3];
  const int lane = threadIdx.x & 31;
  const unsigned FULL = 0xffffffffu, below = (1u << lane) - 1;
  if (blockIdx.x == 0 && threadIdx.x == 0) cur->ray = 0;    // the next extend wave starts at slot 0
  if (threadIdx.x < 3) s_stat[threadIdx.x] = 0;
  if (threadIdx.x < 2) { s_nhit[threadIdx.x] = 0; s_nfree[threadIdx.x] = 0; }
  const unsigned long long per_sample = (unsigned long long)A.n_tiles * 32ull;
  const unsigned long long total = per_sample * (unsigned long long)A.sample_count;
  Counters cnt = {0, 0, 0, 0, 0};
  unsigned n_new = 0, n_seg = 0;
  __syncthreads();

  int buf = 0;
  for (int chunk = blockIdx.x; chunk * CH < pool.n_slots; chunk += gridDim.x, buf ^= 1) {
    const int base = chunk * CH;
#if SRT_SHADE_PREFETCH
    {   // the block's next chunk: pull its 5 x 512 records towards the L2 while this chunk is served (they stream from HBM)
      const long long nb = (long long)(chunk + (int)gridDim.x) * CH;
      if (nb + CH <= (long long)pool.n_slots) {
        // 128-byte lines of the chunk: three 16-byte arrays x CH/8 lines, two 8-byte arrays x CH/16 lines
        constexpr int L16 = CH / 8, L8 = CH / 16;
        for (int t = threadIdx.x; t < 3 * L16 + 2 * L8; t += TS) {
          const void* p;
          if (t < 3 * L16) { const int a = t / L16, line = t % L16; p = (a == 0 ? pool.thr : a == 1 ? pool.ray_o : pool.ray_d) + nb + line * 8; }
          else { const int u = t - 3 * L16, line = u % L8; p = u < L8 ? (const void*)(pool.hit + nb + line * 16) : (const void*)(pool.id + nb + line * 16); }
          asm volatile("prefetch.global.L2 [%0];" :: "l"(p));
        }
      }
    }
#endif
    // ---- pass 0: classify (two slots per thread, all loads first)
    float4 t4[2]; int2 h[2]; uint32_t pix[2];
#pragma unroll
    for (int q = 0; q < 2; q++) {
      const int k = base + q * TS + threadIdx.x;
      t4[q] = make_float4(0, 0, 0, __int_as_float(-2)); h[q] = make_int2(-1, 0); pix[q] = 0;
      if (k < pool.n_slots) { t4[q] = pool.thr[k]; h[q] = pool.hit[k]; pix[q] = pool.id[k].x; }
    }
#pragma unroll
    for (int q = 0; q < 2; q++) {
      const int j = q * TS + threadIdx.x;
      const int bounce = __float_as_int(t4[q].w);
      bool is_hit = false, is_free = false;
      if (bounce == -1) is_free = true;                                    // no path in this slot (-2: beyond the pool)
      else if (bounce >= 0) {
        if (h[q].x < 0) {                                                  // miss -> background, main.cpp:39-40
          float* px = accum + (size_t)pix[q] * 3;
          atomicAdd(px, t4[q].x * A.bg[0]); atomicAdd(px + 1, t4[q].y * A.bg[1]); atomicAdd(px + 2, t4[q].z * A.bg[2]);
          is_free = true;
        } else is_hit = true;
      }
      const unsigned mh = __ballot_sync(FULL, is_hit), mf = __ballot_sync(FULL, is_free);
      unsigned bh = 0, bf = 0;
      if (lane == 0) { if (mh) bh = atomicAdd(&s_nhit[buf], __popc(mh)); if (mf) bf = atomicAdd(&s_nfree[buf], __popc(mf)); }
      bh = __shfl_sync(FULL, bh, 0); bf = __shfl_sync(FULL, bf, 0);
      if (is_hit) s_hit[buf][bh + __popc(mh & below)] = (unsigned short)j;
      if (is_free) s_free[buf][bf + __popc(mf & below)] = (unsigned short)j;
    }
    __syncthreads();
    // ---- pass 1: shade the hits
    const unsigned n_hit = s_nhit[buf];
    for (unsigned t = threadIdx.x; t < n_hit; t += TS) {
      const int j = s_hit[buf][t], k = base + j;
      const float4 th4 = pool.thr[k], o = pool.ray_o[k], d = pool.ray_d[k];
      const int2 hh = pool.hit[k];
      const uint2 id = pool.id[k];
      int bounce = __float_as_int(th4.w);
      Ray r; r.o = mk(o.x, o.y, o.z); r.time = o.w; r.d = mk(d.x, d.y, d.z);
      V3 thr = mk(th4.x, th4.y, th4.z), L = mk(0, 0, 0);
      const bool done = shade_segment<COUNT, true>(sc, A, id.x, id.y, bounce, r, hh.x, __int_as_float(hh.y), L, thr, cnt);
      if (L.x != 0.0f || L.y != 0.0f || L.z != 0.0f) {                      // this segment's emitted term (NaN included)
        float* px = accum + (size_t)id.x * 3;
        atomicAdd(px, L.x); atomicAdd(px + 1, L.y); atomicAdd(px + 2, L.z);
      }
      if (done) s_free[buf][atomicAdd(&s_nfree[buf], 1u)] = (unsigned short)j;
      else {
        pool.ray_o[k] = make_float4(r.o.x, r.o.y, r.o.z, r.time);
        pool.ray_d[k] = make_float4(r.d.x, r.d.y, r.d.z, __int_as_float(1));
        pool.thr[k] = make_float4(thr.x, thr.y, thr.z, __int_as_float(bounce));
        ++n_seg;
      }
    }
    __syncthreads();
    // ---- pass 2: regenerate
    const unsigned n_free = s_nfree[buf];
    if (threadIdx.x == 0) {
      s_base = n_free ? atomicAdd(&cur->path, (unsigned long long)n_free) : total;
      s_nhit[buf ^ 1] = 0; s_nfree[buf ^ 1] = 0;     // the next chunk's counts (its lists were last read a chunk ago)
    }
    __syncthreads();
    const unsigned long long pbase = s_base;
    for (unsigned t = threadIdx.x; t < n_free; t += TS) {
      const int k = base + s_free[buf][t];
      const unsigned long long mine = pbase + t;
      bool made = false;
      if (mine < total) {
        const unsigned q = (unsigned)(mine % per_sample);
        const int s_off = (int)(mine / per_sample);
        const int tile = (int)(q >> 5), l = (int)(q & 31);
        const int x = (tile % A.tiles_x) * 8 + (l & 7), y = (tile / A.tiles_x) * 4 + (l >> 3);
        if (x < A.W && y < A.H) {
          const uint32_t pixel = (uint32_t)(y * A.W + x), sample = (uint32_t)(A.sample_begin + s_off);
          const Ray r = generate_ray<true>(A, x, y, pixel, sample);
          pool.ray_o[k] = make_float4(r.o.x, r.o.y, r.o.z, r.time);
          pool.ray_d[k] = make_float4(r.d.x, r.d.y, r.d.z, __int_as_float(1));
          pool.thr[k] = make_float4(1.0f, 1.0f, 1.0f, __int_as_float(0));
          pool.id[k] = make_uint2(pixel, sample);
          made = true; ++n_new; ++n_seg;
        }
      }
      if (!made) {                                  // no path: extend skips the slot, the next shade pass retries
        pool.ray_d[k] = make_float4(0, 0, 0, __int_as_float(0));
        pool.thr[k] = make_float4(0, 0, 0, __int_as_float(-1));
      }
    }
    // no barrier here: the next chunk fills the other pair of lists, and s_base is not written again before every thread
    // has passed that chunk's second barrier
  }
  unsigned long long v0 = n_new, v1 = n_seg, v2 = cnt.tri_hits;
  for (int o = 16; o > 0; o >>= 1) {
    v0 += __shfl_down_sync(FULL, v0, o); v1 += __shfl_down_sync(FULL, v1, o);
    if (COUNT) v2 += __shfl_down_sync(FULL, v2, o);
  }
  if (lane == 0) { atomicAdd(&s_stat[0], v0); atomicAdd(&s_stat[1], v1); if (COUNT) atomicAdd(&s_stat[2], v2); }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (s_stat[0]) atomicAdd(&counters[0], s_stat[0]);
    if (s_stat[1]) { atomicAdd(&counters[1], s_stat[1]); cur->alive = 1; }
    if (COUNT && s_stat[2]) atomicAdd(&counters[7], s_stat[2]);
  }
}

// ---- SHADE + REGENERATE, sorted phases with carried remainders --------------------------------------------------------
// In k_wf_shade2 a 512-slot chunk yields ~280 hits and ~290 free slots for 256 threads: the second round of passes 1 and 2
// runs with ONE warp while seven wait at the barrier (ncu: 5.3 warps per issue stalled on the barrier, the largest stall).
// Here the lists hold pool-wide slot numbers and outlive the chunk: a pass only serves whole rounds of 256 entries and hands
// the remainder (< 256) to the next chunk's lists; the block's last chunk serves everything.  A slot classified in chunk c
// may thus be shaded or regenerated while chunk c + 1 is classified - nothing else reads it before the next extend launch.
template <bool COUNT>
__global__ void __launch_bounds__(256, SRT_SHADE_MIN_BLOCKS) k_wf_shade2c(Scene sc, RenderArgs A, WfPool pool, WfCursors* cur, float* __restrict__ accum,
                                                    unsigned long long* counters) {
  constexpr int TS = 256, CH = 512;
  constexpr int CAP_H = CH + TS;                // carried (< TS) + this chunk's hits (<= CH)
  constexpr int CAP_F = 2 * CH + 2 * TS;        // carried + this chunk's free slots + the paths pass 1 ended (<= CAP_H)
  __shared__ unsigned s_hit[2][CAP_H], s_free[2][CAP_F];
  __shared__ unsigned s_nhit[2], s_nfree[2];
  __shared__ unsigned long long s_base;
  __shared__ unsigned long long s_stat[3];
  const int lane = threadIdx.x & 31;
  const unsigned FULL = 0xffffffffu, below = (1u << lane) - 1;
  if (blockIdx.x == 0 && threadIdx.x == 0) cur->ray = 0;    // the next extend wave starts at slot 0
  if (threadIdx.x < 3) s_stat[threadIdx.x] = 0;
  if (threadIdx.x < 2) { s_nhit[threadIdx.x] = 0; s_nfree[threadIdx.x] = 0; }
  const unsigned long long per_sample = (unsigned long long)A.n_tiles * 32ull;
  const unsigned long long total = per_sample * (unsigned long long)A.sample_count;
  Counters cnt = {0, 0, 0, 0, 0};
  unsigned n_new = 0, n_seg = 0;
  __syncthreads();

  int buf = 0;
  for (long long chunk = blockIdx.x; chunk * CH < (long long)pool.n_slots; chunk += gridDim.x, buf ^= 1) {
    const long long base = chunk * CH;
    const bool last = (chunk + (long long)gridDim.x) * CH >= (long long)pool.n_slots;
#if SRT_SHADE_PREFETCH
    {   // the block's next chunk: pull its records towards the L2 while this chunk is served (they stream from HBM)
      const long long nb = (chunk + (long long)gridDim.x) * CH;
      if (nb + CH <= (long long)pool.n_slots) {
        constexpr int L16 = CH / 8, L8 = CH / 16;      // 128-byte lines: three 16-byte arrays, two 8-byte arrays
        for (int t = threadIdx.x; t < 3 * L16 + 2 * L8; t += TS) {
          const void* p;
          if (t < 3 * L16) { const int a = t / L16, line = t % L16; p = (a == 0 ? pool.thr : a == 1 ? pool.ray_o : pool.ray_d) + nb + line * 8; }
          else { const int u = t - 3 * L16, line = u % L8; p = u < L8 ? (const void*)(pool.hit + nb + line * 16) : (const void*)(pool.id + nb + line * 16); }
          asm volatile("prefetch.global.L2 [%0];" :: "l"(p));
        }
      }
    }
#endif
    // ---- pass 0: classify (two slots per thread, all loads first); the lists already hold the carried entries
    float4 t4[2]; int2 h[2]; uint32_t pix[2];
#pragma unroll
    for (int q = 0; q < 2; q++) {
      const long long k = base + q * TS + threadIdx.x;
      t4[q] = make_float4(0, 0, 0, __int_as_float(-2)); h[q] = make_int2(-1, 0); pix[q] = 0;
      if (k < (long long)pool.n_slots) { t4[q] = pool.thr[k]; h[q] = pool.hit[k]; pix[q] = pool.id[k].x; }
    }
#pragma unroll
    for (int q = 0; q < 2; q++) {
      const unsigned k = (unsigned)(base + q * TS + threadIdx.x);
      const int bounce = __float_as_int(t4[q].w);
      bool is_hit = false, is_free = false;
      if (bounce == -1) is_free = true;                                    // no path in this slot (-2: beyond the pool)
      else if (bounce >= 0) {
        if (h[q].x < 0) {                                                  // miss -> background, main.cpp:39-40
          float* px = accum + (size_t)pix[q] * 3;
          atomicAdd(px, t4[q].x * A.bg[0]); atomicAdd(px + 1, t4[q].y * A.bg[1]); atomicAdd(px + 2, t4[q].z * A.bg[2]);
          is_free = true;
        } else is_hit = true;
      }
      const unsigned mh = __ballot_sync(FULL, is_hit), mf = __ballot_sync(FULL, is_free);
      unsigned bh = 0, bf = 0;
      if (lane == 0) { if (mh) bh = atomicAdd(&s_nhit[buf], __popc(mh)); if (mf) bf = atomicAdd(&s_nfree[buf], __popc(mf)); }
      bh = __shfl_sync(FULL, bh, 0); bf = __shfl_sync(FULL, bf, 0);
      if (is_hit) s_hit[buf][bh + __popc(mh & below)] = k;
      if (is_free) s_free[buf][bf + __popc(mf & below)] = k;
    }
    __syncthreads();
    // ---- pass 1: shade whole rounds of hits; the remainder moves to the next chunk's list
    const unsigned n_hit = s_nhit[buf];
    const unsigned do_hit = last ? n_hit : n_hit / TS * TS;
    for (unsigned t = threadIdx.x; t < do_hit; t += TS) {
      const unsigned k = s_hit[buf][t];
      const float4 th4 = pool.thr[k], o = pool.ray_o[k], d = pool.ray_d[k];
      const int2 hh = pool.hit[k];
      const uint2 id = pool.id[k];
      int bounce = __float_as_int(th4.w);
      Ray r; r.o = mk(o.x, o.y, o.z); r.time = o.w; r.d = mk(d.x, d.y, d.z);
      V3 thr = mk(th4.x, th4.y, th4.z), L = mk(0, 0, 0);
      const bool done = shade_segment<COUNT, true>(sc, A, id.x, id.y, bounce, r, hh.x, __int_as_float(hh.y), L, thr, cnt);
      if (L.x != 0.0f || L.y != 0.0f || L.z != 0.0f) {                      // this segment's emitted term (NaN included)
        float* px = accum + (size_t)id.x * 3;
        atomicAdd(px, L.x); atomicAdd(px + 1, L.y); atomicAdd(px + 2, L.z);
      }
      if (done) s_free[buf][atomicAdd(&s_nfree[buf], 1u)] = k;
      else {
        pool.ray_o[k] = make_float4(r.o.x, r.o.y, r.o.z, r.time);
        pool.ray_d[k] = make_float4(r.d.x, r.d.y, r.d.z, __int_as_float(1));
        pool.thr[k] = make_float4(thr.x, thr.y, thr.z, __int_as_float(bounce));
        ++n_seg;
      }
    }
    if (threadIdx.x < n_hit - do_hit) s_hit[buf ^ 1][threadIdx.x] = s_hit[buf][do_hit + threadIdx.x];   // (< TS entries)
    __syncthreads();
    // ---- pass 2: regenerate whole rounds of free slots
    const unsigned n_free = s_nfree[buf];
    const unsigned do_free = last ? n_free : n_free / TS * TS;
    if (threadIdx.x == 0) {
      s_base = do_free ? atomicAdd(&cur->path, (unsigned long long)do_free) : total;
      s_nhit[buf ^ 1] = n_hit - do_hit; s_nfree[buf ^ 1] = n_free - do_free;     // the next chunk's lists start with the carry
    }
    if (threadIdx.x < n_free - do_free) s_free[buf ^ 1][threadIdx.x] = s_free[buf][do_free + threadIdx.x];
    __syncthreads();
    const unsigned long long pbase = s_base;
    for (unsigned t = threadIdx.x; t < do_free; t += TS) {
      const unsigned k = s_free[buf][t];
      const unsigned long long mine = pbase + t;
      bool made = false;
      if (mine < total) {
        const unsigned q = (unsigned)(mine % per_sample);
        const int s_off = (int)(mine / per_sample);
        const int tile = (int)(q >> 5), l = (int)(q & 31);
        const int x = (tile % A.tiles_x) * 8 + (l & 7), y = (tile / A.tiles_x) * 4 + (l >> 3);
        if (x < A.W && y < A.H) {
          const uint32_t pixel = (uint32_t)(y * A.W + x), sample = (uint32_t)(A.sample_begin + s_off);
          const Ray r = generate_ray<true>(A, x, y, pixel, sample);
          pool.ray_o[k] = make_float4(r.o.x, r.o.y, r.o.z, r.time);
          pool.ray_d[k] = make_float4(r.d.x, r.d.y, r.d.z, __int_as_float(1));
          pool.thr[k] = make_float4(1.0f, 1.0f, 1.0f, __int_as_float(0));
          pool.id[k] = make_uint2(pixel, sample);
          made = true; ++n_new; ++n_seg;
        }
      }
      if (!made) {                                  // no path: extend skips the slot, the next shade pass retries
        pool.ray_d[k] = make_float4(0, 0, 0, __int_as_float(0));
        pool.thr[k] = make_float4(0, 0, 0, __int_as_float(-1));
      }
    }
    // no barrier here: the next chunk appends to the other pair of lists (whose carried part was written before the last
    // barrier), and s_base is not written again before every thread has passed that chunk's second barrier
  }
  unsigned long long v0 = n_new, v1 = n_seg, v2 = cnt.tri_hits;
  for (int o = 16; o > 0; o >>= 1) {
    v0 += __shfl_down_sync(FULL, v0, o); v1 += __shfl_down_sync(FULL, v1, o);
    if (COUNT) v2 += __shfl_down_sync(FULL, v2, o);
  }
  if (lane == 0) { atomicAdd(&s_stat[0], v0); atomicAdd(&s_stat[1], v1); if (COUNT) atomicAdd(&s_stat[2], v2); }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (s_stat[0]) atomicAdd(&counters[0], s_stat[0]);
    if (s_stat[1]) { atomicAdd(&counters[1], s_stat[1]); cur->alive = 1; }
    if (COUNT && s_stat[2]) atomicAdd(&counters[7], s_stat[2]);
  }
}

// ---- ray ordering for trees that live in HBM -------------------------------------------------------------------------
// After a few bounces the rays of neighbouring slots have nothing in common, and every lane drags its own part of a 1 GB
// scene through the L2.  Sorting the wave's rays by where they start (Morton code of the origin's cell in a 256^3 grid over
// the root box) and roughly where they go (direction octant) makes the rays traced at the same time - one warp, and all the
// CTAs pulling from the same cursor - walk the same neighbourhood of the tree: node and primitive sectors are fetched from
// HBM once and then hit in L2 (config 5: 272 -> 163 GB of DRAM reads per launch, L2 hit rate 55 -> 67 %).  Handing each CTA
// its own run of 1 Ki - 16 Ki sorted rays instead (better L1 locality, 148 neighbourhoods in the L2) measured slower.  Slots without a ray sort to the end.  The sort itself is cub::DeviceRadixSort (a few
// hundred microseconds for 8 Mi pairs against a 160 ms extend launch); results do not depend on the order rays are traced in.
struct WfSortGrid { float org[3], inv_cell[3]; int mode; };
__device__ __forceinline__ unsigned spread3(unsigned x) {          // 8 bits -> every third bit
  x &= 0xffu;
  x = (x | (x << 8)) & 0x0000f00fu;
  x = (x | (x << 4)) & 0x000c30c3u;
  x = (x | (x << 2)) & 0x00249249u;
  return x;
}
__global__ void k_wf_keys(WfPool pool, WfSortGrid g, unsigned* __restrict__ keys, unsigned* __restrict__ vals) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= pool.n_slots) return;
  const float4 d = pool.ray_d[k];
  unsigned key = 0x08000000u;                                      // no ray: after every real key
  if (__float_as_int(d.w) == 1) {
    const float4 o = pool.ray_o[k];
    const int cx = min(255, max(0, (int)((o.x - g.org[0]) * g.inv_cell[0])));
    const int cy = min(255, max(0, (int)((o.y - g.org[1]) * g.inv_cell[1])));
    const int cz = min(255, max(0, (int)((o.z - g.org[2]) * g.inv_cell[2])));
    const unsigned oct = (d.x < 0 ? 1u : 0u) | (d.y < 0 ? 2u : 0u) | (d.z < 0 ? 4u : 0u);
    const unsigned m8 = spread3(cx) | (spread3(cy) << 1) | (spread3(cz) << 2);
    key = g.mode == 1 ? ((m8 << 3) | oct) : ((oct << 24) | m8);   // octant first measured 1 % ahead of cell first (round-1 sweep, profiles/r1_c5_extend_summary.md)
  }
  keys[k] = key;
  vals[k] = (unsigned)k;
}

// marks every slot empty (before the first shade pass of a frame)
__global__ void k_wf_reset(WfPool pool) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= pool.n_slots) return;
  pool.thr[k] = make_float4(0, 0, 0, __int_as_float(-1));
  pool.ray_d[k] = make_float4(0, 0, 0, __int_as_float(0));
}

}  // namespace srt
