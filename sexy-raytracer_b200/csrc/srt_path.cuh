// srt_path.cuh — k_wf_path: the wavefront's EXTEND and SHADE steps in ONE persistent launch per frame (tree in shared memory).
//
// k_wf_extend2 already lets a warp test leaves only when 32 of them are waiting; this kernel treats shading the same way.
// A ray whose walk is over is put on the warp's FINISHED list; when 32 are waiting the whole warp runs main.cpp:39-51 on
// them (shade_segment, srt_kernels.cuh) and regenerates the paths that ended (main.cpp:204-217, generate_ray) in place.
// The new rays go on the warp's READY list and are walked by the next free lanes.  Consequences:
//   * no waves: the two kernels no longer alternate (C4: 994 x (1.4 + 0.5 ms)), and no wave has a tail;
//   * the shade step's memory latency (0.5 ms per wave on its own) is hidden behind the 31 other warps of the SM, which
//     keep walking - what it still costs is its instructions on an issue-bound SM;
//   * path state lives in warp-private slots (PATH_SLOTS per warp, 64 B each: 48 MB for 148 x 32 warps) that are written
//     and read again by the same warp within microseconds - they stay in the L2 instead of streaming through HBM.
// Results: the same paths (Philox keys are (pixel, global sample, bounce)), the same per-ray traversal as k_wf_extend2
// (same theorem, same arithmetic at the leaves); only the order in which a pixel's samples are added changes.
//
// Per-warp shared memory (after the staged nodes), PATH_WARP_BYTES:
//   uint2 sph[64]          rays waiting on a leaf that holds a sphere   (slot << 24 | leaf node, tCur)
//   uint2 tri[PATH_SLOTS]  rays waiting on a leaf of triangles, from index 0 up; from the top down the READY rays
//                          (slot << 24 | next state, tCur) - a ray is in one list at a time, so the two never meet
//   u8    fin[192]         slots whose walk is over (the hit is in pool.hit[slot])
//   u8    free[PATH_SLOTS] slots without a path
// The primitive of the last accepted leaf is not carried with the ray: it is stored to pool.hit[slot] when a leaf accepts.
#pragma once
#include "srt_extend2.cuh"

namespace srt {

#define PATH_SLOTS 160
#define PATH_FIN 192
#define PATH_WARP_BYTES ((64 + PATH_SLOTS) * 8 + PATH_FIN + PATH_SLOTS)      /* 2144, a multiple of 16 */

struct PathPool {
  float4* ray_o;      // (o.xyz, time)
  float4* ray_d;      // (d.xyz, -)
  float4* thr;        // (throughput.xyz, bounce as int bits)
  uint2*  id;         // (pixel, sample)
  int2*   hit;        // (prim or -1, t bits) of the last accepted leaf
};

template <bool COUNT, bool SPH>
__device__ __forceinline__ void path_leaf_phase(const Scene& sc, const PathPool& pool, size_t gbase, uint2* list, int& n_list, uint2* s_top, int& r,
                                                unsigned char* s_fin, int& pf, int lane, int n, int step_scale, Counters& cnt) {
  const unsigned FULL = 0xffffffffu, below = (1u << lane) - 1u;
  const int take = n_list < 32 ? n_list : 32;
  bool push = false, finish = false;
  uint2 res = make_uint2(0, 0);
  unsigned sl = 0;
  if (lane < take) {
    const uint2 e = list[n_list - take + lane];
    sl = e.x >> 24;
    const int node = (int)(e.x & 0xffffffu);
    const size_t g = gbase + sl;
    const float4 d = pool.ray_d[g], o = pool.ray_o[g];
    Ray ry; ry.o = mk(o.x, o.y, o.z); ry.time = o.w; ry.d = mk(d.x, d.y, d.z);
    float tc = __uint_as_float(e.y);
    int pr = -1;
    x2_leaf<COUNT, SPH>(sc, node, ry, tc, pr, cnt);
    if (pr >= 0) pool.hit[g] = make_int2(pr, __float_as_int(tc));       // an accept: (prim, tCur) of the ray from now on
    const int nxt = node + 1;
    if (nxt >= n) finish = true;                                        // the leaf was the last node: the walk is over
    else { push = true; res = make_uint2((sl << 24) | (unsigned)(nxt * step_scale), __float_as_uint(tc)); }
  }
  n_list -= take;
  const unsigned mp = __ballot_sync(FULL, push), mf = __ballot_sync(FULL, finish);   // (orders the reads before the writes)
  if (push) s_top[-1 - r - __popc(mp & below)] = res;
  if (finish) s_fin[pf + __popc(mf & below)] = (unsigned char)sl;
  r += __popc(mp); pf += __popc(mf);
  __syncwarp();
}

// Shade + regenerate the top min(pf, 32) entries of the finished list; the top `n_inj` of them are slots without a path
// (taken from the free list because lanes were idle): they only regenerate.
struct PathIds { unsigned long long next, end; bool no_more; };
template <bool COUNT>
__device__ __forceinline__ void path_shade_phase(const Scene& sc, const RenderArgs& A, const PathPool& pool, size_t gbase, WfCursors* cur,
                                                 float* __restrict__ accum, unsigned char* s_fin, int& pf, int n_inj, unsigned char* s_free, int& n_free,
                                                 uint2* s_top, int& r, PathIds& ids, int chunk, unsigned long long total, unsigned long long per_sample,
                                                 int lane, unsigned& n_new, unsigned& n_seg, Counters& cnt) {
  const unsigned FULL = 0xffffffffu, below = (1u << lane) - 1u;
  const float INF = __int_as_float(0x7f800000);
  const int take = pf < 32 ? pf : 32;
  bool live = false, ended = false;
  unsigned sl = 0;
  size_t g = 0;
  if (lane < take) {
    sl = s_fin[pf - take + lane];
    g = gbase + sl;
    ended = lane >= take - n_inj;
    if (!ended) {
      const int2 h = pool.hit[g];
      const float4 t4 = pool.thr[g];
      const uint2 id = pool.id[g];
      float* px = accum + (size_t)id.x * 3;
      if (h.x < 0) {                                                    // miss -> background, main.cpp:39-40
        atomicAdd(px, t4.x * A.bg[0]); atomicAdd(px + 1, t4.y * A.bg[1]); atomicAdd(px + 2, t4.z * A.bg[2]);
        ended = true;
      } else {
        const float4 o = pool.ray_o[g], d = pool.ray_d[g];
        int bounce = __float_as_int(t4.w);
        Ray ry; ry.o = mk(o.x, o.y, o.z); ry.time = o.w; ry.d = mk(d.x, d.y, d.z);
        V3 thr = mk(t4.x, t4.y, t4.z), L = mk(0, 0, 0);
        const bool done = shade_segment<COUNT, true>(sc, A, id.x, id.y, bounce, ry, h.x, __int_as_float(h.y), L, thr, cnt);
        if (L.x != 0.0f || L.y != 0.0f || L.z != 0.0f) { atomicAdd(px, L.x); atomicAdd(px + 1, L.y); atomicAdd(px + 2, L.z); }
        if (done) ended = true;
        else {
          pool.ray_o[g] = make_float4(ry.o.x, ry.o.y, ry.o.z, ry.time);
          pool.ray_d[g] = make_float4(ry.d.x, ry.d.y, ry.d.z, 0.0f);
          pool.thr[g] = make_float4(thr.x, thr.y, thr.z, __int_as_float(bounce));
          live = true;
        }
      }
    }
  }
  pf -= take;
  // ---- regenerate: the ended paths take the next path ids of the warp's chunk (one global atomic per `chunk` paths)
  const unsigned m_end = __ballot_sync(FULL, ended);
  if (m_end) {
    const int n_end = __popc(m_end), my = __popc(m_end & below);
    long long myid = -1;
    int served = 0;
    while (served < n_end) {
      if (ids.next >= ids.end) {
        if (ids.no_more) break;
        unsigned long long b = 0;
        if (lane == 0) b = atomicAdd(&cur->path, (unsigned long long)chunk);
        b = __shfl_sync(FULL, b, 0);
        if (b >= total) { ids.no_more = true; break; }
        ids.next = b; ids.end = b + (unsigned long long)chunk < total ? b + (unsigned long long)chunk : total;
      }
      const unsigned long long left = ids.end - ids.next;
      const int avail = left < (unsigned long long)(n_end - served) ? (int)left : n_end - served;
      if (ended && my >= served && my < served + avail) myid = (long long)(ids.next + (unsigned long long)(my - served));
      ids.next += (unsigned long long)avail; served += avail;
    }
    if (ended && myid >= 0) {
      const unsigned q = (unsigned)((unsigned long long)myid % per_sample);
      const int s_off = (int)((unsigned long long)myid / per_sample);
      const int tile = (int)(q >> 5), l = (int)(q & 31);
      const int x = (tile % A.tiles_x) * 8 + (l & 7), y = (tile / A.tiles_x) * 4 + (l >> 3);
      if (x < A.W && y < A.H) {
        const uint32_t pixel = (uint32_t)(y * A.W + x), sample = (uint32_t)(A.sample_begin + s_off);
        const Ray ry = generate_ray<true>(A, x, y, pixel, sample);
        pool.ray_o[g] = make_float4(ry.o.x, ry.o.y, ry.o.z, ry.time);
        pool.ray_d[g] = make_float4(ry.d.x, ry.d.y, ry.d.z, 0.0f);
        pool.thr[g] = make_float4(1.0f, 1.0f, 1.0f, __int_as_float(0));
        pool.id[g] = make_uint2(pixel, sample);
        live = true; ++n_new;
      }
    }
  }
  if (live) { pool.hit[g] = make_int2(-1, __float_as_int(INF)); ++n_seg; }
  // slots left without a path go back to the free list; the others are ready to be walked from the root
  const bool freed = lane < take && !live;
  const unsigned mfree = __ballot_sync(FULL, freed), mp = __ballot_sync(FULL, live);
  if (freed) s_free[n_free + __popc(mfree & below)] = (unsigned char)sl;
  if (live) s_top[-1 - r - __popc(mp & below)] = make_uint2(sl << 24, __float_as_uint(INF));   // state 0 = the root
  n_free += __popc(mfree); r += __popc(mp);
  __syncwarp();
}

template <bool COUNT, int FMT, int THREADS>
__global__ void __launch_bounds__(THREADS, 1) k_wf_path(Scene sc, RenderArgs A, PathPool pool, WfCursors* cur, float* __restrict__ accum,
                                                       int th_walk, int chunk, unsigned long long* counters) {
  extern __shared__ uint4 smem_x2[];
  const int n = sc.x2_n;
  for (int k = threadIdx.x; k < n; k += THREADS) x2_encode<FMT>(sc, k, smem_x2);
  const int node_u4 = FMT == 32 ? 2 * n : n;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned char* wbase = reinterpret_cast<unsigned char*>(smem_x2 + node_u4) + (size_t)warp * PATH_WARP_BYTES;
  uint2* s_sph = reinterpret_cast<uint2*>(wbase);
  uint2* s_tri = s_sph + 64;
  uint2* s_top = s_tri + PATH_SLOTS;                 // ready entries: s_top[-1 - k]
  unsigned char* s_fin = reinterpret_cast<unsigned char*>(s_top);
  unsigned char* s_free = s_fin + PATH_FIN;
  for (int k = lane; k < PATH_SLOTS; k += 32) s_free[k] = (unsigned char)(PATH_SLOTS - 1 - k);
  const size_t gbase = ((size_t)blockIdx.x * (THREADS / 32) + (size_t)warp) * PATH_SLOTS;
  const uint32_t nodes_s = (uint32_t)__cvta_generic_to_shared(smem_x2);
  const uint32_t hi_off = (uint32_t)n * 16u;
  __syncthreads();
  const unsigned FULL = 0xffffffffu, below = (1u << lane) - 1u;
  const float INF = __int_as_float(0x7f800000);
  const int step_scale = FMT == 32 ? 16 : 1;
  constexpr int KSH = X2_KIND_SHIFT(FMT);
  const unsigned long long per_sample = (unsigned long long)A.n_tiles * 32ull;
  const unsigned long long total = per_sample * (unsigned long long)A.sample_count;
  Counters cnt = {0, 0, 0, 0, 0};
  unsigned n_new = 0, n_seg = 0;

  Aux2 x = {0, 0, 0, 0, 0, 0, 0};
  int i = X2_DONE, slot = -1;
  float tCur = INF;
  int pt = 0, ps = 0, r = 0, pf = 0, n_free = PATH_SLOTS;     // warp-uniform
  PathIds ids = {0, 0, n <= 0};

  for (;;) {
    // ---- A. lanes whose walk ended hand their slot to the finished list; lanes stopped on a leaf park their ray
    {
      const bool has = slot >= 0;
      const bool stp = has && i <= -2, fin = has && i == X2_DONE;
      const int code = -2 - i;
      const bool stp_s = stp && (code >> KSH) != 0;
      const unsigned ms = __ballot_sync(FULL, stp), mss = __ballot_sync(FULL, stp_s), mf = __ballot_sync(FULL, fin);
      if (stp) {
        const uint2 e = make_uint2(((unsigned)slot << 24) | (unsigned)(code & ((1 << KSH) - 1)), __float_as_uint(tCur));
        if (stp_s) s_sph[ps + __popc(mss & below)] = e;
        else s_tri[pt + __popc(ms & ~mss & below)] = e;
      }
      if (fin) s_fin[pf + __popc(mf & below)] = (unsigned char)slot;
      if (stp || fin) { slot = -1; i = X2_DONE; }
      ps += __popc(mss); pt += __popc(ms & ~mss); pf += __popc(mf);
      __syncwarp();
    }
    // ---- B. phases.  Round 0: every list with 32 waiting rays.  Round 1 only if idle lanes remain that the ready list
    // cannot fill: new paths while the frame has some (free slots join the finished list as entries that only regenerate),
    // else whatever is waiting, with fewer than 32 lanes.  (One call site per phase: the shade code is ~4 k instructions.)
#pragma unroll 1
    for (int round = 0; round < 2; round++) {
      bool do_t = pt >= 32, do_s = ps >= 32, do_f = pf >= 32;
      int n_inj = 0;
      if (round == 1) {
        const int n_empty = __popc(__ballot_sync(FULL, slot < 0));
        if (r >= n_empty) break;
        if (!ids.no_more && n_free > 0) {
          n_inj = n_empty - r < n_free ? n_empty - r : n_free;
          if (lane < n_inj) s_fin[pf + lane] = s_free[n_free - 1 - lane];
          pf += n_inj; n_free -= n_inj;
          __syncwarp();
          do_t = do_s = false; do_f = true;
        } else {
          // the end of the frame (or of the warp's slots): as in k_wf_extend2, let the walkers run until all have stopped
          // before serving a few rays per phase
          if (n_empty < 32) break;
          do_t = pt > 0; do_s = ps > 0; do_f = pf > 0;
        }
      }
      if (do_t) path_leaf_phase<COUNT, false>(sc, pool, gbase, s_tri, pt, s_top, r, s_fin, pf, lane, n, step_scale, cnt);
      if (do_s) path_leaf_phase<COUNT, true>(sc, pool, gbase, s_sph, ps, s_top, r, s_fin, pf, lane, n, step_scale, cnt);
      if (do_f) path_shade_phase<COUNT>(sc, A, pool, gbase, cur, accum, s_fin, pf, n_inj, s_free, n_free, s_top, r, ids, chunk, total, per_sample,
                                        lane, n_new, n_seg, cnt);
    }
    // ---- C. idle lanes take a ready ray
    const unsigned m_empty = __ballot_sync(FULL, slot < 0);
    if (m_empty && r > 0) {
      const int n_empty = __popc(m_empty), q = __popc(m_empty & below);
      const int take = r < n_empty ? r : n_empty;
      if (slot < 0 && q < take) {
        const uint2 e = s_top[-1 - (r - take + q)];
        const int sl = (int)(e.x >> 24), link = (int)(e.x & 0xffffffu);
        const size_t g = gbase + (size_t)sl;
        const float4 d = pool.ray_d[g], o = pool.ray_o[g];
        Ray ry; ry.o = mk(o.x, o.y, o.z); ry.time = o.w; ry.d = mk(d.x, d.y, d.z);
        tCur = __uint_as_float(e.y);
        slot = sl;
        if (make_aux2(ry, sc.x2_centre, x)) i = link;
        else {                                       // only a fresh ray (link = root) can land here: the exact ordered walk
          int pr; float t;
          traverse<COUNT>(sc, sc.nodes, ry, 0.001f, INF, pr, t, &cnt);
          pool.hit[g] = make_int2(pr, __float_as_int(t));
          i = X2_DONE;
        }
      }
      r -= take;
    }
    __syncwarp();
    // ---- D. walk while enough lanes are walking
    const unsigned mW = __ballot_sync(FULL, i >= 0);
    if (!mW) {
      if (__ballot_sync(FULL, slot >= 0)) continue;          // finished lanes only: serve them
      if (pt == 0 && ps == 0 && pf == 0 && r == 0 && (ids.no_more || n_free == 0)) break;
      continue;
    }
    const bool drain = r == 0 && (ids.no_more || n_free == 0);
    const int th = drain ? 1 : (__popc(mW) < th_walk ? __popc(mW) : th_walk);
    do {
#pragma unroll
      for (int rep = 0; rep < SRT_X2_REPS; rep++)
        if (i >= 0) walk2_step<COUNT, FMT>(nodes_s, hi_off, x, tCur, i, cnt);
    } while (__popc(__ballot_sync(FULL, i >= 0)) >= th);
  }

  unsigned long long v0 = n_new, v1 = n_seg, v2 = cnt.tri_hits;
  for (int o = 16; o > 0; o >>= 1) {
    v0 += __shfl_down_sync(FULL, v0, o); v1 += __shfl_down_sync(FULL, v1, o);
    if (COUNT) {
      v2 += __shfl_down_sync(FULL, v2, o);
      cnt.nodes += __shfl_down_sync(FULL, cnt.nodes, o); cnt.leaves += __shfl_down_sync(FULL, cnt.leaves, o);
      cnt.exact += __shfl_down_sync(FULL, cnt.exact, o); cnt.sph_leaves += __shfl_down_sync(FULL, cnt.sph_leaves, o);
    }
  }
  if (lane == 0) {
    if (v0) atomicAdd(&counters[0], v0);
    if (v1) atomicAdd(&counters[1], v1);
    if (COUNT) {
      atomicAdd(&counters[2], cnt.nodes); atomicAdd(&counters[3], cnt.leaves); atomicAdd(&counters[4], cnt.exact);
      atomicAdd(&counters[6], cnt.sph_leaves); atomicAdd(&counters[7], v2);
    }
  }
}

}  // namespace srt
