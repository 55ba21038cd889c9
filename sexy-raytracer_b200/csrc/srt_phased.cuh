// srt_phased.cuh — k_render_phased: the persistent megakernel re-scheduled INSIDE the warp.
//
// Why (profiles/r1_v1_megakernel_summary.md): k_render_mega runs one traversal `while` per lane and tests
// leaves inside it, so a warp waits for its longest traversal (mean 54 nodes, long tail) and leaf tests
// / shading run with 2-3 lanes: 10.6 of 32 lanes active on average.  Here every lane is a small state
// machine and the WARP votes on what to execute next, so that each of the three code bodies runs with
// as many lanes as possible:
//     T  traversing          : one box test per step (the hot body, ~35 instructions)
//     P  leaf pending        : box of a leaf node was hit; its one or two primitives wait to be tested
//     F  traversal finished  : the segment's hit is known; waits for shading (+ regeneration)
//     D  dry                 : no samples left in this work unit
// A lane that finishes a path starts the next sample of its pixel in the same shade phase, so lanes
// never idle across segments.  Traversal ORDER per ray is untouched (pre-order, running tCur, SURVEY
// 8.0(1)): the voting only changes when a lane takes its next step, never which step it is — results are
// bit-identical to k_render_mega (tests/test_gpu_parity.py::test_variants_agree).
#pragma once
#include "srt_kernels.cuh"

namespace srt {

// thresholds of the vote (lanes out of 32); RenderArgs carries them so they can be swept without rebuilding
struct PhaseTune { int th_shade, th_leaf, th_trav; };

template <bool SMEM>
__device__ __forceinline__ void load_node(const Node* nodes, int i, float4& lo, float4& hi) {
  if (SMEM) { lo = nodes[i].lo; hi = nodes[i].hi; }
  else { lo = __ldg(&nodes[i].lo); hi = __ldg(&nodes[i].hi); }
}

template <bool COUNT, bool SMEM, int THREADS>
__global__ void __launch_bounds__(THREADS, 1) k_render_phased(Scene sc, RenderArgs A, PhaseTune tune, float* __restrict__ accum,
                                                             unsigned long long* counters) {
  extern __shared__ float4 smem_nodes[];
  const Node* nodes = stage_nodes(sc, SMEM, smem_nodes);
  const int lane = threadIdx.x & 31;
  const unsigned FULL = 0xffffffffu;
  const int n = sc.n_nodes;
  const float INF = __int_as_float(0x7f800000);
  unsigned long long n_paths = 0, n_segments = 0;
  Counters cnt = {0, 0, 0, 0, 0};

  for (;;) {
    long long unit = 0;
    if (lane == 0) unit = (long long)atomicAdd(&counters[5], 1ULL);
    unit = __shfl_sync(FULL, unit, 0);
    if (unit >= A.n_units) break;
    const int tile = (int)(unit % A.n_tiles), chunk = (int)(unit / A.n_tiles);
    const int x = (tile % A.tiles_x) * 8 + (lane & 7), y = (tile / A.tiles_x) * 4 + (lane >> 3);
    const bool valid = x < A.W && y < A.H;
    const uint32_t pixel = (uint32_t)(y * A.W + x);
    int s = A.sample_begin + chunk * A.chunk;
    const int s_end = valid ? min(s + A.chunk, A.sample_begin + A.sample_count) : s;

    V3 sum = mk(0, 0, 0), L = mk(0, 0, 0), thr = mk(1, 1, 1);
    Ray r; r.o = r.d = mk(0, 0, 0); r.time = 0;
    RayAux aux = make_aux(r, 0);
    uint32_t sample = 0;
    int bounce = -1;                 // -1: no path in flight
    int i = n;                       // next node in pre-order; n = traversal finished
    int pend0 = -1, pend1 = -1;      // primitives of a hit leaf node, still to be tested (in this order)
    int prim = -1;
    float tCur = INF;

    for (;;) {
      const bool stT = i < n && pend0 < 0;
      const bool stP = pend0 >= 0;
      const bool stF = !stT && !stP && (bounce >= 0 || s < s_end);
      const unsigned mT = __ballot_sync(FULL, stT), mP = __ballot_sync(FULL, stP), mF = __ballot_sync(FULL, stF);
      if (!(mT | mP | mF)) break;

      if (__popc(mF) >= tune.th_shade || !(mT | mP)) {
        // ---- shade phase: main.cpp:44-51 for finished segments, then regeneration (main.cpp:204-217)
        if (stF) {
          bool done = true;
          if (bounce >= 0) {
            done = shade_segment<COUNT>(sc, A, pixel, sample, bounce, r, prim, tCur, L, thr, cnt);
            if (done) { sum = sum + L; bounce = -1; }
          }
          if (done && s < s_end) {                                  // next sample of this lane's pixel
            sample = (uint32_t)s++;
            r = generate_ray(A, x, y, pixel, sample);
            L = mk(0, 0, 0); thr = mk(1, 1, 1); bounce = 0;
            ++n_paths;
          }
          if (bounce >= 0) {                                        // a segment starts: world.hit, main.cpp:39
            ++n_segments;
            aux = make_aux(r, sc.filter_ok);
            i = 0; tCur = INF; prim = -1;
          }
        }
        continue;
      }

      if (__popc(mP) >= tune.th_leaf || !mT) {
        // ---- leaf phase: triangle::hit / sphere::hit of the pending primitives, in reference order
        if (stP) {
#pragma unroll 1
          for (int k = 0; k < 2; k++) {
            const int p = k == 0 ? pend0 : pend1;
            if (p < 0) break;
            if (COUNT) { ++cnt.leaves; if (p >= sc.n_tri) ++cnt.sph_leaves; }
            float t;
            const bool acc = p < sc.n_tri ? tri_hit(sc.tris, p, r, 0.001f, t) : sph_hit(sc.sphs, p - sc.n_tri, r, 0.001f, tCur, t);
            if (acc) { tCur = t; prim = p; }
          }
          pend0 = pend1 = -1;
        }
        continue;
      }

      // ---- traverse phase: box tests until too few lanes are still walking
      bool walking = stT;
      do {
        if (walking) {
          float4 lo, hi;
          load_node<SMEM>(nodes, i, lo, hi);
          const int w6 = __float_as_int(hi.z), w7 = __float_as_int(hi.w);
          if (COUNT) ++cnt.nodes;
          const bool hit = (w7 == NODE_NO_BOX) ? true : box_hit<COUNT>(lo, hi, r, aux, 0.001f, tCur, COUNT ? &cnt.exact : nullptr);
          if (w7 == NODE_INNER) {
            i = hit ? i + 1 : w6;
          } else {
            i = i + 1;
            if (hit) { pend0 = w6; pend1 = w7; walking = false; }
          }
          if (i >= n) walking = false;
        }
      } while (__popc(__ballot_sync(FULL, walking)) >= tune.th_trav);
    }

    if (valid) {
      float* px = accum + (size_t)pixel * 3;
      if (A.n_chunks == 1) { px[0] += sum.x; px[1] += sum.y; px[2] += sum.z; }      // sole owner of the pixel
      else { atomicAdd(px, sum.x); atomicAdd(px + 1, sum.y); atomicAdd(px + 2, sum.z); }
    }
  }
  // statistics: one atomic per warp
  for (int o = 16; o > 0; o >>= 1) {
    n_paths += __shfl_down_sync(FULL, n_paths, o);
    n_segments += __shfl_down_sync(FULL, n_segments, o);
    if (COUNT) {
      cnt.nodes += __shfl_down_sync(FULL, cnt.nodes, o);
      cnt.leaves += __shfl_down_sync(FULL, cnt.leaves, o);
      cnt.exact += __shfl_down_sync(FULL, cnt.exact, o);
      cnt.sph_leaves += __shfl_down_sync(FULL, cnt.sph_leaves, o);
      cnt.tri_hits += __shfl_down_sync(FULL, cnt.tri_hits, o);
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

}  // namespace srt
