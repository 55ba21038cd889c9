// srt_launch.cuh — compute entry points of the C ABI (included at the end of srt_b200.cu).
#pragma once

namespace {

struct DevBuf {   // scoped device allocation
  void* p = nullptr;
  ~DevBuf() { if (p) cudaFree(p); }
  cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 16); }
  template <typename T> T* as() { return static_cast<T*>(p); }
};

// sampler 1: the table of in-pixel positions of global samples [begin, begin + count) on the device (grow-only buffer)
static srt_status ensure_grid(srt_device* dev, int begin, int count, RenderArgs& A) {
  CK(cudaSetDevice(dev->gpu));
  if ((size_t)count > dev->grid_cap) {
    if (dev->d_grid) CK(cudaFree(dev->d_grid));
    dev->d_grid = nullptr; dev->grid_cap = 0;
    CK(cudaMalloc(&dev->d_grid, (size_t)count * sizeof(float2)));
    dev->grid_cap = (size_t)count;
  }
  std::vector<float2> tab((size_t)count);
  for (int k = 0; k < count; k++) srt_rotated_grid(begin + k, &tab[k].x, &tab[k].y);
  CK(cudaMemcpyAsync(dev->d_grid, tab.data(), (size_t)count * sizeof(float2), cudaMemcpyHostToDevice, dev->stream));
  CK(cudaStreamSynchronize(dev->stream));
  A.sampler = 1; A.grid = dev->d_grid; A.grid_begin = begin;
  return SRT_OK;
}

// Tuning overrides (launch shapes, tree modes) are read from the environment ONLY when SRT_DEBUG_TUNING=1: a stray variable
// must not change a production launch.  They exist for A/B measurements and for tests that push small scenes through the
// large-scene paths; results never depend on them.
//   SRT_WF="slots,threads,th_walk"   SRT_WF_TREE=full|compact   SRT_WF_SORT=0|1   SRT_WF_KEY=0|1   SRT_WF_SHADE=1|2
//   SRT_WF_X2=0|16|32 (k_wf_extend2 off / 16-byte / 32-byte nodes)   SRT_WF_L1=<carve-out %>   SRT_WF_CTAS_PER_SM=<k>
//   SRT_TUNE="threads,th_shade,th_leaf,th_trav" (variant 3)   SRT_FORCE_GLOBAL_TREE=1 (srt_scene_create)
static const char* tune_env(const char* name) {
  const char* on = getenv("SRT_DEBUG_TUNING");
  if (!on || strcmp(on, "1") != 0) return nullptr;
  return getenv(name);
}

struct RenderPlan {
  RenderArgs A; int blocks, threads; size_t smem; int variant; PhaseTune tune; const void* fn;
  int wf_slots, wf_th_walk, wf_shade_chunk, wf_shade_threads; const void* wf_shade;     // variant 2
  int x2_fmt, x2_park;                                                // variant 2 with k_wf_extend2: node bytes (0 = k_wf_extend), entries per warp
  int wf_walk;                                                        // k_wf_extend over a tree in HBM walks Scene::x2_nodes / x2_cnodes
  int fused, path_chunk;                                              // variant 2 as ONE launch of k_wf_path (srt_path.cuh); path ids per atomic
};

typedef void (*path_fn)(Scene, RenderArgs, PathPool, WfCursors*, float*, int, int, unsigned long long*);
template <int THREADS>
path_fn pick_path(bool count, int fmt) {
  if (count) return fmt == 16 ? k_wf_path<true, 16, THREADS> : k_wf_path<true, 32, THREADS>;
  return fmt == 16 ? k_wf_path<false, 16, THREADS> : k_wf_path<false, 32, THREADS>;
}
#ifndef SRT_WF_SLOTS_SMEM
#define SRT_WF_SLOTS_SMEM (1 << 25)
#endif
#ifndef SRT_SHADE_DEFAULT
#define SRT_SHADE_DEFAULT 3       /* 2: k_wf_shade2, 3: k_wf_shade2c (carried remainders; -7 % per wave, B200) */
#endif
#ifndef SRT_SHADE2_THREADS
#define SRT_SHADE2_THREADS 256
#endif
#ifndef SRT_PATH_DEFAULT
#define SRT_PATH_DEFAULT 0        /* 1: variant 2 runs as k_wf_path whenever k_wf_extend2 would be used and the lists fit */
#endif

typedef void (*extend2_fn)(Scene, WfPool, WfCursors*, int, int, unsigned long long*);
template <int THREADS>
extend2_fn pick_extend2(bool count, int fmt) {
  if (count) return fmt == 16 ? k_wf_extend2<true, 16, THREADS> : k_wf_extend2<true, 32, THREADS>;
  return fmt == 16 ? k_wf_extend2<false, 16, THREADS> : k_wf_extend2<false, 32, THREADS>;
}
// shared memory of k_wf_extend2: the staged nodes + `park` 16-byte entries per warp
static size_t x2_smem_bytes(int n_nodes, int fmt, int threads, int park) {
  return (size_t)n_nodes * (size_t)fmt + (size_t)(threads / 32) * (size_t)park * 16;
}

typedef void (*extend_fn)(Scene, WfPool, WfCursors*, int, unsigned long long*);
template <int THREADS>
extend_fn pick_extend(bool count, int mode) {     // mode 0: tree in shared memory, 1: 32-byte nodes in HBM, 2: compact tree in HBM
  if (count) return mode == 0 ? k_wf_extend<true, 0, THREADS> : mode == 1 ? k_wf_extend<true, 1, THREADS> : k_wf_extend<true, 2, THREADS>;
  return mode == 0 ? k_wf_extend<false, 0, THREADS> : mode == 1 ? k_wf_extend<false, 1, THREADS> : k_wf_extend<false, 2, THREADS>;
}

// variant 1: k_render_mega (flat per-lane loop); variant 3: k_render_phased (warp-voted phases); 0 = auto = 3
typedef void (*phased_fn)(Scene, RenderArgs, PhaseTune, float*, unsigned long long*);
template <int THREADS>
phased_fn pick_phased(bool count, bool smem) {
  if (count) return smem ? k_render_phased<true, true, THREADS> : k_render_phased<true, false, THREADS>;
  return smem ? k_render_phased<false, true, THREADS> : k_render_phased<false, false, THREADS>;
}

// variant 2: the launch shape of the EXTEND and SHADE kernels for this scene (`total` = paths of the call)
static cudaError_t plan_wavefront(srt_scene* s, long long total, RenderPlan& plan, int& per_sm) {
  srt_device* dev = s->dev;
  cudaError_t e;
  plan.smem = s->nodes_in_smem ? s->node_bytes : 0;
  // SRT_WF="slots,threads,th_walk" (tuning sweeps only)
  // Pool: 8 Mi slots (537 MB).  When the tree is walked from shared memory a wave of k_wf_extend2 is only ~1 ms per 8 Mi rays, and
  // the fixed cost of a wave (two launches, node staging, the drain at its end) is worth amortising over more slots - as long as
  // the frame still refills the pool several times, because the last refill runs with a pool that empties: slots = paths / 8,
  // between 8 Mi and 32 Mi (2.1 GB).  Measured on B200, M segments/s: config 4 (4.6 G paths) 8 / 12 / 16 / 24 / 32 Mi slots ->
  // 5 051 / 5 289 / 5 413 / 5 558 / 5 630; a 200-spp 720p frame (184 M paths) 8 / 16 / 32 Mi -> 4 812 / 4 975 / 4 842; config 3
  // (102 M paths) 8 / 16 / 32 Mi -> 4 662 / 4 705 / 4 436, and on two GPUs (51 M paths each) 8 / 32 Mi -> 8 611 / 6 860.
  plan.wf_slots = 1 << 23; plan.threads = 1024; plan.wf_th_walk = 16;
  if (s->nodes_in_smem) {
    const long long want = total / 8;
    plan.wf_slots = (int)std::min<long long>(std::max<long long>(want, 1ll << 23), (long long)SRT_WF_SLOTS_SMEM) / 256 * 256;
  }
  // trees beyond the L2 (sorted rays, srt_wavefront.cuh: k_wf_keys): 2 x 768 threads per SM at 40 registers measured +4 %
  // over 1 x 1024 at 56 (61.2 vs 58.8 M segments/s, round-1 sweep recorded in profiles/r1_c5_extend_summary.md); below 1024 the rate falls linearly
  // ... and a pool of 64 Mi slots (5.4 GB + 1 GB of sort buffers): the more rays one sort orders, the smaller the region of
  // the scene the rays in flight cover - 8 / 32 / 64 / 128 Mi slots: 63.7 / 70.3 / 71.3 / 71.1 M segments/s (profiles/r1_c5_extend_summary.md)
  if (!s->nodes_in_smem && s->node_bytes > dev->l2_bytes) { plan.threads = 768; plan.wf_slots = 1 << 26; }
  bool wf_env = false;
  if (const char* t = tune_env("SRT_WF")) {
    int a = 0, b = 0, c = 0;
    if (sscanf(t, "%d,%d,%d", &a, &b, &c) == 3 && a >= 1024 && (b == 256 || b == 512 || b == 768 || b == 1024) && c >= 1) {
      plan.wf_slots = a; plan.threads = b; plan.wf_th_walk = c; wf_env = true;
    }
  }
    if ((long long)plan.wf_slots > total) plan.wf_slots = (int)((total + 255) / 256 * 256);
  // Trees in HBM: the compact tree pays off once the 32-byte nodes no longer fit the L2 (config 5: 378 MB of nodes,
  // 47.6 vs 36.7 M segments/s); an L2-resident tree is faster through its full nodes (1 M triangles: 236 vs 221).
  // SRT_WF_TREE=full|compact overrides (A/B measurements, tests).
  const char* tree = tune_env("SRT_WF_TREE");
  bool compact = s->sc.cnodes && s->node_bytes > dev->l2_bytes;
  if (tree && !strcmp(tree, "full")) compact = false;
  if (tree && !strcmp(tree, "compact")) compact = s->sc.cnodes != nullptr;
  const int mode = s->nodes_in_smem ? 0 : compact ? 2 : 1;
  // Trees in HBM: k_wf_extend's exact (filtered) box test is also a conservative one, so it can walk the scene's walk tree
  // (Scene::x2_nodes: a surface-area sweep over the reference's leaf sequence, srt_scene_create) instead of the reference's
  // inner nodes - same leaves, same order, same hits, about half the nodes to fetch.  SRT_WF_WALKTREE=0: the reference's tree.
  plan.wf_walk = mode != 0 && s->sc.x2_ok && s->sc.x2_nodes && (mode == 1 || s->sc.x2_cnodes);
  if (const char* w = tune_env("SRT_WF_WALKTREE")) if (atoi(w) == 0) plan.wf_walk = 0;
  // Tree in shared memory: k_wf_extend2 (conservative walk + parked leaf tests, srt_extend2.cuh) whenever the scene allows
  // it and its nodes + park lists fit; 16-byte nodes first (half the shared-memory wavefronts per step), 32-byte otherwise.
  plan.x2_fmt = 0; plan.x2_park = SRT_X2_PARK;
  if (mode == 0 && s->sc.x2_ok) {
    int want = SRT_X2_DEFAULT_FMT;
    if (const char* e = tune_env("SRT_WF_X2")) want = atoi(e);
    cudaFuncAttributes fa2{};
    for (int fmt : {want, want == 16 ? 32 : 0}) {
      if (fmt != 16 && fmt != 32) continue;
      if (s->sc.x2_n > X2_MAX_NODES(fmt)) continue;                      // links + the leaf-kind bit of the stop code
      const void* f2 = plan.threads == 1024 ? (const void*)pick_extend2<1024>(dev->counters, fmt)
                       : plan.threads == 768 ? (const void*)pick_extend2<768>(dev->counters, fmt)
                       : plan.threads == 256 ? (const void*)pick_extend2<256>(dev->counters, fmt)
                                             : (const void*)pick_extend2<512>(dev->counters, fmt);
      if (cudaFuncGetAttributes(&fa2, f2) != cudaSuccess) { cudaGetLastError(); continue; }
      const size_t need = x2_smem_bytes(s->sc.x2_n, fmt, plan.threads, SRT_X2_PARK);
      if (need + fa2.sharedSizeBytes > dev->smem_optin) continue;
      plan.x2_park = SRT_X2_PARK;
      plan.x2_fmt = fmt; plan.fn = f2; plan.smem = need;
      if (!wf_env) plan.wf_th_walk = SRT_X2_DEFAULT_TH;
      break;
    }
  }
  // ... and as ONE launch per frame (k_wf_path: the shade step inside the persistent kernel) when its lists fit beside the tree
  plan.fused = 0; plan.path_chunk = 1024;
  if (plan.x2_fmt && (plan.threads == 1024 || plan.threads == 512)) {
    int want = SRT_PATH_DEFAULT;
    if (const char* e = tune_env("SRT_WF_FUSED")) want = atoi(e);
    if (want) {
      const void* fp = plan.threads == 1024 ? (const void*)pick_path<1024>(dev->counters, plan.x2_fmt) : (const void*)pick_path<512>(dev->counters, plan.x2_fmt);
      cudaFuncAttributes fa3{};
      const size_t need = (size_t)s->sc.x2_n * (size_t)plan.x2_fmt + (size_t)(plan.threads / 32) * PATH_WARP_BYTES;
      if (cudaFuncGetAttributes(&fa3, fp) == cudaSuccess && need + fa3.sharedSizeBytes <= dev->smem_optin && s->sc.x2_n * 16ll < (1ll << 24)) {
        plan.fused = 1; plan.fn = fp; plan.smem = need;
      } else cudaGetLastError();
    }
  }
  if (!plan.x2_fmt) {
    extend_fn fn = plan.threads == 1024 ? pick_extend<1024>(dev->counters, mode)
                   : plan.threads == 768 ? pick_extend<768>(dev->counters, mode)
                   : plan.threads == 256 ? pick_extend<256>(dev->counters, mode)
                                         : pick_extend<512>(dev->counters, mode);
    plan.fn = (const void*)fn;
  }
  {   // SRT_WF_SHADE=1: one thread per slot; 2: sorted passes over 512-slot chunks; 3: the same with carried remainders
    int kind = SRT_SHADE_DEFAULT;
    if (const char* sh = tune_env("SRT_WF_SHADE")) { const int v = atoi(sh); if (v >= 1 && v <= 3) kind = v; }
    int ts = SRT_SHADE2_THREADS;
    if (const char* t = tune_env("SRT_WF_SHADE_THREADS")) ts = atoi(t) == 128 ? 128 : 256;
    plan.wf_shade_threads = kind == 2 ? ts : 256;
    plan.wf_shade_chunk = kind == 2 ? 2 * ts : kind == 3 ? 512 : 256;
    if (kind == 1) plan.wf_shade = dev->counters ? (const void*)k_wf_shade<true> : (const void*)k_wf_shade<false>;
    else if (kind == 3) plan.wf_shade = dev->counters ? (const void*)k_wf_shade2c<true> : (const void*)k_wf_shade2c<false>;
    else if (ts == 128) plan.wf_shade = dev->counters ? (const void*)k_wf_shade2<true, 128> : (const void*)k_wf_shade2<false, 128>;
    else plan.wf_shade = dev->counters ? (const void*)k_wf_shade2<true, 256> : (const void*)k_wf_shade2<false, 256>;
  }
  cudaFuncAttributes fa{};
  e = cudaFuncGetAttributes(&fa, plan.fn);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(plan.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(dev->smem_optin - fa.sharedSizeBytes));
  if (e == cudaSuccess && mode != 0 && tune_env("SRT_WF_L1"))      // tree in HBM: ask for the largest L1 the SM can give (A/B)
    e = cudaFuncSetAttribute(plan.fn, cudaFuncAttributePreferredSharedMemoryCarveout, atoi(tune_env("SRT_WF_L1")));
  if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, plan.fn, plan.threads, plan.smem);
  return e;
}

srt_status plan_render(srt_scene* s, const srt_camera* cam, const srt_render_params* P, RenderPlan& plan) {
  if (!s || !cam || !P) return fail(SRT_ERR_INVALID, "NULL argument");
  if (P->width < 2 || P->height < 2) return fail(SRT_ERR_INVALID, "width and height must be >= 2 (main.cpp:210-211 divide by W-1, H-1)");
  if ((int64_t)P->width * P->height > (int64_t)INT32_MAX / 4) return fail(SRT_ERR_UNSUPPORTED, "frame too large");
  if (P->spp <= 0 || P->max_bounce < 0 || P->sample_begin < 0 || P->sample_count < 0) return fail(SRT_ERR_INVALID, "bad sample/bounce counts");
  if (P->variant < 0 || P->variant > 3) return fail(SRT_ERR_INVALID, "render variant must be 0 (auto), 1, 2 or 3");
  RenderArgs& A = plan.A;
  A.cam = *cam; A.W = P->width; A.H = P->height;
  A.sample_begin = P->sample_begin; A.sample_count = P->sample_count > 0 ? P->sample_count : P->spp;
  A.max_bounce = P->max_bounce; A.rr_begin = P->rr_begin;
  if (P->sampler != 0 && P->sampler != 1) return fail(SRT_ERR_INVALID, "sampler must be 0 (random jitter) or 1 (4x rotated grid)");
  A.sampler = P->sampler;
  if (A.sampler == 1) { srt_status rc = ensure_grid(s->dev, A.sample_begin, A.sample_count, A); if (rc != SRT_OK) return rc; }
  A.bg[0] = P->background[0]; A.bg[1] = P->background[1]; A.bg[2] = P->background[2];
  A.seed = P->seed;
  A.tiles_x = (A.W + 7) / 8;
  A.n_tiles = A.tiles_x * ((A.H + 3) / 4);

  srt_device* dev = s->dev;
  plan.threads = 512;
  plan.smem = s->nodes_in_smem ? s->node_bytes : 0;
  // auto: the wavefront needs a few waves over a large slot pool to amortise its launches; small frames stay in one megakernel
  // (measured, B200: C3/C4 1.7x faster as waves; C1 equal; C2's 11-node tree 2x faster in the megakernel)
  plan.variant = P->variant ? P->variant
                            : (((long long)A.n_tiles * 32 * A.sample_count >= (4ll << 20) && s->sc.n_nodes >= 256) ? 2 : 1);
  plan.tune = PhaseTune{16, 12, 20};
  plan.fn = nullptr;
  int per_sm = 0;
  cudaError_t e;
  if (plan.variant == 2) {
    e = plan_wavefront(s, (long long)A.n_tiles * 32 * A.sample_count, plan, per_sm);
  } else if (plan.variant == 3) {
    // SRT_TUNE="threads,th_shade,th_leaf,th_trav" (tuning sweeps only; results do not depend on it)
    if (const char* t = tune_env("SRT_TUNE")) {
      int a = 0, b = 0, c = 0, d = 0;
      if (sscanf(t, "%d,%d,%d,%d", &a, &b, &c, &d) == 4 && (a == 256 || a == 512 || a == 768 || a == 1024) && b >= 1 && c >= 1 && d >= 1) {
        plan.threads = a; plan.tune = PhaseTune{b, c, d};
      }
    }
    phased_fn fn = plan.threads == 1024 ? pick_phased<1024>(dev->counters, s->nodes_in_smem)
                   : plan.threads == 768 ? pick_phased<768>(dev->counters, s->nodes_in_smem)
                   : plan.threads == 256 ? pick_phased<256>(dev->counters, s->nodes_in_smem)
                                         : pick_phased<512>(dev->counters, s->nodes_in_smem);
    plan.fn = (const void*)fn;
    e = cudaFuncSetAttribute(plan.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dev->smem_optin);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, plan.fn, plan.threads, plan.smem);
  } else if (dev->counters) {
    e = cudaFuncSetAttribute(k_render_mega<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dev->smem_optin);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_render_mega<true>, plan.threads, plan.smem);
  } else {
    e = cudaFuncSetAttribute(k_render_mega<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dev->smem_optin);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_render_mega<false>, plan.threads, plan.smem);
  }
  if (e != cudaSuccess) { cudaGetLastError(); return fail(SRT_ERR_CUDA, std::string("occupancy query: ") + cudaGetErrorString(e)); }
  if (per_sm < 1) return fail(SRT_ERR_CUDA, "render kernel does not fit on an SM");
  if (const char* b = tune_env("SRT_WF_CTAS_PER_SM")) { const int k = atoi(b); if (plan.variant == 2 && k >= 1 && k < per_sm) per_sm = k; }   // A/B runs
  plan.blocks = dev->sm_count * per_sm;           // persistent grid: every CTA resident, a multiple of the SM count

  // work units: enough (tile, chunk) pairs per resident warp that the tail is ~1 %, chunks of >= 8 samples
  const long long warps = (long long)plan.blocks * (plan.threads / 32);
  const long long want_units = warps * 64;
  long long n_chunks = (want_units + A.n_tiles - 1) / A.n_tiles;
  if (n_chunks < 1) n_chunks = 1;
  int chunk = (int)((A.sample_count + n_chunks - 1) / n_chunks);
  if (chunk < 8) chunk = 8;
  if (chunk > A.sample_count) chunk = A.sample_count;
  A.chunk = chunk;
  A.n_chunks = (A.sample_count + chunk - 1) / chunk;
  A.n_units = (long long)A.n_tiles * A.n_chunks;
  return SRT_OK;
}

// variant 2: SHADE(+regenerate) / EXTEND waves over the slot pool until the frame's paths are used up
srt_status run_wavefront(srt_scene* s, const RenderPlan& plan, float* accum_dev, srt_stats* stats) {
  srt_device* dev = s->dev;
  const size_t per_slot = 3 * sizeof(float4) + sizeof(int2) + sizeof(uint2);   // ray_o, ray_d, thr | hit, id = 64 bytes
  const size_t N = (size_t)plan.wf_slots;
  if (dev->wf_capacity < N) {
    if (dev->wf.ray_o) { CK(cudaFree(dev->wf.ray_o)); dev->wf.ray_o = nullptr; dev->wf_capacity = 0; }
    void* base = nullptr;
    cudaError_t e = cudaMalloc(&base, N * per_slot);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(SRT_ERR_NOMEM, "wavefront slot pool: " + std::string(cudaGetErrorString(e))); }
    dev->wf.ray_o = (float4*)base;
    dev->wf_capacity = N;
  }
  if (!dev->d_wf_cur) {
    CK(cudaMalloc(&dev->d_wf_cur, sizeof(WfCursors)));
    CK(cudaMallocHost(&dev->h_wf_ring, WF_RING * sizeof(WfCursors)));
    for (int k = 0; k < WF_RING; k++) CK(cudaEventCreateWithFlags(&dev->wf_ev[k], cudaEventDisableTiming));
    for (int k = 0; k < WF_TRING; k++) for (int q = 0; q < 3; q++) CK(cudaEventCreate(&dev->wf_t[k][q]));
  }
  WfPool pool = dev->wf;
  {  // carve the SoA arrays out of the one allocation (capacity-strided so the layout survives smaller frames)
    char* b = (char*)dev->wf.ray_o; const size_t C = dev->wf_capacity;
    pool.ray_o = (float4*)b; b += C * sizeof(float4);
    pool.ray_d = (float4*)b; b += C * sizeof(float4);
    pool.thr = (float4*)b;   b += C * sizeof(float4);
    pool.hit = (int2*)b;     b += C * sizeof(int2);
    pool.id = (uint2*)b;
    pool.n_slots = (int)N;
    pool.order = nullptr;
  }
  // ray ordering (srt_wavefront.cuh: k_wf_keys) for trees in HBM; SRT_WF_SORT=0|1 overrides (A/B runs)
  bool sort_rays = !s->nodes_in_smem && s->node_bytes > dev->l2_bytes;
  if (const char* e = tune_env("SRT_WF_SORT")) sort_rays = atoi(e) != 0;
  WfSortGrid grid{};
  if (const char* e = tune_env("SRT_WF_KEY")) grid.mode = atoi(e);
  unsigned *keys_in = nullptr, *keys_out = nullptr, *vals_in = nullptr, *vals_out = nullptr;
  if (sort_rays) {
    if (dev->sort_cap < N) {
      if (dev->d_sort) { CK(cudaFree(dev->d_sort)); dev->d_sort = nullptr; dev->sort_cap = 0; }
      if (dev->d_sort_tmp) { CK(cudaFree(dev->d_sort_tmp)); dev->d_sort_tmp = nullptr; dev->sort_tmp_bytes = 0; }
      cudaError_t e = cudaMalloc(&dev->d_sort, 4 * N * sizeof(unsigned));
      if (e != cudaSuccess) { cudaGetLastError(); return fail(SRT_ERR_NOMEM, "ray-order buffers: " + std::string(cudaGetErrorString(e))); }
      dev->sort_cap = N;
      size_t tmp = 0;
      CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp, (const unsigned*)nullptr, (unsigned*)nullptr, (const unsigned*)nullptr,
                                         (unsigned*)nullptr, (int)N, 0, 28, dev->stream));
      CK(cudaMalloc(&dev->d_sort_tmp, tmp));
      dev->sort_tmp_bytes = tmp;
    }
    keys_in = dev->d_sort; keys_out = keys_in + dev->sort_cap; vals_in = keys_out + dev->sort_cap; vals_out = vals_in + dev->sort_cap;
    for (int k = 0; k < 3; k++) {
      const float lo = s->root_lo[k], ext = s->root_hi[k] - s->root_lo[k];
      grid.org[k] = lo; grid.inv_cell[k] = ext > 0 ? 256.0f / ext : 0.0f;
    }
    pool.order = vals_out;
  }
  const unsigned long long total = (unsigned long long)plan.A.n_tiles * 32ull * (unsigned long long)plan.A.sample_count;
  Scene sc = s->sc; RenderArgs A = plan.A; WfCursors* cur = dev->d_wf_cur; unsigned long long* ctr = dev->d_counters;
  int th_walk = plan.wf_th_walk;
  void* shade_args[] = {&sc, &A, &pool, &cur, &accum_dev, &ctr};
  int x2_park = plan.x2_park;
  Scene sce = sc;                                      // what the extend kernel walks
  if (plan.wf_walk) { sce.nodes = sc.x2_nodes; sce.n_nodes = sc.x2_n; if (sc.x2_cnodes) sce.cnodes = sc.x2_cnodes; }
  void* extend_args_v1[] = {&sce, &pool, &cur, &th_walk, &ctr};
  void* extend_args_x2[] = {&sce, &pool, &cur, &th_walk, &x2_park, &ctr};
  void** extend_args = plan.x2_fmt ? extend_args_x2 : extend_args_v1;
  const unsigned reset_blocks = (unsigned)((N + 255) / 256);
  unsigned shade_blocks = (unsigned)dev->sm_count * 2 * SRT_SHADE_MIN_BLOCKS * (256 / plan.wf_shade_threads);   // persistent: resident CTAs per SM x two rounds
  {
    const unsigned chunks = (unsigned)((N + plan.wf_shade_chunk - 1) / plan.wf_shade_chunk);
    if (shade_blocks > chunks) shade_blocks = chunks;
  }
  int launches = 0;

  CK(cudaMemsetAsync(dev->d_counters, 0, 8 * sizeof(unsigned long long), dev->stream));
  CK(cudaMemsetAsync(cur, 0, sizeof(WfCursors), dev->stream));
  CK(cudaEventRecord(dev->ev0, dev->stream));
  k_wf_reset<<<reset_blocks, 256, 0, dev->stream>>>(pool);
  ++launches;
  // The host learns (alive, path cursor) of wave j when it is about to launch wave j + LAG: the queue never drains,
  // and at most LAG empty waves run after the frame is finished.
  const int LAG = 4;
  double shade_ms = 0, extend_ms = 0;
  long long waves = 0, timed = 0;
  auto collect = [&](long long w) -> cudaError_t {      // wave w is known to be complete
    cudaEvent_t* t = dev->wf_t[w % WF_TRING];
    float a = 0, b = 0;
    cudaError_t e = cudaEventElapsedTime(&a, t[0], t[1]);
    if (e == cudaSuccess) e = cudaEventElapsedTime(&b, t[1], t[2]);
    shade_ms += a; extend_ms += b; ++timed;
    return e;
  };
  for (long long j = 0;; j++) {
    if (j >= LAG) {
      const int slot = (int)((j - LAG) % WF_RING);
      CK(cudaEventSynchronize(dev->wf_ev[slot]));       // after wave j-LAG's shade => wave j-LAG-1 is complete
      if (j >= LAG + 1) CK(collect(j - LAG - 1));
      const WfCursors& c = dev->h_wf_ring[slot];
      if (c.alive == 0 && c.path >= total) break;
    }
    cudaEvent_t* t = dev->wf_t[j % WF_TRING];
    CK(cudaEventRecord(t[0], dev->stream));
    CK(cudaLaunchKernel(plan.wf_shade, dim3(shade_blocks), dim3(plan.wf_shade_threads), shade_args, 0, dev->stream));
    CK(cudaEventRecord(t[1], dev->stream));
    CK(cudaMemcpyAsync(&dev->h_wf_ring[j % WF_RING], cur, sizeof(WfCursors), cudaMemcpyDeviceToHost, dev->stream));
    CK(cudaEventRecord(dev->wf_ev[j % WF_RING], dev->stream));
    if (sort_rays) {
      k_wf_keys<<<reset_blocks, 256, 0, dev->stream>>>(pool, grid, keys_in, vals_in);
      size_t tmp = dev->sort_tmp_bytes;
      CK(cub::DeviceRadixSort::SortPairs(dev->d_sort_tmp, tmp, (const unsigned*)keys_in, keys_out, (const unsigned*)vals_in, vals_out,
                                         (int)N, 0, 28, dev->stream));
      ++launches;                                        // k_wf_keys; the library's sort passes are not counted as ours
    }
    CK(cudaLaunchKernel(plan.fn, dim3(plan.blocks), dim3(plan.threads), extend_args, plan.smem, dev->stream));
    CK(cudaEventRecord(t[2], dev->stream));
    launches += 2; ++waves;
  }
  CK(cudaEventRecord(dev->ev1, dev->stream));
  unsigned long long h[8];
  CK(cudaMemcpyAsync(h, dev->d_counters, sizeof h, cudaMemcpyDeviceToHost, dev->stream));
  CK(cudaStreamSynchronize(dev->stream));
  if (stats) {
    memset(stats, 0, sizeof *stats);
    stats->paths = h[0]; stats->segments = h[1]; stats->node_visits = h[2]; stats->leaf_tests = h[3];
    float ms = 0; CK(cudaEventElapsedTime(&ms, dev->ev0, dev->ev1));
    stats->kernel_ms = ms; stats->launches = launches;
    for (long long w = timed; w < waves; w++) CK(collect(w));     // the last LAG+1 waves (the stream is idle now)
    stats->extend_ms = (float)extend_ms; stats->shade_ms = (float)shade_ms;
    stats->extend_launches = (int)waves; stats->variant = 2; stats->extend_kernel = plan.x2_fmt;
  }
  return SRT_OK;
}

// variant 2, fused: one persistent launch traces, shades and regenerates until the frame's paths are used up (srt_path.cuh)
srt_status run_path(srt_scene* s, const RenderPlan& plan, float* accum_dev, srt_stats* stats) {
  srt_device* dev = s->dev;
  const size_t N = (size_t)plan.blocks * (size_t)(plan.threads / 32) * PATH_SLOTS;     // warp-private slots
  const size_t per_slot = 3 * sizeof(float4) + sizeof(int2) + sizeof(uint2);
  if (dev->path_capacity < N) {
    if (dev->d_path) { CK(cudaFree(dev->d_path)); dev->d_path = nullptr; dev->path_capacity = 0; }
    cudaError_t e = cudaMalloc(&dev->d_path, N * per_slot);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(SRT_ERR_NOMEM, "path slots: " + std::string(cudaGetErrorString(e))); }
    dev->path_capacity = N;
  }
  if (!dev->d_wf_cur) {
    CK(cudaMalloc(&dev->d_wf_cur, sizeof(WfCursors)));
    CK(cudaMallocHost(&dev->h_wf_ring, WF_RING * sizeof(WfCursors)));
    for (int k = 0; k < WF_RING; k++) CK(cudaEventCreateWithFlags(&dev->wf_ev[k], cudaEventDisableTiming));
    for (int k = 0; k < WF_TRING; k++) for (int q = 0; q < 3; q++) CK(cudaEventCreate(&dev->wf_t[k][q]));
  }
  PathPool pool;
  {
    char* b = (char*)dev->d_path; const size_t C = dev->path_capacity;
    pool.ray_o = (float4*)b; b += C * sizeof(float4);
    pool.ray_d = (float4*)b; b += C * sizeof(float4);
    pool.thr = (float4*)b;   b += C * sizeof(float4);
    pool.hit = (int2*)b;     b += C * sizeof(int2);
    pool.id = (uint2*)b;
  }
  // path ids per global atomic: 1024 (32 pixel tiles of one sample index), fewer when the frame is small against the grid
  const unsigned long long total = (unsigned long long)plan.A.n_tiles * 32ull * (unsigned long long)plan.A.sample_count;
  const unsigned long long warps = (unsigned long long)plan.blocks * (unsigned long long)(plan.threads / 32);
  int chunk = (int)std::min<unsigned long long>(1024, std::max<unsigned long long>(32, total / (warps * 4) / 32 * 32));
  Scene sc = s->sc; RenderArgs A = plan.A; WfCursors* cur = dev->d_wf_cur; unsigned long long* ctr = dev->d_counters;
  int th_walk = plan.wf_th_walk;
  void* args[] = {&sc, &A, &pool, &cur, &accum_dev, &th_walk, &chunk, &ctr};
  CK(cudaMemsetAsync(dev->d_counters, 0, 8 * sizeof(unsigned long long), dev->stream));
  CK(cudaMemsetAsync(cur, 0, sizeof(WfCursors), dev->stream));
  CK(cudaEventRecord(dev->ev0, dev->stream));
  CK(cudaLaunchKernel(plan.fn, dim3(plan.blocks), dim3(plan.threads), args, plan.smem, dev->stream));
  CK(cudaEventRecord(dev->ev1, dev->stream));
  unsigned long long h[8];
  CK(cudaMemcpyAsync(h, dev->d_counters, sizeof h, cudaMemcpyDeviceToHost, dev->stream));
  CK(cudaStreamSynchronize(dev->stream));
  if (stats) {
    memset(stats, 0, sizeof *stats);
    stats->paths = h[0]; stats->segments = h[1]; stats->node_visits = h[2]; stats->leaf_tests = h[3];
    float ms = 0; CK(cudaEventElapsedTime(&ms, dev->ev0, dev->ev1));
    stats->kernel_ms = ms; stats->launches = 1;
    stats->extend_ms = ms; stats->shade_ms = 0; stats->extend_launches = 1; stats->variant = 2; stats->extend_kernel = plan.x2_fmt;
    stats->fused = 1;
  }
  return SRT_OK;
}

srt_status run_render(srt_scene* s, const RenderPlan& plan, float* accum_dev, srt_stats* stats) {
  srt_device* dev = s->dev;
  if (plan.A.max_bounce <= 0) {     // rayColor returns (0,0,0) before world.hit (main.cpp:36-37): nothing to add, no segment traced
    if (stats) {
      memset(stats, 0, sizeof *stats);
      stats->paths = (uint64_t)plan.A.W * plan.A.H * (uint64_t)plan.A.sample_count;
      stats->variant = plan.variant;
    }
    return SRT_OK;
  }
  if (plan.variant == 2) return plan.fused ? run_path(s, plan, accum_dev, stats) : run_wavefront(s, plan, accum_dev, stats);
  CK(cudaMemsetAsync(dev->d_counters, 0, 8 * sizeof(unsigned long long), dev->stream));
  CK(cudaEventRecord(dev->ev0, dev->stream));
  if (plan.variant == 3) {
    Scene sc = s->sc; RenderArgs A = plan.A; PhaseTune tune = plan.tune; unsigned long long* ctr = dev->d_counters;
    void* args[] = {&sc, &A, &tune, &accum_dev, &ctr};
    CK(cudaLaunchKernel(plan.fn, dim3(plan.blocks), dim3(plan.threads), args, plan.smem, dev->stream));
  } else if (dev->counters)
    k_render_mega<true><<<plan.blocks, plan.threads, plan.smem, dev->stream>>>(s->sc, s->nodes_in_smem, plan.A, accum_dev, dev->d_counters);
  else
    k_render_mega<false><<<plan.blocks, plan.threads, plan.smem, dev->stream>>>(s->sc, s->nodes_in_smem, plan.A, accum_dev, dev->d_counters);
  CK(cudaGetLastError());
  CK(cudaEventRecord(dev->ev1, dev->stream));
  unsigned long long h[8];
  CK(cudaMemcpyAsync(h, dev->d_counters, sizeof h, cudaMemcpyDeviceToHost, dev->stream));
  CK(cudaStreamSynchronize(dev->stream));
  if (stats) {
    memset(stats, 0, sizeof *stats);
    stats->paths = h[0]; stats->segments = h[1]; stats->node_visits = h[2]; stats->leaf_tests = h[3];
    float ms = 0; CK(cudaEventElapsedTime(&ms, dev->ev0, dev->ev1));
    stats->kernel_ms = ms; stats->launches = 1;
    stats->extend_ms = ms; stats->shade_ms = 0; stats->extend_launches = 1; stats->variant = plan.variant;
  }
  return SRT_OK;
}

}  // namespace

extern "C" {

srt_status srt_render_accumulate(srt_scene* s, const srt_camera* cam, const srt_render_params* P, float* accum_dev,
                                 srt_stats* stats) try {
  if (!accum_dev) return fail(SRT_ERR_INVALID, "accum_dev is NULL");
  RenderPlan plan;
  srt_status rc = plan_render(s, cam, P, plan);
  if (rc) return rc;
  CK(cudaSetDevice(s->dev->gpu));
  return run_render(s, plan, accum_dev, stats);
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_resolve(srt_device* dev, const float* sum_dev, int32_t W, int32_t H, int32_t spp, uint8_t* rgba8_dev) try {
  if (!dev || !sum_dev || !rgba8_dev || W <= 0 || H <= 0 || spp <= 0) return fail(SRT_ERR_INVALID, "bad argument");
  CK(cudaSetDevice(dev->gpu));
  int n = W * H;
  k_resolve<<<(n + 255) / 256, 256, 0, dev->stream>>>(sum_dev, n, spp, reinterpret_cast<uchar4*>(rgba8_dev));
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(dev->stream));
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_render(srt_scene* s, const srt_camera* cam, const srt_render_params* P, float* out_sum_rgb,
                      uint8_t* out_rgba8, srt_stats* stats) try {
  RenderPlan plan;
  srt_status rc = plan_render(s, cam, P, plan);
  if (rc) return rc;
  srt_device* dev = s->dev;
  CK(cudaSetDevice(dev->gpu));
  const size_t n = (size_t)P->width * P->height;
  // the frame's device buffers live in the device object (grow-only): a frame costs no cudaMalloc / cudaFree
  if (dev->frame_cap < n) {
    if (dev->d_frame_sum) { CK(cudaFree(dev->d_frame_sum)); dev->d_frame_sum = nullptr; }
    if (dev->d_frame_rgba) { CK(cudaFree(dev->d_frame_rgba)); dev->d_frame_rgba = nullptr; }
    dev->frame_cap = 0;
    CK(cudaMalloc(&dev->d_frame_sum, n * 3 * sizeof(float)));
    CK(cudaMalloc(&dev->d_frame_rgba, n * 4));
    dev->frame_cap = n;
  }
  float* acc = dev->d_frame_sum;
  CK(cudaMemsetAsync(acc, 0, n * 3 * sizeof(float), dev->stream));
  rc = run_render(s, plan, acc, stats);
  if (rc) return rc;
  if (out_rgba8) {
    k_resolve<<<(unsigned)((n + 255) / 256), 256, 0, dev->stream>>>(acc, (int)n, P->spp, dev->d_frame_rgba);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out_rgba8, dev->d_frame_rgba, n * 4, cudaMemcpyDeviceToHost, dev->stream));
    if (stats) stats->launches += 1;
  }
  if (out_sum_rgb) CK(cudaMemcpyAsync(out_sum_rgb, acc, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, dev->stream));
  CK(cudaStreamSynchronize(dev->stream));
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_trace_rays(srt_scene* s, const srt_ray* rays, int64_t n, float t_min, float t_max, srt_hit* out_hits) try {
  if (!s || n < 0 || (n > 0 && (!rays || !out_hits))) return fail(SRT_ERR_INVALID, "bad argument");
  if (n == 0) return SRT_OK;
  srt_device* dev = s->dev;
  CK(cudaSetDevice(dev->gpu));
  DevBuf dr, dh;
  CK(dr.alloc((size_t)n * sizeof(srt_ray)));
  CK(dh.alloc((size_t)n * sizeof(srt_hit)));
  CK(cudaMemcpyAsync(dr.p, rays, (size_t)n * sizeof(srt_ray), cudaMemcpyHostToDevice, dev->stream));
  const int threads = 256;
  size_t smem = s->nodes_in_smem ? s->node_bytes : 0;
  long long want = (n + threads - 1) / threads;
  int blocks = (int)(want < (long long)dev->sm_count * 8 ? want : (long long)dev->sm_count * 8);
  CK(cudaMemsetAsync(dev->d_counters, 0, 8 * sizeof(unsigned long long), dev->stream));
  if (dev->counters) {
    CK(cudaFuncSetAttribute(k_trace_rays<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dev->smem_optin));
    k_trace_rays<true><<<blocks, threads, smem, dev->stream>>>(s->sc, s->nodes_in_smem, dr.as<srt_ray>(), n, t_min, t_max, dh.as<srt_hit>(), dev->d_counters);
  } else {
    CK(cudaFuncSetAttribute(k_trace_rays<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dev->smem_optin));
    k_trace_rays<false><<<blocks, threads, smem, dev->stream>>>(s->sc, s->nodes_in_smem, dr.as<srt_ray>(), n, t_min, t_max, dh.as<srt_hit>(), dev->d_counters);
  }
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(out_hits, dh.p, (size_t)n * sizeof(srt_hit), cudaMemcpyDeviceToHost, dev->stream));
  CK(cudaStreamSynchronize(dev->stream));
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

// hittableList::hit for explicit rays through the PRODUCTION extend kernel of the wavefront (the launch shape, tree mode and
// ray ordering plan_wavefront picks for this scene): rays are loaded into pool slots, one EXTEND wave runs, (prim, t) come back.
srt_status srt_trace_rays_wavefront(srt_scene* s, const srt_ray* rays, int64_t n, int32_t* out_prim, float* out_t) try {
  if (!s || n < 0 || (n > 0 && (!rays || !out_prim || !out_t))) return fail(SRT_ERR_INVALID, "bad argument");
  if (n == 0) return SRT_OK;
  if (n > (int64_t)1 << 30) return fail(SRT_ERR_UNSUPPORTED, "too many rays for one wave");
  srt_device* dev = s->dev;
  CK(cudaSetDevice(dev->gpu));
  RenderPlan plan;
  int per_sm = 0;
  cudaError_t e = plan_wavefront(s, (long long)n, plan, per_sm);
  if (e != cudaSuccess) { cudaGetLastError(); return fail(SRT_ERR_CUDA, std::string("occupancy query: ") + cudaGetErrorString(e)); }
  if (per_sm < 1) return fail(SRT_ERR_CUDA, "extend kernel does not fit on an SM");
  if (const char* b = tune_env("SRT_WF_CTAS_PER_SM")) { const int k = atoi(b); if (k >= 1 && k < per_sm) per_sm = k; }
  const int blocks = dev->sm_count * per_sm;
  DevBuf dr, dro, drd, dh, dcur, dsort, dtmp;
  CK(dr.alloc((size_t)n * sizeof(srt_ray)));
  CK(dro.alloc((size_t)n * sizeof(float4))); CK(drd.alloc((size_t)n * sizeof(float4))); CK(dh.alloc((size_t)n * sizeof(int2)));
  CK(dcur.alloc(sizeof(WfCursors)));
  CK(cudaMemcpyAsync(dr.p, rays, (size_t)n * sizeof(srt_ray), cudaMemcpyHostToDevice, dev->stream));
  CK(cudaMemsetAsync(dcur.p, 0, sizeof(WfCursors), dev->stream));
  CK(cudaMemsetAsync(dev->d_counters, 0, 8 * sizeof(unsigned long long), dev->stream));
  WfPool pool{};
  pool.ray_o = dro.as<float4>(); pool.ray_d = drd.as<float4>(); pool.hit = dh.as<int2>(); pool.n_slots = (int)n; pool.order = nullptr;
  const unsigned nb = (unsigned)((n + 255) / 256);
  k_wf_load_rays<<<nb, 256, 0, dev->stream>>>(dr.as<srt_ray>(), (long long)n, pool);
  bool sort_rays = !s->nodes_in_smem && s->node_bytes > dev->l2_bytes;
  if (const char* ev = tune_env("SRT_WF_SORT")) sort_rays = atoi(ev) != 0;
  if (sort_rays) {
    WfSortGrid grid{};
    for (int k = 0; k < 3; k++) {
      const float lo = s->root_lo[k], ext = s->root_hi[k] - s->root_lo[k];
      grid.org[k] = lo; grid.inv_cell[k] = ext > 0 ? 256.0f / ext : 0.0f;
    }
    CK(dsort.alloc(4 * (size_t)n * sizeof(unsigned)));
    unsigned *keys_in = dsort.as<unsigned>(), *keys_out = keys_in + n, *vals_in = keys_out + n, *vals_out = vals_in + n;
    size_t tmp = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp, (const unsigned*)nullptr, (unsigned*)nullptr, (const unsigned*)nullptr, (unsigned*)nullptr,
                                       (int)n, 0, 28, dev->stream));
    CK(dtmp.alloc(tmp));
    k_wf_keys<<<nb, 256, 0, dev->stream>>>(pool, grid, keys_in, vals_in);
    CK(cub::DeviceRadixSort::SortPairs(dtmp.p, tmp, (const unsigned*)keys_in, keys_out, (const unsigned*)vals_in, vals_out, (int)n, 0, 28, dev->stream));
    pool.order = vals_out;
  }
  Scene sc = s->sc; WfCursors* cur = dcur.as<WfCursors>(); unsigned long long* ctr = dev->d_counters;
  if (plan.wf_walk) { sc.nodes = sc.x2_nodes; sc.n_nodes = sc.x2_n; if (sc.x2_cnodes) sc.cnodes = sc.x2_cnodes; }
  int th_walk = plan.wf_th_walk, x2_park = plan.x2_park;
  void* args_v1[] = {&sc, &pool, &cur, &th_walk, &ctr};
  void* args_x2[] = {&sc, &pool, &cur, &th_walk, &x2_park, &ctr};
  CK(cudaLaunchKernel(plan.fn, dim3(blocks), dim3(plan.threads), plan.x2_fmt ? args_x2 : args_v1, plan.smem, dev->stream));
  std::vector<int2> h((size_t)n);
  CK(cudaMemcpyAsync(h.data(), dh.p, (size_t)n * sizeof(int2), cudaMemcpyDeviceToHost, dev->stream));
  CK(cudaStreamSynchronize(dev->stream));
  for (int64_t i = 0; i < n; i++) { out_prim[i] = h[i].x; memcpy(&out_t[i], &h[i].y, 4); }
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

// aabb::hit (aabb.h:11-27) decisions of the device's box predicates on explicit (ray, box, tMin, tMax) rows - see k_aabb_kat.
// boxes[n*6] = min3 max3; tminmax[n*2]; out_path optional.  Predicates 2 and 3 (the conservative cull of k_wf_extend2) use the
// walk's constant tMin = 0.001 and return 1 = kept / 0 = culled: a row the reference HITS must never come back 0.
srt_status srt_aabb_kat(srt_device* dev, const srt_ray* rays, const float* boxes, const float* tminmax, int64_t n, int32_t predicate,
                        int32_t* out_decision, int32_t* out_path) try {
  if (!dev || n < 0 || predicate < 0 || predicate > 3 || (n > 0 && (!rays || !boxes || !tminmax || !out_decision)))
    return fail(SRT_ERR_INVALID, "bad argument");
  if (n == 0) return SRT_OK;
  CK(cudaSetDevice(dev->gpu));
  KatGrid g{};
  {   // the 16-bit grid over the union of the boxes (predicate 1) and the centre of the walk's coordinates (2, 3)
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    for (int64_t i = 0; i < n; i++) for (int k = 0; k < 3; k++) {
      const float a = boxes[6 * i + k], b = boxes[6 * i + 3 + k];
      if (std::isfinite(a) && a < lo[k]) lo[k] = a;
      if (std::isfinite(b) && b > hi[k]) hi[k] = b;
    }
    for (int k = 0; k < 3; k++) {
      if (!(hi[k] >= lo[k])) { lo[k] = 0; hi[k] = 1; }
      float c = (float)((hi[k] - lo[k]) / 65534.0);
      c = std::nextafter(std::max(c, 1e-30f), std::numeric_limits<float>::infinity());
      g.org[k] = (float)lo[k]; if ((double)g.org[k] > lo[k]) g.org[k] = std::nextafter(g.org[k], -std::numeric_limits<float>::infinity());
      g.cell[k] = c;
      g.centre[k] = (float)(0.5 * (lo[k] + hi[k]));
    }
  }
  DevBuf dr, db, dt, dd, dp;
  CK(dr.alloc((size_t)n * sizeof(srt_ray))); CK(db.alloc((size_t)n * 24)); CK(dt.alloc((size_t)n * 8)); CK(dd.alloc((size_t)n * 4)); CK(dp.alloc((size_t)n * 4));
  CK(cudaMemcpyAsync(dr.p, rays, (size_t)n * sizeof(srt_ray), cudaMemcpyHostToDevice, dev->stream));
  CK(cudaMemcpyAsync(db.p, boxes, (size_t)n * 24, cudaMemcpyHostToDevice, dev->stream));
  CK(cudaMemcpyAsync(dt.p, tminmax, (size_t)n * 8, cudaMemcpyHostToDevice, dev->stream));
  k_aabb_kat<<<(unsigned)((n + 127) / 128), 128, 0, dev->stream>>>(dr.as<srt_ray>(), db.as<float>(), dt.as<float>(), (long long)n, predicate, g,
                                                                  dd.as<int>(), dp.as<int>());
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(out_decision, dd.p, (size_t)n * 4, cudaMemcpyDeviceToHost, dev->stream));
  if (out_path) CK(cudaMemcpyAsync(out_path, dp.p, (size_t)n * 4, cudaMemcpyDeviceToHost, dev->stream));
  CK(cudaStreamSynchronize(dev->stream));
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

// counters of the last srt_trace_rays or render call (only with srt_set_counters(dev, 1)):
// node visits, leaf tests, exact (unfiltered) box tests, sphere leaf tests, segments ending on a triangle
srt_status srt_last_counters(srt_device* dev, uint64_t* out5) try {
  uint64_t* out3 = out5;
  if (!dev || !out5) return fail(SRT_ERR_INVALID, "NULL argument");
  unsigned long long h[8];
  CK(cudaSetDevice(dev->gpu));
  CK(cudaMemcpy(h, dev->d_counters, sizeof h, cudaMemcpyDeviceToHost));
  out3[0] = h[2]; out3[1] = h[3]; out3[2] = h[4]; out3[3] = h[6]; out3[4] = h[7];
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_shade_kat(srt_scene* s, const srt_ray* rays_in, const srt_hit* hits, const float* xi, int64_t n,
                         float* out_att, float* out_emit, srt_ray* out_rays, int32_t* out_scattered) try {
  if (!s || n < 0) return fail(SRT_ERR_INVALID, "bad argument");
  if (n == 0) return SRT_OK;
  for (int64_t i = 0; i < n; i++) if (hits[i].mat < 0 || hits[i].mat >= s->sc.n_mat) return fail(SRT_ERR_INVALID, "hit material id out of range");
  srt_device* dev = s->dev;
  CK(cudaSetDevice(dev->gpu));
  DevBuf dr, dh, dx, da, de, dout, ds;
  CK(dr.alloc(n * sizeof(srt_ray))); CK(dh.alloc(n * sizeof(srt_hit))); CK(dx.alloc(n * 16)); CK(da.alloc(n * 12));
  CK(de.alloc(n * 12)); CK(dout.alloc(n * sizeof(srt_ray))); CK(ds.alloc(n * 4));
  CK(cudaMemcpyAsync(dr.p, rays_in, n * sizeof(srt_ray), cudaMemcpyHostToDevice, dev->stream));
  CK(cudaMemcpyAsync(dh.p, hits, n * sizeof(srt_hit), cudaMemcpyHostToDevice, dev->stream));
  CK(cudaMemcpyAsync(dx.p, xi, n * 16, cudaMemcpyHostToDevice, dev->stream));
  k_shade_kat<<<(unsigned)((n + 127) / 128), 128, 0, dev->stream>>>(s->sc, dr.as<srt_ray>(), dh.as<srt_hit>(), dx.as<float>(), n,
                                                                   da.as<float>(), de.as<float>(), dout.as<srt_ray>(), ds.as<int>());
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(out_att, da.p, n * 12, cudaMemcpyDeviceToHost, dev->stream));
  CK(cudaMemcpyAsync(out_emit, de.p, n * 12, cudaMemcpyDeviceToHost, dev->stream));
  CK(cudaMemcpyAsync(out_rays, dout.p, n * sizeof(srt_ray), cudaMemcpyDeviceToHost, dev->stream));
  CK(cudaMemcpyAsync(out_scattered, ds.p, n * 4, cudaMemcpyDeviceToHost, dev->stream));
  CK(cudaStreamSynchronize(dev->stream));
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_texture_kat(srt_scene* s, int32_t tex, const float* uvp, int64_t n, float* out_rgb) try {
  if (!s || n < 0 || tex < 0 || tex >= s->sc.n_tex) return fail(SRT_ERR_INVALID, "bad argument");
  if (n == 0) return SRT_OK;
  srt_device* dev = s->dev;
  CK(cudaSetDevice(dev->gpu));
  DevBuf di, dout;
  CK(di.alloc(n * 20)); CK(dout.alloc(n * 12));
  CK(cudaMemcpyAsync(di.p, uvp, n * 20, cudaMemcpyHostToDevice, dev->stream));
  k_texture_kat<<<(unsigned)((n + 127) / 128), 128, 0, dev->stream>>>(s->sc, tex, di.as<float>(), n, dout.as<float>());
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(out_rgb, dout.p, n * 12, cudaMemcpyDeviceToHost, dev->stream));
  CK(cudaStreamSynchronize(dev->stream));
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_getray_kat(srt_device* dev, const srt_camera* cam, const float* in, int64_t n, srt_ray* out_rays) try {
  if (!dev || !cam || n < 0) return fail(SRT_ERR_INVALID, "bad argument");
  if (n == 0) return SRT_OK;
  CK(cudaSetDevice(dev->gpu));
  DevBuf di, dout;
  CK(di.alloc(n * 20)); CK(dout.alloc(n * sizeof(srt_ray)));
  CK(cudaMemcpyAsync(di.p, in, n * 20, cudaMemcpyHostToDevice, dev->stream));
  k_getray_kat<<<(unsigned)((n + 127) / 128), 128, 0, dev->stream>>>(*cam, di.as<float>(), n, dout.as<srt_ray>());
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(out_rays, dout.p, n * sizeof(srt_ray), cudaMemcpyDeviceToHost, dev->stream));
  CK(cudaStreamSynchronize(dev->stream));
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_primary_rays(srt_device* dev, const srt_camera* cam, const srt_render_params* P, const int32_t* pixel_xy,
                            const int32_t* sample, int64_t n, srt_ray* out_rays) try {
  if (!dev || !cam || !P || n < 0) return fail(SRT_ERR_INVALID, "bad argument");
  if (n == 0) return SRT_OK;
  CK(cudaSetDevice(dev->gpu));
  RenderArgs A; memset(&A, 0, sizeof A);
  A.cam = *cam; A.W = P->width; A.H = P->height; A.seed = P->seed;
  if (P->sampler == 1) {
    int top = 0;
    for (int64_t i = 0; i < n; i++) { if (sample[i] < 0) return fail(SRT_ERR_INVALID, "negative sample index"); top = sample[i] > top ? sample[i] : top; }
    srt_status rc = ensure_grid(dev, 0, top + 1, A);
    if (rc != SRT_OK) return rc;
  }
  DevBuf dp, dsm, dout;
  CK(dp.alloc(n * 8)); CK(dsm.alloc(n * 4)); CK(dout.alloc(n * sizeof(srt_ray)));
  CK(cudaMemcpyAsync(dp.p, pixel_xy, n * 8, cudaMemcpyHostToDevice, dev->stream));
  CK(cudaMemcpyAsync(dsm.p, sample, n * 4, cudaMemcpyHostToDevice, dev->stream));
  k_primary_rays<<<(unsigned)((n + 127) / 128), 128, 0, dev->stream>>>(A, dp.as<int>(), dsm.as<int>(), n, dout.as<srt_ray>());
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(out_rays, dout.p, n * sizeof(srt_ray), cudaMemcpyDeviceToHost, dev->stream));
  CK(cudaStreamSynchronize(dev->stream));
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

}  // extern "C"

// ---- progressive accumulation: the role of glDevice::rtFrame as a per-frame loop (main.cpp:195-199, gl.h:269-320), headless ----
struct srt_film {
  srt_device* dev = nullptr;
  int W = 0, H = 0;
  int done = 0;                 // samples per pixel accumulated so far = the next global sample index
  float* sum = nullptr;         // device, [H*W*3]
};

extern "C" {

srt_status srt_film_create(srt_device* dev, int32_t W, int32_t H, srt_film** out) try {
  if (!dev || !out || W <= 0 || H <= 0) return fail(SRT_ERR_INVALID, "bad argument");
  *out = nullptr;
  CK(cudaSetDevice(dev->gpu));
  srt_film* f = new srt_film;
  f->dev = dev; f->W = W; f->H = H;
  const size_t nb = (size_t)W * H * 3 * sizeof(float);
  cudaError_t e = cudaMalloc(&f->sum, nb);
  if (e != cudaSuccess) { cudaGetLastError(); delete f; return fail(e == cudaErrorMemoryAllocation ? SRT_ERR_NOMEM : SRT_ERR_CUDA, cudaGetErrorString(e)); }
  e = cudaMemsetAsync(f->sum, 0, nb, dev->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(dev->stream);
  if (e != cudaSuccess) { cudaGetLastError(); cudaFree(f->sum); delete f; return fail(SRT_ERR_CUDA, cudaGetErrorString(e)); }
  *out = f;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

void srt_film_destroy(srt_film* f) {
  if (!f) return;
  cudaSetDevice(f->dev->gpu);
  if (f->sum) cudaFree(f->sum);
  delete f;
}

int32_t srt_film_samples(const srt_film* f) { return f ? f->done : 0; }

srt_status srt_film_add_samples(srt_film* f, srt_scene* s, const srt_camera* cam, const srt_render_params* P, int32_t count,
                                srt_stats* stats) try {
  if (!f || !s || !P || count <= 0) return fail(SRT_ERR_INVALID, "bad argument");
  if (s->dev != f->dev) return fail(SRT_ERR_INVALID, "film and scene live on different devices");
  if (P->width != f->W || P->height != f->H) return fail(SRT_ERR_INVALID, "render size differs from the film's");
  srt_render_params q = *P;
  q.sample_begin = f->done; q.sample_count = count;
  q.spp = f->done + count;                      // only resolve divides by spp; the sample range is what the kernels use
  srt_status rc = srt_render_accumulate(s, cam, &q, f->sum, stats);
  if (rc) return rc;
  f->done += count;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_film_resolve(srt_film* f, uint8_t* out_rgba8, float* out_sum_rgb) try {
  if (!f) return fail(SRT_ERR_INVALID, "film is NULL");
  if (f->done <= 0) return fail(SRT_ERR_INVALID, "no samples accumulated yet");
  srt_device* dev = f->dev;
  CK(cudaSetDevice(dev->gpu));
  const size_t n = (size_t)f->W * f->H;
  if (out_rgba8) {
    DevBuf rgba;
    CK(rgba.alloc(n * 4));
    k_resolve<<<(unsigned)((n + 255) / 256), 256, 0, dev->stream>>>(f->sum, (int)n, f->done, rgba.as<uchar4>());
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out_rgba8, rgba.p, n * 4, cudaMemcpyDeviceToHost, dev->stream));
    CK(cudaStreamSynchronize(dev->stream));
  }
  if (out_sum_rgb) {
    CK(cudaMemcpyAsync(out_sum_rgb, f->sum, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, dev->stream));
    CK(cudaStreamSynchronize(dev->stream));
  }
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_status srt_film_load(srt_film* f, const float* sum_rgb, int32_t samples_done) try {
  if (!f || !sum_rgb || samples_done < 0) return fail(SRT_ERR_INVALID, "bad argument");
  CK(cudaSetDevice(f->dev->gpu));
  CK(cudaMemcpyAsync(f->sum, sum_rgb, (size_t)f->W * f->H * 3 * sizeof(float), cudaMemcpyHostToDevice, f->dev->stream));
  CK(cudaStreamSynchronize(f->dev->stream));
  f->done = samples_done;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

}  // extern "C"
