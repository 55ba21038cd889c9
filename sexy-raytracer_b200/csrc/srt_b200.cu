// srt_b200.cu — host side of the C ABI (include/srt.h) + kernel launches.  Built into libsrt_b200.so:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false ...   (see build.py)
// No torch types, no CPU fallback: every compute entry point needs a CUDA device.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <cub/device/device_radix_sort.cuh>

#include "srt_kernels.cuh"
#include "srt_phased.cuh"
#include "srt_wavefront.cuh"
#include "srt_extend2.cuh"
#include "srt_path.cuh"

using namespace srt;

static thread_local std::string g_err;

static srt_status fail(srt_status s, const std::string& msg) { g_err = msg; return s; }
#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      cudaGetLastError();                                                                          \
      return fail(SRT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));               \
    }                                                                                              \
  } while (0)

// ---- scratch of srt_trace_rays_wavefront_static (srt_noalloc.cuh): an entry point that allocates nothing ---------------------
// A file-scope device array: every GPU that loads this library holds SRT_STATIC_SCRATCH_BYTES (2.5 MiB + 256 B) until the process
// exits.  Layout: WfCursors (256 bytes reserved) | ray_o[MAX] float4 | ray_d[MAX] float4 | hit[MAX] int2.
// In HOST code this name is only a stand-in for the symbol, not a device address: srt_device_create asks the runtime for the address
// on each device (cudaGetSymbolAddress with that device current) and keeps it in srt_device::static_scratch.
#define SRT_STATIC_SCRATCH_BYTES (256 + (size_t)SRT_STATIC_MAX_RAYS * (16 + 16 + 8))
static __device__ __align__(256) unsigned char g_static_scratch[SRT_STATIC_SCRATCH_BYTES];

struct srt_device {
  int gpu = 0;
  int sm_count = 0;
  size_t smem_optin = 0;
  size_t l2_bytes = 0;
  cudaStream_t stream = nullptr;       // the stream kernels are launched on
  cudaStream_t own_stream = nullptr;   // created by srt_device_create
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  unsigned long long* d_counters = nullptr;   // 8 x u64, see k_render_mega
  int counters = 0;
  // wavefront variant: path-slot pool (grown on demand), cursors, pinned read-back ring
  WfPool wf{};
  size_t wf_capacity = 0;
  WfCursors* d_wf_cur = nullptr;
  WfCursors* h_wf_ring = nullptr;             // pinned, WF_RING entries
  cudaEvent_t wf_ev[8] = {};
  cudaEvent_t wf_t[16][3] = {};               // per-wave timing ring: before shade, after shade, after extend
  float2* d_grid = nullptr; size_t grid_cap = 0;   // srt_render_params::sampler == 1 (srt_launch.cuh: ensure_grid)
  float* d_frame_sum = nullptr; uchar4* d_frame_rgba = nullptr; size_t frame_cap = 0;   // srt_render's device buffers (pixels)
  unsigned* d_sort = nullptr; size_t sort_cap = 0;  // ray ordering (k_wf_keys): keys in/out, slots in/out, 4 x sort_cap words
  void* d_sort_tmp = nullptr; size_t sort_tmp_bytes = 0;
  void* d_path = nullptr; size_t path_capacity = 0;  // k_wf_path's warp-private path slots (srt_path.cuh)
  void* static_scratch = nullptr;                  // device address of g_static_scratch on THIS gpu (nullptr: see static_scratch_err)
  std::string static_scratch_err;
  int async_alloc = 0;                             // scene arenas through cudaMallocAsync / cudaFreeAsync on `stream`
  void* h_stage = nullptr; size_t h_stage_cap = 0;  // pinned staging of srt_scene_create's arena (scenes up to SCENE_STAGE_MAX bytes)
};
static const int WF_RING = 8;
static const int WF_TRING = 16;

struct srt_scene {
  srt_device* dev = nullptr;
  Scene sc{};
  std::vector<void*> buffers;
  int async_alloc = 0;
  int64_t bytes = 0;
  int nodes_in_smem = 0;
  size_t node_bytes = 0;
  float root_lo[3] = {0, 0, 0}, root_hi[3] = {0, 0, 0};   // node 0's box
};

extern "C" {

int srt_abi_version(void) { return SRT_ABI_VERSION; }
const char* srt_last_error(void) { return g_err.c_str(); }
const char* srt_status_string(srt_status s) {
  switch (s) {
    case SRT_OK: return "ok";
    case SRT_ERR_INVALID: return "invalid argument or scene description";
    case SRT_ERR_CUDA: return "CUDA error";
    case SRT_ERR_UNSUPPORTED: return "unsupported";
    case SRT_ERR_NOMEM: return "out of memory";
    default: return "unknown status";
  }
}

srt_status srt_device_create(int gpu_id, srt_device** out) try {
  if (!out) return fail(SRT_ERR_INVALID, "out is NULL");
  *out = nullptr;
  int n = 0;
  CK(cudaGetDeviceCount(&n));
  if (gpu_id < 0 || gpu_id >= n) return fail(SRT_ERR_INVALID, "gpu_id out of range");
  CK(cudaSetDevice(gpu_id));
  srt_device* d = new srt_device;
  d->gpu = gpu_id;
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, gpu_id));
  d->sm_count = prop.multiProcessorCount;
  d->smem_optin = prop.sharedMemPerBlockOptin;
  d->l2_bytes = (size_t)prop.l2CacheSize;
  CK(cudaStreamCreateWithFlags(&d->own_stream, cudaStreamNonBlocking));
  d->stream = d->own_stream;
  CK(cudaEventCreate(&d->ev0));
  CK(cudaEventCreate(&d->ev1));
  CK(cudaMalloc(&d->d_counters, 8 * sizeof(unsigned long long)));
  {  // scene arenas come from the device's stream-ordered pool and go back to it (cudaMallocAsync / cudaFreeAsync): a scene per
     // frame - the reference's flow - then costs no cudaFree, which took 1-200 ms on this pool's boxes.  Keep what was freed.
    cudaMemPool_t pool = nullptr;
    if (cudaDeviceGetDefaultMemPool(&pool, gpu_id) == cudaSuccess && pool) {
      uint64_t keep = UINT64_MAX;
      if (cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep) == cudaSuccess) d->async_alloc = 1;
    }
    cudaGetLastError();
  }
  {  // the address of the static scratch array on this device (this device is current); a failure only disables
     // srt_trace_rays_wavefront_static, which then reports it
    void* p = nullptr;
    const cudaError_t es = cudaGetSymbolAddress(&p, g_static_scratch);
    if (es == cudaSuccess && p) d->static_scratch = p;
    else { d->static_scratch_err = std::string("cudaGetSymbolAddress(g_static_scratch): ") + cudaGetErrorString(es); cudaGetLastError(); }
  }
  *out = d;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

void srt_device_destroy(srt_device* d) {
  if (!d) return;
  cudaSetDevice(d->gpu);
  if (d->d_counters) cudaFree(d->d_counters);
  if (d->d_grid) cudaFree(d->d_grid);
  if (d->d_sort) cudaFree(d->d_sort);
  if (d->d_frame_sum) cudaFree(d->d_frame_sum);
  if (d->d_frame_rgba) cudaFree(d->d_frame_rgba);
  if (d->d_sort_tmp) cudaFree(d->d_sort_tmp);
  if (d->wf.ray_o) cudaFree(d->wf.ray_o);
  if (d->d_wf_cur) cudaFree(d->d_wf_cur);
  if (d->d_path) cudaFree(d->d_path);
  if (d->h_wf_ring) cudaFreeHost(d->h_wf_ring);
  if (d->h_stage) cudaFreeHost(d->h_stage);
  for (int k = 0; k < WF_RING; k++) if (d->wf_ev[k]) cudaEventDestroy(d->wf_ev[k]);
  for (int k = 0; k < WF_TRING; k++) for (int q = 0; q < 3; q++) if (d->wf_t[k][q]) cudaEventDestroy(d->wf_t[k][q]);
  if (d->ev0) cudaEventDestroy(d->ev0);
  if (d->ev1) cudaEventDestroy(d->ev1);
  if (d->own_stream) cudaStreamDestroy(d->own_stream);
  delete d;
}

srt_status srt_set_counters(srt_device* d, int enabled) try {
  if (!d) return fail(SRT_ERR_INVALID, "device is NULL");
  d->counters = enabled ? 1 : 0;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }
void* srt_device_stream(srt_device* d) { return d ? (void*)d->stream : nullptr; }
srt_status srt_device_set_stream(srt_device* d, void* cuda_stream) try {
  if (!d) return fail(SRT_ERR_INVALID, "device is NULL");
  d->stream = cuda_stream ? (cudaStream_t)cuda_stream : d->own_stream;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

}  // extern "C"

// ---- scene upload ---------------------------------------------------------------------------------
static inline float i2f(int32_t v) { float f; memcpy(&f, &v, 4); return f; }

// Threads the caller's tree (srt_bvh_node, root 0) into the skip-link form of srt_core.cuh.
static srt_status thread_bvh(const srt_scene_desc* d, std::vector<Node>& out, bool& filter_ok) {
  const int64_t n = d->n_nodes, n_prim = d->n_tri + d->n_sph;
  out.clear();
  filter_ok = true;
  if (n == 0) return SRT_OK;
  std::vector<uint8_t> seen((size_t)n, 0);
  struct Frame { int32_t node; int stage; int32_t out_idx; };
  std::vector<Frame> st;
  auto leaf_prim = [&](int32_t child, int32_t& prim) -> bool { prim = ~child; return prim >= 0 && prim < n_prim; };
  auto push_node = [&](int32_t i) -> srt_status {
    if (i < 0 || i >= n) return fail(SRT_ERR_INVALID, "BVH child index out of range");
    if (seen[i]) return fail(SRT_ERR_INVALID, "BVH node reachable twice (not a tree)");
    seen[i] = 1;
    const srt_bvh_node& s = d->nodes[i];
    Node nd;
    nd.lo = make_float4(s.bmin[0], s.bmin[1], s.bmin[2], s.bmax[0]);
    nd.hi = make_float4(s.bmax[1], s.bmax[2], 0.f, 0.f);
    for (int k = 0; k < 3; k++)
      if (!(std::fabs(s.bmin[k]) <= 1.0995116e12f) || !(std::fabs(s.bmax[k]) <= 1.0995116e12f)) filter_ok = false;
    st.push_back({i, 0, (int32_t)out.size()});
    out.push_back(nd);
    return SRT_OK;
  };
  srt_status rc = push_node(0);
  if (rc) return rc;
  while (!st.empty()) {
    Frame f = st.back();
    const srt_bvh_node& s = d->nodes[f.node];
    const bool none = s.right == SRT_BVH_NONE;
    if (f.stage == 0) {
      int32_t pa, pb;
      if (s.left < 0 && (none || s.right < 0)) {            // leaf node: one or two primitives
        if (!leaf_prim(s.left, pa) || (!none && !leaf_prim(s.right, pb))) return fail(SRT_ERR_INVALID, "BVH primitive id out of range");
        out[f.out_idx].hi.z = i2f(pa);
        out[f.out_idx].hi.w = i2f(none ? NODE_NO_B : pb);
        st.pop_back();
        continue;
      }
      if (none) return fail(SRT_ERR_UNSUPPORTED, "SRT_BVH_NONE needs a primitive as the left child");
      st.back().stage = 1;
      if (s.left >= 0) { rc = push_node(s.left); if (rc) return rc; }
      else {                                                 // primitive hanging under an inner node
        if (!leaf_prim(s.left, pa)) return fail(SRT_ERR_INVALID, "BVH primitive id out of range");
        Node p; p.lo = make_float4(0, 0, 0, 0); p.hi = make_float4(0, 0, i2f(pa), i2f(NODE_NO_BOX)); out.push_back(p);
      }
    } else if (f.stage == 1) {
      st.back().stage = 2;
      int32_t pb;
      if (s.right >= 0) { rc = push_node(s.right); if (rc) return rc; }
      else {
        if (!leaf_prim(s.right, pb)) return fail(SRT_ERR_INVALID, "BVH primitive id out of range");
        Node p; p.lo = make_float4(0, 0, 0, 0); p.hi = make_float4(0, 0, i2f(pb), i2f(NODE_NO_BOX)); out.push_back(p);
      }
    } else {
      out[f.out_idx].hi.z = i2f((int32_t)out.size());        // skip link: first node after this subtree
      out[f.out_idx].hi.w = i2f(NODE_INNER);
      st.pop_back();
    }
  }
  if (out.size() > (size_t)INT32_MAX) return fail(SRT_ERR_UNSUPPORTED, "BVH too large");
  return SRT_OK;
}

// The device side of a scene is ONE allocation: every table is laid out in it at a 256-byte boundary, gathered in a pinned
// staging buffer of the device object and copied with one cudaMemcpyAsync; the set-up kernels follow on the same stream and
// srt_scene_create synchronises once, at its end.  (A dozen cudaMalloc + blocking copies cost 4.5 ms per scene, and as many
// cudaFree calls at destroy - per frame in the reference's flow, glDevice::rtFrame.)  Scenes beyond SCENE_STAGE_MAX bytes
// copy table by table from the caller's pageable memory instead of doubling their size in pinned memory.
static const size_t SCENE_STAGE_MAX = (size_t)64 << 20;
struct SceneArena {
  struct Item { const void* src; size_t nb, off; const void** dst; };
  std::vector<Item> items;
  size_t total = 0;
  template <typename T>
  void add(const std::vector<T>& v, const T** dptr) { reserve(v.data(), v.size() * sizeof(T), reinterpret_cast<const void**>(dptr)); }
  size_t reserve(const void* src, size_t nb, const void** dptr) {
    items.push_back({src, nb, total, dptr});
    total += (std::max<size_t>(nb, 16) + 255) & ~(size_t)255;
    return items.size() - 1;
  }
  char* base = nullptr;
  // one allocation; every item's device address is known afterwards (tables that hold device pointers are patched by the
  // caller between alloc and upload)
  srt_status alloc(srt_scene* s) {
    void* b = nullptr;
    cudaError_t e = s->dev->async_alloc ? cudaMallocAsync(&b, total ? total : 256, s->dev->stream) : cudaMalloc(&b, total ? total : 256);
    s->async_alloc = s->dev->async_alloc;
    if (e != cudaSuccess) { cudaGetLastError(); return fail(e == cudaErrorMemoryAllocation ? SRT_ERR_NOMEM : SRT_ERR_CUDA, cudaGetErrorString(e)); }
    s->buffers.push_back(b);
    base = static_cast<char*>(b);
    for (const Item& it : items) {
      if (it.dst) *it.dst = base + it.off;
      s->bytes += (int64_t)it.nb;
    }
    return SRT_OK;
  }
  char* ptr(size_t item) const { return base + items[item].off; }
  // enqueue the copies on the device's stream; the sources must stay alive until the stream is synchronised
  srt_status upload(srt_scene* s) {
    srt_device* dev = s->dev;
    const bool staged = total <= SCENE_STAGE_MAX;
    if (staged && dev->h_stage_cap < total) {
      if (dev->h_stage) { CK(cudaFreeHost(dev->h_stage)); dev->h_stage = nullptr; dev->h_stage_cap = 0; }
      const size_t cap = std::max<size_t>(total, (size_t)8 << 20);
      CK(cudaMallocHost(&dev->h_stage, cap));
      dev->h_stage_cap = cap;
    }
    // staged: one copy covering the items that have a source (reserved-only items at the end of the arena are left out)
    size_t end = 0;
    for (const Item& it : items) {
      if (!it.src || !it.nb) continue;
      if (staged) { memcpy(static_cast<char*>(dev->h_stage) + it.off, it.src, it.nb); end = std::max(end, it.off + it.nb); }
      else CK(cudaMemcpyAsync(base + it.off, it.src, it.nb, cudaMemcpyHostToDevice, dev->stream));
    }
    if (staged && end) CK(cudaMemcpyAsync(base, dev->h_stage, end, cudaMemcpyHostToDevice, dev->stream));
    return SRT_OK;
  }
};

static bool tex_is_image(const srt_scene_desc* d, int id) {
  if (id < 0) return false;
  const srt_texture& t = d->textures[id];
  if (t.kind == SRT_TEX_IMAGE) return t.data != nullptr;
  if (t.kind == SRT_TEX_CHECKER) return tex_is_image(d, t.even) || tex_is_image(d, t.odd);
  return false;
}

extern "C" {

void srt_scene_destroy(srt_scene* s) {
  if (!s) return;
  cudaSetDevice(s->dev->gpu);
  // (every call that used the scene has synchronised its stream before it returned)
  for (auto p : s->buffers) { if (s->async_alloc) cudaFreeAsync(p, s->dev->stream); else cudaFree(p); }
  delete s;
}

int64_t srt_scene_device_bytes(const srt_scene* s) { return s ? s->bytes : 0; }

srt_status srt_scene_create(srt_device* dev, const srt_scene_desc* d, srt_scene** out) try {
  if (!dev || !d || !out) return fail(SRT_ERR_INVALID, "NULL argument");
  *out = nullptr;
  if (d->n_tri < 0 || d->n_sph < 0 || d->n_nodes < 0 || d->n_mat < 0 || d->n_tex < 0) return fail(SRT_ERR_INVALID, "negative count");
  if (d->n_tri + d->n_sph > (int64_t)INT32_MAX || d->n_nodes > (int64_t)INT32_MAX) return fail(SRT_ERR_UNSUPPORTED, "scene too large");
  CK(cudaSetDevice(dev->gpu));

  // tables first: primitives are validated against them
  for (int i = 0; i < d->n_tex; i++) {
    const srt_texture& t = d->textures[i];
    if (t.kind == SRT_TEX_CHECKER) {
      for (int c : {t.even, t.odd})
        if (c < 0 || c >= d->n_tex || d->textures[c].kind == SRT_TEX_CHECKER) return fail(SRT_ERR_UNSUPPORTED, "checker children must be solid or image textures");
    } else if (t.kind == SRT_TEX_IMAGE) {
      if (t.data && (t.width <= 0 || t.height <= 0 || t.bpp <= 0)) return fail(SRT_ERR_INVALID, "image texture with non-positive size");
    } else if (t.kind != SRT_TEX_SOLID) return fail(SRT_ERR_INVALID, "unknown texture kind");
  }
  for (int i = 0; i < d->n_mat; i++) {
    const srt_material& m = d->materials[i];
    if (m.kind < SRT_MAT_PBR || m.kind > SRT_MAT_LIGHT) return fail(SRT_ERR_INVALID, "unknown material kind");
    for (int t : {m.albedo_tex, m.normal_tex, m.metallic_tex, m.roughness_tex})
      if (t >= d->n_tex) return fail(SRT_ERR_INVALID, "material texture id out of range");
    if (m.kind == SRT_MAT_LIGHT && m.albedo_tex < 0) return fail(SRT_ERR_INVALID, "light without an emit texture");
  }
  for (int64_t i = 0; i < d->n_tri; i++) if (d->tri_mat[i] < 0 || d->tri_mat[i] >= d->n_mat) return fail(SRT_ERR_INVALID, "triangle material id out of range");
  for (int64_t i = 0; i < d->n_sph; i++) if (d->sph_mat[i] < 0 || d->sph_mat[i] >= d->n_mat) return fail(SRT_ERR_INVALID, "sphere material id out of range");

  std::vector<Node> nodes;
  bool filter_ok = true;
  srt_status rc = thread_bvh(d, nodes, filter_ok);
  if (rc) return rc;

  srt_scene* s = new srt_scene;
  s->dev = dev;
#define SCK(expr) do { srt_status rc_ = (expr); if (rc_) { srt_scene_destroy(s); return rc_; } } while (0)
  const int n_real = (int)nodes.size();
  {  // sentinel at index n (not counted in n_nodes): a box-less leaf without primitives, see walk_step
    Node t; t.lo = make_float4(0, 0, 0, 0); t.hi = make_float4(0, 0, i2f(-1), i2f(NODE_NO_BOX));
    for (int k = 0; k < 4; k++) nodes.push_back(t);      // + 3 of padding: the walk may prefetch up to 3 nodes ahead
  }
  SceneArena arena;
  arena.add(nodes, &s->sc.nodes);
  s->sc.n_nodes = n_real;
  s->sc.filter_ok = filter_ok ? 1 : 0;
  s->node_bytes = (size_t)(n_real + 1) * sizeof(Node);
  if (n_real > 0) {
    const Node& r0 = nodes[0];
    s->root_lo[0] = r0.lo.x; s->root_lo[1] = r0.lo.y; s->root_lo[2] = r0.lo.z;
    s->root_hi[0] = r0.lo.w; s->root_hi[1] = r0.hi.x; s->root_hi[2] = r0.hi.y;
  }
  s->nodes_in_smem = s->node_bytes > 0 && s->node_bytes <= dev->smem_optin;
  {  // k_wf_extend2's preconditions (srt_extend2.cuh): nested boxes, no box-less primitives, coordinates within 2^40
    bool ok = filter_ok && n_real > 0;
    for (int64_t i = 0; ok && i < d->n_nodes; i++) {
      const srt_bvh_node& p = d->nodes[i];
      const bool leaf = p.left < 0 && (p.right == SRT_BVH_NONE || p.right < 0);
      if (leaf) continue;
      for (int32_t c : {p.left, p.right}) {
        if (c < 0) { ok = false; break; }            // a primitive hanging directly under an inner node (NODE_NO_BOX)
        const srt_bvh_node& q = d->nodes[c];
        for (int k = 0; k < 3; k++) if (!(q.bmin[k] >= p.bmin[k]) || !(q.bmax[k] <= p.bmax[k])) ok = false;
      }
    }
    s->sc.x2_ok = ok ? 1 : 0;
    std::vector<float> cx[3];
    for (int i = 0; i < n_real; i++) {
      int32_t w7; memcpy(&w7, &nodes[i].hi.w, 4);
      if (w7 == NODE_INNER || w7 == NODE_NO_BOX) continue;
      cx[0].push_back(0.5f * (nodes[i].lo.x + nodes[i].lo.w)); cx[1].push_back(0.5f * (nodes[i].lo.y + nodes[i].hi.x));
      cx[2].push_back(0.5f * (nodes[i].lo.z + nodes[i].hi.y));
    }
    for (int k = 0; k < 3; k++) {
      float c = 0.0f;
      if (!cx[k].empty()) { std::nth_element(cx[k].begin(), cx[k].begin() + cx[k].size() / 2, cx[k].end()); c = cx[k][cx[k].size() / 2]; }
      s->sc.x2_centre[k] = std::isfinite(c) ? c : 0.0f;
    }
  }
  // The tree k_wf_extend2 walks (Scene::x2_nodes).  Its inner nodes only cull, and they may cull conservatively (DESIGN
  // section 4): what has to be the reference's is the SEQUENCE of leaves and the test at each leaf.  Any hierarchy of boxes over
  // contiguous runs of that sequence is therefore admissible - a run whose union box is certainly missed holds no leaf whose
  // own box could be hit, and skipping it changes no tCur.  So the walk does not use the reference's inner nodes (random axis,
  // median split: bvh.h:55-95) but its own over the same leaves, in two steps:
  //   1. top-down over the leaf sequence, each run split where  area(left) * count(left) + area(right) * count(right)  is
  //      smallest (a surface-area-heuristic sweep along the one order parity allows);
  //   2. an inner node that would be hit most of the times it is tested is left out - its children hang under the kept ancestor.
  //      Leaving it out saves its own test and costs the tests of children it would have culled: break-even at a hit
  //      probability of 1/2, estimated by the area ratio to the kept ancestor (threshold `tau`, 0.6 measured best).  The root
  //      goes too (rays start inside it).
  // Config 3/4, box tests per ray (CPU model over primary + secondary rays, profiles/r2_walk_tree.md): reference tree 55.3,
  // step 2 alone on the reference's inner nodes 39.2, step 1 alone 32.5, both 26.5.  Leaves are copied unchanged.
  std::vector<Node> x2nodes;
  s->sc.x2_nodes = nullptr; s->sc.x2_n = 0; s->sc.x2_cnodes = nullptr;
  if (s->sc.x2_ok) {
    float tau = 0.6f;
    int rebuild = 1;
    if (const char* t = getenv("SRT_DEBUG_TUNING")) if (!strcmp(t, "1")) {
      if (const char* v = getenv("SRT_X2_TAU")) tau = (float)atof(v);
      if (const char* v = getenv("SRT_X2_SAH")) rebuild = atoi(v);
    }
    auto w7_of = [&](int i) { int32_t w; memcpy(&w, &nodes[i].hi.w, 4); return w; };
    auto w6_of = [&](int i) { int32_t w; memcpy(&w, &nodes[i].hi.z, 4); return w; };
    struct Box { float lo[3], hi[3]; };
    auto box_of = [&](int i) { Box b = {{nodes[i].lo.x, nodes[i].lo.y, nodes[i].lo.z}, {nodes[i].lo.w, nodes[i].hi.x, nodes[i].hi.y}}; return b; };
    auto grow = [](Box& a, const Box& b) { for (int k = 0; k < 3; k++) { a.lo[k] = std::min(a.lo[k], b.lo[k]); a.hi[k] = std::max(a.hi[k], b.hi[k]); } };
    auto area = [](const Box& b) {
      const double dx = std::max(0.0, (double)b.hi[0] - b.lo[0]), dy = std::max(0.0, (double)b.hi[1] - b.lo[1]), dz = std::max(0.0, (double)b.hi[2] - b.lo[2]);
      return 2.0 * (dx * dy + dy * dz + dz * dx);
    };
    auto inner_node = [&](const Box& b) {
      Node nd; nd.lo = make_float4(b.lo[0], b.lo[1], b.lo[2], b.hi[0]); nd.hi = make_float4(b.hi[1], b.hi[2], 0.f, i2f(NODE_INNER)); return nd;
    };
    if (rebuild) {
      std::vector<int> leaves;                               // the leaf nodes in the reference's order
      for (int i = 0; i < n_real; i++) if (w7_of(i) != NODE_INNER) leaves.push_back(i);
      const int L = (int)leaves.size();
      std::vector<Box> lb((size_t)L), suf((size_t)L);
      for (int k = 0; k < L; k++) lb[k] = box_of(leaves[k]);
      const int exact_len = 4096;   // (sweeping only runs of <= 64 leaves position by position: 0.3 s less per 5.9 M-leaf scene, 5 % slower frames)
      struct Job { int a, b; double anc_sa; int close; };     // close >= 0: "subtree of emitted node `close` ends here"
      std::vector<Job> st;
      st.push_back({0, L, -1.0, -1});
      bool top = true;
      while (!st.empty()) {
        const Job j = st.back(); st.pop_back();
        if (j.close >= 0) { x2nodes[j.close].hi.z = i2f((int32_t)x2nodes.size()); continue; }
        if (j.b - j.a == 1) { x2nodes.push_back(nodes[leaves[j.a]]); top = false; continue; }
        // the sweep: every position of a run of up to 4096 leaves; of a longer run (trees of millions of leaves) 128 evenly
        // spaced positions
        const int len = j.b - j.a;
        double best = std::numeric_limits<double>::infinity();
        int cut = j.a + 1;
        Box all;
        if (len <= exact_len) {
          suf[j.b - 1] = lb[j.b - 1];
          for (int k = j.b - 2; k >= j.a; k--) { suf[k] = lb[k]; grow(suf[k], suf[k + 1]); }
          Box pre = lb[j.a];
          for (int k = j.a + 1; k < j.b; k++) {               // left = [a, k), right = [k, b)
            const double c = area(pre) * (k - j.a) + area(suf[k]) * (j.b - k);
            if (c < best) { best = c; cut = k; }
            grow(pre, lb[k]);
          }
          all = suf[j.a];
        } else {
          constexpr int NBMAX = 128;
          const int NB = std::min(NBMAX, len / 4);
          Box bin[NBMAX], bsuf[NBMAX];
          int edge[NBMAX + 1];
          for (int q = 0; q <= NB; q++) edge[q] = j.a + (int)((long long)len * q / NB);
          for (int q = 0; q < NB; q++) { bin[q] = lb[edge[q]]; for (int k = edge[q] + 1; k < edge[q + 1]; k++) grow(bin[q], lb[k]); }
          bsuf[NB - 1] = bin[NB - 1];
          for (int q = NB - 2; q >= 0; q--) { bsuf[q] = bin[q]; grow(bsuf[q], bsuf[q + 1]); }
          Box pre = bin[0];
          for (int q = 1; q < NB; q++) {
            const double c = area(pre) * (edge[q] - j.a) + area(bsuf[q]) * (j.b - edge[q]);
            if (c < best) { best = c; cut = edge[q]; }
            grow(pre, bin[q]);
          }
          all = bsuf[0];
        }
        const double sa = area(all);
        const bool drop = tau < 1.0f && (top || (j.anc_sa > 0.0 && std::isfinite(sa) && sa / j.anc_sa > (double)tau));
        double child_anc = j.anc_sa;
        if (!drop) {
          const int pos = (int)x2nodes.size();
          x2nodes.push_back(inner_node(all));
          st.push_back({0, 0, 0.0, pos});
          child_anc = sa;
        }
        top = false;
        st.push_back({cut, j.b, child_anc, -1});
        st.push_back({j.a, cut, child_anc, -1});
      }
    } else {                                                  // step 2 on the reference's own inner nodes (A/B runs)
      struct Kept { int end; double sa; };
      std::vector<Kept> anc;
      std::vector<int> skip_orig;
      std::vector<int> first_at(n_real + 1, -1);
      for (int i = 0; i < n_real; i++) {
        while (!anc.empty() && i >= anc.back().end) anc.pop_back();
        const bool inner = w7_of(i) == NODE_INNER;
        if (inner) {
          const double sa = area(box_of(i));
          const bool drop = tau < 1.0f && n_real > 1 &&
                            (anc.empty() ? i == 0 : (anc.back().sa > 0.0 && std::isfinite(sa) && sa / anc.back().sa > (double)tau));
          if (drop) continue;
          anc.push_back({w6_of(i), sa});
        }
        first_at[i] = (int)x2nodes.size();
        x2nodes.push_back(nodes[i]);
        skip_orig.push_back(inner ? w6_of(i) : -1);
      }
      first_at[n_real] = (int)x2nodes.size();
      for (int i = n_real - 1; i >= 0; i--) if (first_at[i] < 0) first_at[i] = first_at[i + 1];
      for (size_t k = 0; k < x2nodes.size(); k++)
        if (skip_orig[k] >= 0) x2nodes[k].hi.z = i2f(first_at[std::min(skip_orig[k], n_real)]);
    }
    s->sc.x2_n = (int)x2nodes.size();
    {  // the same sentinel + padding as `nodes` carries (k_wf_extend walks this tree too when it lives in HBM)
      Node t; t.lo = make_float4(0, 0, 0, 0); t.hi = make_float4(0, 0, i2f(-1), i2f(NODE_NO_BOX));
      for (int k = 0; k < 4; k++) x2nodes.push_back(t);
    }
    arena.add(x2nodes, &s->sc.x2_nodes);
  }
  if (const char* t = getenv("SRT_DEBUG_TUNING")) if (!strcmp(t, "1"))
    if (const char* f = getenv("SRT_FORCE_GLOBAL_TREE")) if (atoi(f)) s->nodes_in_smem = 0;   // tests: small scenes through the HBM paths

  // compact tree + leaf-ordered primitives for trees that stay in HBM (srt_core.cuh)
  std::vector<int32_t> lp_src;            // leaf order -> original primitive id
  std::vector<uint4> cn, cn2;
  bool have_cn = false;
  s->sc.cnodes = nullptr; s->sc.lprims = nullptr;
  if (!s->nodes_in_smem && filter_ok && n_real > 0 && d->n_tri + d->n_sph < (int64_t)CNODE_NO_PRIM - 1) {
    const Node& root = nodes[0];
    const float rlo[3] = {root.lo.x, root.lo.y, root.lo.z}, rhi[3] = {root.lo.w, root.hi.x, root.hi.y};
    double org[3], cell[3];
    for (int k = 0; k < 3; k++) {
      // float cell, rounded up so that 65534 cells span the root box; org is the root minimum itself
      float c = (float)(((double)rhi[k] - (double)rlo[k]) / 65534.0);
      c = std::nextafter(std::max(c, 1e-30f), std::numeric_limits<float>::infinity());
      s->sc.q_org[k] = rlo[k]; s->sc.q_cell[k] = c;
      org[k] = rlo[k]; cell[k] = c;
    }
    auto snap = [&](double v, int k, bool up) -> uint32_t {
      double g = (v - org[k]) / cell[k];
      g = up ? std::ceil(g) : std::floor(g);
      if (!(g >= 0.0)) g = 0.0;              // also NaN
      if (g > 65535.0) g = 65535.0;
      return (uint32_t)g;
    };
    bool ok = true;
    auto make_compact = [&](const std::vector<Node>& src, std::vector<uint4>& cn, std::vector<int32_t>& lp_src) {
    cn.resize(src.size());
    for (size_t i = 0; i < src.size(); i++) {
      const Node& nd = src[i];
      int32_t w6, w7; memcpy(&w6, &nd.hi.z, 4); memcpy(&w7, &nd.hi.w, 4);
      uint4 q = make_uint4(0, 0, 0, 0);
      if (w7 != NODE_NO_BOX) {
        const float lo[3] = {nd.lo.x, nd.lo.y, nd.lo.z}, hi[3] = {nd.lo.w, nd.hi.x, nd.hi.y};
        uint32_t L[3], H[3];
        for (int k = 0; k < 3; k++) {
          // a box outside the root box (malformed tree) cannot be snapped conservatively: keep the 32-byte path
          if (!(lo[k] >= rlo[k]) || !(hi[k] <= rhi[k])) ok = false;
          L[k] = snap(lo[k], k, false); H[k] = snap(hi[k], k, true);
        }
        q.x = L[0] | (L[1] << 16); q.y = L[2] | (H[0] << 16); q.z = H[1] | (H[2] << 16);
      }
      if (w7 == NODE_INNER) q.w = CNODE_INNER | (uint32_t)w6;
      else {
        uint32_t w = w7 == NODE_NO_BOX ? CNODE_NO_BOX : 0u;
        if (w6 < 0) w |= CNODE_NO_PRIM;                                        // the sentinel
        else {
          w |= (uint32_t)lp_src.size();
          lp_src.push_back(w6);
          if (w7 >= 0) { w |= CNODE_TWO; lp_src.push_back(w7); }
        }
        q.w = w;
      }
      cn[i] = q;
    }
    };
    make_compact(nodes, cn, lp_src);
    if (ok && !lp_src.empty()) { arena.add(cn, &s->sc.cnodes); have_cn = true; }
    else lp_src.clear();
    if (have_cn && s->sc.x2_ok && !x2nodes.empty()) {     // the walk tree's compact image: same leaves in the same order, so the
      std::vector<int32_t> lp2;                           // leaf words come out as the same indices into lprims
      make_compact(x2nodes, cn2, lp2);
      if (ok && lp2 == lp_src) arena.add(cn2, &s->sc.x2_cnodes);
    }
  }

  std::vector<Tri> tris((size_t)d->n_tri);
  std::vector<TriUV> uvs((size_t)d->n_tri);
  for (int64_t i = 0; i < d->n_tri; i++) {
    const float *a = d->tri_v0 + 3 * i, *b = d->tri_v1 + 3 * i, *c = d->tri_v2 + 3 * i;
    tris[i].a = make_float4(a[0], a[1], a[2], 0); tris[i].b = make_float4(b[0], b[1], b[2], 0);
    tris[i].c = make_float4(c[0], c[1], c[2], 0); tris[i].d = make_float4(0, i2f(d->tri_mat[i]), 0, 0);
    uvs[i].a = make_float4(d->tri_uv0[2 * i], d->tri_uv0[2 * i + 1], d->tri_uv1[2 * i], d->tri_uv1[2 * i + 1]);
    uvs[i].b = make_float4(d->tri_uv2[2 * i], d->tri_uv2[2 * i + 1], 0, 0);
  }
  arena.add(tris, &s->sc.tris);
  arena.add(uvs, &s->sc.tri_uv);
  s->sc.n_tri = (int)d->n_tri;
  s->sc.tri_vn = nullptr;
  std::vector<float4> vn;
  if (d->n_tri > 0 && (d->tri_n0 || d->tri_n1 || d->tri_n2)) {
    if (!d->tri_n0 || !d->tri_n1 || !d->tri_n2) { srt_scene_destroy(s); return fail(SRT_ERR_INVALID, "tri_n0, tri_n1, tri_n2: all three or none"); }
    vn.resize((size_t)d->n_tri * 3);
    for (int64_t i = 0; i < d->n_tri; i++) {
      const float* n[3] = {d->tri_n0 + 3 * i, d->tri_n1 + 3 * i, d->tri_n2 + 3 * i};
      for (int k = 0; k < 3; k++) vn[3 * i + k] = make_float4(n[k][0], n[k][1], n[k][2], 0);
    }
    arena.add(vn, &s->sc.tri_vn);
  }
  s->sc.tri_vt = nullptr;
  std::vector<float4> vt;
  if (d->n_tri > 0 && (d->tri_tan0 || d->tri_tan1 || d->tri_tan2)) {
    if (!d->tri_tan0 || !d->tri_tan1 || !d->tri_tan2) { srt_scene_destroy(s); return fail(SRT_ERR_INVALID, "tri_tan0, tri_tan1, tri_tan2: all three or none"); }
    vt.resize((size_t)d->n_tri * 3);
    for (int64_t i = 0; i < d->n_tri; i++) {
      const float* t[3] = {d->tri_tan0 + 4 * i, d->tri_tan1 + 4 * i, d->tri_tan2 + 4 * i};
      for (int k = 0; k < 3; k++) vt[3 * i + k] = make_float4(t[k][0], t[k][1], t[k][2], t[k][3]);
    }
    arena.add(vt, &s->sc.tri_vt);
  }
  std::vector<Sph> sphs((size_t)d->n_sph);
  for (int64_t j = 0; j < d->n_sph; j++) {
    const float *a = d->sph_c0 + 3 * j, *b = d->sph_c1 + 3 * j;
    bool moving = a[0] != b[0] || a[1] != b[1] || a[2] != b[2];      // center0 != center1, sphere.h:48
    sphs[j].a = make_float4(a[0], a[1], a[2], d->sph_r[j]);
    sphs[j].b = make_float4(b[0], b[1], b[2], d->sph_t0[j]);
    sphs[j].c = make_float4(d->sph_t1[j], i2f(d->sph_mat[j]), moving ? 1.0f : 0.0f, 0);
  }
  arena.add(sphs, &s->sc.sphs);
  s->sc.n_sph = (int)d->n_sph;

  // textures: texel (i,j) = (px[0],px[1],px[2],255), px = data + j*bpp*w + i*bpp; bytes past the end read 0
  std::vector<Tex> texs((size_t)d->n_tex);
  std::vector<std::pair<int, size_t>> images;          // distinct images: (first texture id, arena item of the raw bytes)
  std::vector<int> tex_image((size_t)d->n_tex, -1);    // texture id -> index into `images`
  for (int i = 0; i < d->n_tex; i++) {
    const srt_texture& t = d->textures[i];
    Tex x; memset(&x, 0, sizeof x);
    x.kind = t.kind; x.r = t.color[0]; x.g = t.color[1]; x.b = t.color[2]; x.even = t.even; x.odd = t.odd;
    x.w = t.width; x.h = t.height; x.missing = (t.kind == SRT_TEX_IMAGE && !t.data) ? 1 : 0;
    if (t.kind == SRT_TEX_IMAGE && t.data) {
      // the raw bytes go into the arena once per distinct image (the same PNG behind several materials is common); the
      // device expands them to uchar4 texels (k_expand_texels) - no cudaArray, no texture object, no per-texel host loop
      const size_t n = (size_t)t.width * t.height * t.bpp;
      int u = -1;
      for (size_t q = 0; q < images.size() && u < 0; q++) {
        const srt_texture& o = d->textures[images[q].first];
        if (o.width == t.width && o.height == t.height && o.bpp == t.bpp && (o.data == t.data || memcmp(o.data, t.data, n) == 0)) u = (int)q;
      }
      if (u < 0) { u = (int)images.size(); images.push_back({i, arena.reserve(t.data, n, nullptr)}); }
      tex_image[i] = u;
    }
    texs[i] = x;
  }
  arena.add(texs, &s->sc.texs);
  s->sc.n_tex = d->n_tex;

  std::vector<Mat> mats((size_t)d->n_mat);
  for (int i = 0; i < d->n_mat; i++) {
    const srt_material& m = d->materials[i];
    Mat x; memset(&x, 0, sizeof x);
    x.kind = m.kind; x.albedo_tex = m.albedo_tex; x.normal_tex = m.normal_tex; x.metallic_tex = m.metallic_tex;
    x.roughness_tex = m.roughness_tex; x.metalness = m.metalness; x.roughness = m.roughness;
    for (int k = 0; k < 4; k++) x.albedo[k] = m.albedo[k];
    x.fuzz = m.fuzz; x.ir = m.ir;
    if (m.kind != SRT_MAT_PBR) { x.normal_tex = x.metallic_tex = x.roughness_tex = -1; if (m.kind != SRT_MAT_LIGHT) x.albedo_tex = -1; }
    if (tex_is_image(d, x.albedo_tex) || tex_is_image(d, x.normal_tex) || tex_is_image(d, x.metallic_tex) || tex_is_image(d, x.roughness_tex))
      x.flags |= MAT_NEEDS_UV;
    if (x.normal_tex >= 0) x.flags |= MAT_NEEDS_TBN;
    if (m.kind == SRT_MAT_PBR && (m.flags & SRT_MATF_PACKED_MR) && x.metallic_tex >= 0) { x.flags |= MAT_PACKED_MR; x.roughness_tex = -1; }
    mats[i] = x;
  }
  arena.add(mats, &s->sc.mats);
  s->sc.n_mat = d->n_mat;
  // leaf-ordered copies of the (set-up) primitives for the compact tree: source ids + the output table, both in the arena
  const int32_t* d_src = nullptr;
  const LPrim* d_lp = nullptr;
  if (have_cn) {
    arena.add(lp_src, &d_src);
    arena.reserve(nullptr, lp_src.size() * sizeof(LPrim), reinterpret_cast<const void**>(&d_lp));
  }
  std::vector<size_t> texel_item(images.size());
  for (size_t q = 0; q < images.size(); q++) {           // outputs of k_expand_texels: reserved behind everything that is copied
    const srt_texture& t = d->textures[images[q].first];
    texel_item[q] = arena.reserve(nullptr, (size_t)t.width * t.height * sizeof(uchar4), nullptr);
  }
  SCK(arena.alloc(s));
  for (int i = 0; i < d->n_tex; i++)
    if (tex_image[i] >= 0) texs[i].texels = reinterpret_cast<const uchar4*>(arena.ptr(texel_item[tex_image[i]]));
  SCK(arena.upload(s));
#undef SCK
  for (size_t q = 0; q < images.size(); q++) {
    const srt_texture& t = d->textures[images[q].first];
    const size_t px = (size_t)t.width * t.height;
    k_expand_texels<<<(unsigned)((px + 255) / 256), 256, 0, dev->stream>>>(reinterpret_cast<const unsigned char*>(arena.ptr(images[q].second)),
                                                                          (size_t)t.width * t.height * t.bpp, t.width, t.height, t.bpp,
                                                                          reinterpret_cast<uchar4*>(arena.ptr(texel_item[q])));
  }

  // set-up kernels behind the copy, one synchronisation for the whole scene
  if (d->n_tri > 0) k_setup_tris<<<(unsigned)((d->n_tri + 255) / 256), 256, 0, dev->stream>>>(const_cast<Tri*>(s->sc.tris), (int)d->n_tri);
  if (have_cn) {
    s->sc.lprims = d_lp;
    k_build_lprims<<<(unsigned)((lp_src.size() + 255) / 256), 256, 0, dev->stream>>>(s->sc, d_src, (int)lp_src.size(), const_cast<LPrim*>(d_lp));
  } else {
    s->sc.cnodes = nullptr;
  }
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaStreamSynchronize(dev->stream);
  if (e != cudaSuccess) { cudaGetLastError(); srt_scene_destroy(s); return fail(SRT_ERR_CUDA, std::string("scene set-up: ") + cudaGetErrorString(e)); }
  *out = s;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

}  // extern "C"

#include "srt_launch.cuh"
#include "srt_group.cuh"
#include "srt_noalloc.cuh"
