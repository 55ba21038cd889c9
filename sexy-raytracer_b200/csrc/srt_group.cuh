// srt_group.cuh — several GPUs behind ONE handle (included at the end of srt_b200.cu).
//
// The reference's device boundary is one object that is handed a scene and asked for frames (glDevice::init / rtFrame,
// gl.h:28-31, main.cpp:190-199).  A device group keeps that shape for a whole box: srt_scene_group_create uploads a replica
// of the flattened scene to every GPU, srt_render_group splits the samples of every pixel across them (Philox
// is keyed by the GLOBAL sample index, so the frame does not depend on the split), one host thread per GPU drives its
// render, and the per-GPU fp32 radiance buffers are summed on the first GPU - peer copies over NVLink into a staging
// buffer + one add kernel per peer, 11 MB each at 720p - before writeColorTarget runs there.  There is no collective
// inside the render (paths are independent), so nothing but this one reduction crosses the links.
// One process per GPU with an NCCL reduce (bench.py, sexy_raytracer_b200/dist.py) remains the other route.
#pragma once
#include <thread>

namespace srt {
__global__ void k_add_buffer(float* __restrict__ dst, const float* __restrict__ src, size_t n) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) dst[i] += src[i];
}
}  // namespace srt

struct srt_device_group {
  std::vector<srt_device*> devs;
  std::vector<float*> d_sum;        // per device: its frame sums (grow-only)
  float* d_stage = nullptr;         // on devs[0]: where a peer's buffer lands
  uchar4* d_rgba = nullptr;         // on devs[0]
  size_t cap = 0;                   // pixels the buffers above hold
};
struct srt_scene_group {
  srt_device_group* group = nullptr;
  std::vector<srt_scene*> scenes;
};

namespace {
// global samples [begin, begin + count) -> the part of device k of n (the remainder goes to the low ranks)
inline void group_split(int begin, int count, int n, int k, int& b, int& c) {
  const int base = count / n, rem = count % n;
  c = base + (k < rem ? 1 : 0);
  b = begin + k * base + (k < rem ? k : rem);
}
}  // namespace

extern "C" {

// the share of device k of n of the global samples [begin, begin + count): count / n each, the remainder to the first ones
// (what srt_render_group and sexy_raytracer_b200/dist.py:sample_split both use; no GPU needed)
srt_status srt_sample_split(int32_t begin, int32_t count, int32_t n, int32_t k, int32_t* out_begin, int32_t* out_count) {
  if (begin < 0 || count < 0 || n < 1 || k < 0 || k >= n || !out_begin || !out_count) return fail(SRT_ERR_INVALID, "bad sample split");
  int b, c;
  group_split(begin, count, n, k, b, c);
  *out_begin = b; *out_count = c;
  return SRT_OK;
}

void srt_device_group_destroy(srt_device_group* g) {
  if (!g) return;
  if (!g->devs.empty()) {
    cudaSetDevice(g->devs[0]->gpu);
    if (g->d_stage) cudaFree(g->d_stage);
    if (g->d_rgba) cudaFree(g->d_rgba);
  }
  for (size_t k = 0; k < g->devs.size(); k++) {
    if (k < g->d_sum.size() && g->d_sum[k]) { cudaSetDevice(g->devs[k]->gpu); cudaFree(g->d_sum[k]); }
    srt_device_destroy(g->devs[k]);
  }
  delete g;
}

srt_status srt_device_group_create(const int* gpu_ids, int n_gpus, srt_device_group** out) try {
  if (!out || n_gpus < 0 || (n_gpus > 0 && !gpu_ids)) return fail(SRT_ERR_INVALID, "bad argument");
  *out = nullptr;
  std::vector<int> ids;
  if (n_gpus == 0) {                          // every visible GPU
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    for (int k = 0; k < n; k++) ids.push_back(k);
  } else ids.assign(gpu_ids, gpu_ids + n_gpus);
  if (ids.empty()) return fail(SRT_ERR_CUDA, "no CUDA device");
  srt_device_group* g = new srt_device_group;
  for (int id : ids) {
    srt_device* d = nullptr;
    const srt_status rc = srt_device_create(id, &d);
    if (rc) { srt_device_group_destroy(g); return rc; }
    g->devs.push_back(d);
  }
  g->d_sum.assign(g->devs.size(), nullptr);
  // direct peer copies where the hardware allows (NVLink / NVSwitch); cudaMemcpyPeerAsync falls back to staging otherwise
  cudaSetDevice(g->devs[0]->gpu);
  for (size_t k = 1; k < g->devs.size(); k++) {
    int can = 0;
    if (g->devs[k]->gpu != g->devs[0]->gpu && cudaDeviceCanAccessPeer(&can, g->devs[0]->gpu, g->devs[k]->gpu) == cudaSuccess && can) {
      const cudaError_t e = cudaDeviceEnablePeerAccess(g->devs[k]->gpu, 0);
      if (e != cudaSuccess) cudaGetLastError();          // already enabled, or refused: the copy still works
    }
  }
  *out = g;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

int32_t srt_device_group_size(const srt_device_group* g) { return g ? (int32_t)g->devs.size() : 0; }
srt_device* srt_device_group_device(srt_device_group* g, int32_t k) { return (g && k >= 0 && (size_t)k < g->devs.size()) ? g->devs[k] : nullptr; }

void srt_scene_group_destroy(srt_scene_group* sg) {
  if (!sg) return;
  for (srt_scene* s : sg->scenes) srt_scene_destroy(s);
  delete sg;
}

srt_status srt_scene_group_create(srt_device_group* g, const srt_scene_desc* desc, srt_scene_group** out) try {
  if (!g || !desc || !out) return fail(SRT_ERR_INVALID, "NULL argument");
  *out = nullptr;
  const size_t n = g->devs.size();
  srt_scene_group* sg = new srt_scene_group;
  sg->group = g;
  sg->scenes.assign(n, nullptr);
  std::vector<srt_status> rc(n, SRT_OK);
  std::vector<std::string> err(n);
  auto one = [&](size_t k) { rc[k] = srt_scene_create(g->devs[k], desc, &sg->scenes[k]); if (rc[k]) err[k] = g_err; };
  std::vector<std::thread> th;
  for (size_t k = 1; k < n; k++) th.emplace_back(one, k);          // replicas upload side by side
  one(0);
  for (auto& t : th) t.join();
  for (size_t k = 0; k < n; k++)
    if (rc[k]) { const srt_status r = rc[k]; const std::string m = err[k]; srt_scene_group_destroy(sg); return fail(r, "GPU " + std::to_string(g->devs[k]->gpu) + ": " + m); }
  *out = sg;
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

srt_scene* srt_scene_group_scene(srt_scene_group* sg, int32_t k) { return (sg && k >= 0 && (size_t)k < sg->scenes.size()) ? sg->scenes[k] : nullptr; }

// One frame on all GPUs of the group: glDevice::rtFrame + main.cpp:200-227 for a whole box.  Same outputs as srt_render;
// stats: paths / segments / launches summed, kernel_ms / extend_ms / shade_ms of the slowest GPU.
srt_status srt_render_group(srt_scene_group* sg, const srt_camera* cam, const srt_render_params* P, float* out_sum_rgb,
                            uint8_t* out_rgba8, srt_stats* stats) try {
  if (!sg || !cam || !P) return fail(SRT_ERR_INVALID, "NULL argument");
  srt_device_group* g = sg->group;
  const size_t n = g->devs.size();
  if (P->width < 2 || P->height < 2 || (int64_t)P->width * P->height > (int64_t)INT32_MAX / 4) return fail(SRT_ERR_INVALID, "bad frame size");
  const size_t px = (size_t)P->width * P->height, nf = px * 3;
  srt_device* d0 = g->devs[0];
  if (g->cap < px) {
    for (size_t k = 0; k < n; k++) {
      CK(cudaSetDevice(g->devs[k]->gpu));
      if (g->d_sum[k]) { CK(cudaFree(g->d_sum[k])); g->d_sum[k] = nullptr; }
      CK(cudaMalloc(&g->d_sum[k], nf * sizeof(float)));
    }
    CK(cudaSetDevice(d0->gpu));
    if (g->d_stage) { CK(cudaFree(g->d_stage)); g->d_stage = nullptr; }
    if (g->d_rgba) { CK(cudaFree(g->d_rgba)); g->d_rgba = nullptr; }
    g->cap = 0;
    if (n > 1) CK(cudaMalloc(&g->d_stage, nf * sizeof(float)));
    CK(cudaMalloc(&g->d_rgba, px * 4));
    g->cap = px;
  }
  const int begin = P->sample_begin, count = P->sample_count > 0 ? P->sample_count : P->spp;
  std::vector<srt_status> rc(n, SRT_OK);
  std::vector<std::string> err(n);
  std::vector<srt_stats> st(n);
  auto one = [&](size_t k) {
    srt_render_params q = *P;
    int b, c;
    group_split(begin, count, (int)n, (int)k, b, c);
    q.sample_begin = b; q.sample_count = c;
    memset(&st[k], 0, sizeof st[k]);
    srt_device* d = g->devs[k];
    cudaError_t e = cudaSetDevice(d->gpu);
    if (e == cudaSuccess) e = cudaMemsetAsync(g->d_sum[k], 0, nf * sizeof(float), d->stream);
    if (e != cudaSuccess) { cudaGetLastError(); rc[k] = SRT_ERR_CUDA; err[k] = cudaGetErrorString(e); return; }
    if (c > 0) { rc[k] = srt_render_accumulate(sg->scenes[k], cam, &q, g->d_sum[k], &st[k]); if (rc[k]) err[k] = g_err; }
    else { e = cudaStreamSynchronize(d->stream); if (e != cudaSuccess) { cudaGetLastError(); rc[k] = SRT_ERR_CUDA; err[k] = cudaGetErrorString(e); } }
  };
  std::vector<std::thread> th;
  for (size_t k = 1; k < n; k++) th.emplace_back(one, k);
  one(0);
  for (auto& t : th) t.join();
  for (size_t k = 0; k < n; k++) if (rc[k]) return fail(rc[k], "GPU " + std::to_string(g->devs[k]->gpu) + ": " + err[k]);
  // the one exchange of the frame: every peer's sums -> GPU 0 (each render has synchronised its stream), added in rank order
  CK(cudaSetDevice(d0->gpu));
  for (size_t k = 1; k < n; k++) {
    CK(cudaMemcpyPeerAsync(g->d_stage, d0->gpu, g->d_sum[k], g->devs[k]->gpu, nf * sizeof(float), d0->stream));
    srt::k_add_buffer<<<(unsigned)((nf + 255) / 256), 256, 0, d0->stream>>>(g->d_sum[0], g->d_stage, nf);
    CK(cudaGetLastError());
  }
  int extra = (int)n - 1;
  if (out_rgba8) {
    k_resolve<<<(unsigned)((px + 255) / 256), 256, 0, d0->stream>>>(g->d_sum[0], (int)px, P->spp, g->d_rgba);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out_rgba8, g->d_rgba, px * 4, cudaMemcpyDeviceToHost, d0->stream));
    ++extra;
  }
  if (out_sum_rgb) CK(cudaMemcpyAsync(out_sum_rgb, g->d_sum[0], nf * sizeof(float), cudaMemcpyDeviceToHost, d0->stream));
  CK(cudaStreamSynchronize(d0->stream));
  if (stats) {
    srt_stats t = st[0];
    for (size_t k = 1; k < n; k++) {
      t.paths += st[k].paths; t.segments += st[k].segments; t.node_visits += st[k].node_visits; t.leaf_tests += st[k].leaf_tests;
      t.launches += st[k].launches; t.extend_launches += st[k].extend_launches;
      t.kernel_ms = std::max(t.kernel_ms, st[k].kernel_ms); t.extend_ms = std::max(t.extend_ms, st[k].extend_ms);
      t.shade_ms = std::max(t.shade_ms, st[k].shade_ms);
      if (!t.variant) { t.variant = st[k].variant; t.extend_kernel = st[k].extend_kernel; }
    }
    t.launches += extra;
    *stats = t;
  }
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

}  // extern "C"
