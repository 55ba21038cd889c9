// srt_noalloc.cuh — srt_trace_rays_wavefront_static: srt_trace_rays_wavefront for callers that may not allocate.
//
// No counterpart in the reference.  Some callers cannot call cudaMalloc: code under
// CUDA-graph capture, tight serving loops, harnesses that forbid it.  srt_trace_rays_wavefront allocates its scratch (the pool
// slots the rays are unpacked into, the hit records, the wave's cursors) per call; this variant keeps that scratch in the
// file-scope device array g_static_scratch (srt_b200.cu) and works on DEVICE pointers, so it calls no cudaMalloc / cudaFree / cub
// routine and makes no host copy of the rays.  The kernels, their launch shape and the tree they walk are those of
// srt_trace_rays_wavefront; the only difference is that rays are never sorted before the wave (trees beyond the L2 sort them
// there through cub::DeviceRadixSort, which needs temporary storage) - results do not depend on the order rays are traced in.
//
// Limits and sharing: at most SRT_STATIC_MAX_RAYS rays per call (include/srt.h); every call on a GPU uses the same array, which
// keeps the last call's contents - the cursors are cleared here, every ray slot and hit record of [0, n) is rewritten by
// k_wf_load_rays before the wave reads it; two calls that overlap on one GPU (two host threads, two streams) race on it.
#pragma once

namespace srt {

// (prim, t bits) of the pool's hit records -> the caller's two arrays
__global__ void k_static_unpack_hits(const int2* __restrict__ hit, long long n, int32_t* __restrict__ out_prim, float* __restrict__ out_t) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int2 h = hit[i];
  out_prim[i] = h.x;
  out_t[i] = __int_as_float(h.y);
}

}  // namespace srt

extern "C" {

int64_t srt_static_max_rays(void) { return (int64_t)SRT_STATIC_MAX_RAYS; }

srt_status srt_trace_rays_wavefront_static(srt_scene* s, const srt_ray* rays_dev, int64_t n, int32_t* out_prim_dev, float* out_t_dev) try {
  if (!s || n < 0 || (n > 0 && (!rays_dev || !out_prim_dev || !out_t_dev))) return fail(SRT_ERR_INVALID, "bad argument");
  if (n > (int64_t)SRT_STATIC_MAX_RAYS)
    return fail(SRT_ERR_UNSUPPORTED, "srt_trace_rays_wavefront_static: " + std::to_string(n) + " rays exceed the static scratch (SRT_STATIC_MAX_RAYS = " +
                                         std::to_string((long long)SRT_STATIC_MAX_RAYS) + "); split the batch or call srt_trace_rays_wavefront");
  if (n == 0) return SRT_OK;
  srt_device* dev = s->dev;
  if (!dev->static_scratch) return fail(SRT_ERR_CUDA, "static scratch unavailable on GPU " + std::to_string(dev->gpu) + ": " + dev->static_scratch_err);
  CK(cudaSetDevice(dev->gpu));
  RenderPlan plan;
  int per_sm = 0;
  cudaError_t e = plan_wavefront(s, (long long)n, plan, per_sm);            // attribute / occupancy queries only
  if (e != cudaSuccess) { cudaGetLastError(); return fail(SRT_ERR_CUDA, std::string("occupancy query: ") + cudaGetErrorString(e)); }
  if (per_sm < 1) return fail(SRT_ERR_CUDA, "extend kernel does not fit on an SM");
  const int blocks = dev->sm_count * per_sm;
  // carve the array (the DEVICE address obtained in srt_device_create - never the symbol's host-side name)
  unsigned char* base = static_cast<unsigned char*>(dev->static_scratch);
  WfCursors* cur = reinterpret_cast<WfCursors*>(base);
  WfPool pool{};
  pool.ray_o = reinterpret_cast<float4*>(base + 256);
  pool.ray_d = pool.ray_o + SRT_STATIC_MAX_RAYS;
  pool.hit = reinterpret_cast<int2*>(pool.ray_d + SRT_STATIC_MAX_RAYS);
  pool.n_slots = (int)n; pool.order = nullptr;
  CK(cudaMemsetAsync(cur, 0, sizeof(WfCursors), dev->stream));               // what must start at zero
  CK(cudaMemsetAsync(dev->d_counters, 0, 8 * sizeof(unsigned long long), dev->stream));
  const unsigned nb = (unsigned)((n + 255) / 256);
  k_wf_load_rays<<<nb, 256, 0, dev->stream>>>(rays_dev, (long long)n, pool);
  CK(cudaGetLastError());
  Scene sc = s->sc; unsigned long long* ctr = dev->d_counters;
  if (plan.wf_walk) { sc.nodes = sc.x2_nodes; sc.n_nodes = sc.x2_n; if (sc.x2_cnodes) sc.cnodes = sc.x2_cnodes; }
  int th_walk = plan.wf_th_walk, x2_park = plan.x2_park;
  void* args_v1[] = {&sc, &pool, &cur, &th_walk, &ctr};
  void* args_x2[] = {&sc, &pool, &cur, &th_walk, &x2_park, &ctr};
  CK(cudaLaunchKernel(plan.fn, dim3(blocks), dim3(plan.threads), plan.x2_fmt ? args_x2 : args_v1, plan.smem, dev->stream));
  k_static_unpack_hits<<<nb, 256, 0, dev->stream>>>(pool.hit, (long long)n, out_prim_dev, out_t_dev);
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(dev->stream));
  return SRT_OK;
} catch (const std::bad_alloc&) { return fail(SRT_ERR_NOMEM, "out of memory"); } catch (const std::exception& e_) { return fail(SRT_ERR_INVALID, e_.what()); } catch (...) { return fail(SRT_ERR_INVALID, "unknown exception"); }

}  // extern "C"
