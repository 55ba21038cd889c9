// rand_access_bench.cu — what the B200 memory system gives for the access pattern of the config-5 walk: every lane reads
// ONE 32-byte (or 16-byte) record at an unpredictable index of a large array, and the next index depends on the data.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/rand_access_bench tools/rand_access_bench.cu
//   tools/rand_access_bench    -> table: footprint x record size x chains per lane x threads per SM -> G records/s
// `chains` independent dependent-chains per lane = the memory-level parallelism a lane offers.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

template <int BYTES, int CHAINS>
__global__ void k_chase(const uint4* __restrict__ a, uint32_t mask, int steps, uint32_t* out) {
  uint32_t idx[CHAINS];
  uint32_t acc = 0;
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
#pragma unroll
  for (int c = 0; c < CHAINS; c++) idx[c] = (t * 2654435761u + c * 40503u) & mask;
  for (int s = 0; s < steps; s++) {
#pragma unroll
    for (int c = 0; c < CHAINS; c++) {
      uint32_t v;
      if (BYTES == 32) {
        uint32_t r0, r1, r2, r3, r4, r5, r6, r7;
        asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3), "=r"(r4), "=r"(r5), "=r"(r6), "=r"(r7)
                     : "l"(a + 2 * (size_t)idx[c]));
        v = r0 ^ r7;
      } else {
        const uint4 q = __ldg(a + idx[c]);
        v = q.x ^ q.w;
      }
      acc += v;
      uint32_t z = v + idx[c] * 0x9E3779B9u + t * 0x632BE5ABu + (uint32_t)s;   // depends on the loaded data; t, s keep the chains of different lanes from merging
      z ^= z >> 16; z *= 0x85EBCA6Bu; z ^= z >> 13; z *= 0xC2B2AE35u; z ^= z >> 16;
      idx[c] = z & mask;
    }
  }
  if (acc == 0x12345678u) out[0] = acc;
}

template <int BYTES, int CHAINS>
double run(const uint4* a, size_t records, int threads, int blocks_per_sm, int sms, uint32_t* out) {
  const uint32_t mask = (uint32_t)(records - 1);
  const int steps = 2000 / CHAINS;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_chase<BYTES, CHAINS><<<sms * blocks_per_sm, threads>>>(a, mask, 50, out);
  cudaEventRecord(e0);
  k_chase<BYTES, CHAINS><<<sms * blocks_per_sm, threads>>>(a, mask, steps, out);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  return (double)sms * blocks_per_sm * threads * steps * CHAINS / (ms * 1e-3) / 1e9;
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount;
  const size_t max_bytes = 2ull << 30;
  uint4* a; uint32_t* out;
  cudaMalloc(&a, max_bytes); cudaMalloc(&out, 4);
  {  // pseudo-random contents
    uint32_t* h = (uint32_t*)malloc(max_bytes);
    for (size_t i = 0; i < max_bytes / 4; i++) {
      uint32_t z = (uint32_t)i * 0x9E3779B9u + 0x7F4A7C15u;
      z ^= z >> 16; z *= 0x85EBCA6Bu; z ^= z >> 13; z *= 0xC2B2AE35u; z ^= z >> 16;
      h[i] = z;
    }
    cudaMemcpy(a, h, max_bytes, cudaMemcpyHostToDevice);
    free(h);
  }
  printf("%s, %d SMs, L2 %d MB\n", p.name, sms, p.l2CacheSize >> 20);
  printf("%10s %6s %7s %8s %12s %10s\n", "footprint", "bytes", "chains", "thr/SM", "G records/s", "GB/s");
  const size_t foot[] = {64ull << 20, 128ull << 20, 256ull << 20, 512ull << 20, 1ull << 30, 2ull << 30};
  for (size_t f : foot) {
    for (int tpsm : {1024}) {
      const int bps = tpsm / 1024;
      double g;
      g = run<32, 1>(a, f / 32, 1024, bps, sms, out); printf("%8zu MB %6d %7d %8d %12.1f %10.0f\n", f >> 20, 32, 1, tpsm, g, g * 32);
      g = run<32, 2>(a, f / 32, 1024, bps, sms, out); printf("%8zu MB %6d %7d %8d %12.1f %10.0f\n", f >> 20, 32, 2, tpsm, g, g * 32);
      g = run<32, 4>(a, f / 32, 1024, bps, sms, out); printf("%8zu MB %6d %7d %8d %12.1f %10.0f\n", f >> 20, 32, 4, tpsm, g, g * 32);
      g = run<16, 1>(a, f / 16, 1024, bps, sms, out); printf("%8zu MB %6d %7d %8d %12.1f %10.0f\n", f >> 20, 16, 1, tpsm, g, g * 16);
      g = run<16, 4>(a, f / 16, 1024, bps, sms, out); printf("%8zu MB %6d %7d %8d %12.1f %10.0f\n", f >> 20, 16, 4, tpsm, g, g * 16);
    }
  }
  return 0;
}
