// fp32_peak.cu — the FP32 ceilings SURVEY 8.4(d) asks for, measured on the box: fused multiply-add (2 flop / instruction)
// and the un-fused mul + add the parity-critical arithmetic is made of (1 flop / instruction), plus FMNMX (the other half
// of a slab test).  Prints one JSON line with TFLOP/s and the SM clock sampled around the kernels.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp32_peak tools/fp32_peak.cu && tools/fp32_peak
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__global__ void __launch_bounds__(1024, 2) k_flops(float* out, int iters, float a, float b) {
  float x[8];
#pragma unroll
  for (int k = 0; k < 8; k++) x[k] = (float)(threadIdx.x + k) * 1e-3f;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int r = 0; r < 8; r++) {
#pragma unroll
      for (int k = 0; k < 8; k++) {
        if (OP == 0) x[k] = __fmaf_rn(x[k], a, b);                      // FFMA
        else if (OP == 1) x[k] = __fadd_rn(__fmul_rn(x[k], a), b);      // FMUL + FADD
        else x[k] = fminf(fmaxf(x[k], a), b + x[(k + 1) & 7]);          // FMNMX + FADD + FMNMX
      }
    }
  }
  float s = 0;
#pragma unroll
  for (int k = 0; k < 8; k++) s += x[k];
  if (s == 123.456f) out[0] = s;
}

template <int OP>
static double run(int sms, float* out, double instr_per_inner) {
  const int iters = 4000;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_flops<OP><<<sms * 2, 1024>>>(out, 100, 1.0001f, 0.0001f);
  cudaEventRecord(e0);
  k_flops<OP><<<sms * 2, 1024>>>(out, iters, 1.0001f, 0.0001f);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  return (double)sms * 2 * 1024 * iters * 64 * instr_per_inner / (ms * 1e-3) / 1e12;   // T instructions/s (per lane)
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  float* out; cudaMalloc(&out, 4);
  const int sms = p.multiProcessorCount;
  for (int rep = 0; rep < 2; rep++) {                       // first round warms the clocks
    const double ffma = run<0>(sms, out, 1), muladd = run<1>(sms, out, 2), mnmx = run<2>(sms, out, 3);
    if (rep == 1)
      printf("{\"gpu\": \"%s\", \"sms\": %d, \"ffma_tinstr_s\": %.2f, \"ffma_tflops\": %.2f, \"unfused_mul_add_tflops\": %.2f, "
             "\"fmnmx_fadd_tinstr_s\": %.2f, \"implied_sm_mhz_ffma\": %.0f}\n",
             p.name, sms, ffma, 2 * ffma, muladd, mnmx, ffma * 1e12 / (sms * 128.0) / 1e6);
  }
  return 0;
}
