// FP32 FMA and MUFU (ex2 / lg2 / rcp) issue-rate microbenchmarks for the B200 (SURVEY.md 8(d) "Binding roofline":
// the fills are bound by FP32/SFU issue, not by HBM; these are the peaks that bound is quoted against).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench/peaks tools/microbench/peaks.cu
// Every thread runs 8 independent dependency chains so that the pipes, not latency, are the limit.
#include <cstdio>
#include <cuda_runtime.h>

#define ITER 4096

template <int OP>
__global__ void __launch_bounds__(1024) peak_kernel(float* out, float seed) {
  float a[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) a[k] = seed + 0.001f * (threadIdx.x + k);
  const float m = 0.999f, c = 0.001f;
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (OP == 0) a[k] = fmaf(a[k], m, c);
      else if (OP == 1) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[k]));
      else if (OP == 2) asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(a[k]));
      else asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(a[k]));
    }
  }
  float s = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += a[k];
  if (s == 12345.678f) out[0] = s;   // keeps the chains alive
}

template <int OP>
double run(int sms, int clock_khz, const char* name, double flop_per_op) {
  float* d;
  cudaMalloc(&d, 4);
  const int blocks = sms * 2;
  peak_kernel<OP><<<blocks, 1024>>>(d, 1.0f);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best = 1e30;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0);
    peak_kernel<OP><<<blocks, 1024>>>(d, 1.0f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  const double ops = (double)blocks * 1024 * 8 * ITER;
  const double rate = ops / (best * 1e-3);
  printf("\"%s\": {\"gops\": %.1f, \"per_sm_per_clk_at_max_clock\": %.2f},\n", name, rate / 1e9, rate / sms / (clock_khz * 1e3));
  if (flop_per_op > 1) printf("\"%s_tflops\": %.2f,\n", name, rate * flop_per_op / 1e12);
  cudaFree(d);
  return rate;
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  int clk = 0;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"max_clock_mhz\": %d,\n", p.name, p.multiProcessorCount, clk / 1000);
  run<0>(p.multiProcessorCount, clk, "fp32_fma", 2.0);
  run<1>(p.multiProcessorCount, clk, "mufu_ex2", 1.0);
  run<2>(p.multiProcessorCount, clk, "mufu_lg2", 1.0);

  printf("\"iterations_per_thread\": %d, \"chains_per_thread\": 8, \"threads\": \"2 x 1024 per SM\"}\n", ITER);
  return 0;
}
