// tools/ffma_peak.cu -- measures the FP32 CUDA-core peak of the GPU it runs on (the FP32
// roofline denominator is not in MEASURED_PEAKS.json).  Three variants:
//   ffma      : 3-register FFMA, 16 independent accumulators per thread
//   ffma_imm  : FFMA with an immediate multiplier
//   ffma2     : packed fma.rn.f32x2 (sm_100+), 8 independent pairs per thread
// Prints one JSON line.   build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ffma_peak tools/ffma_peak.cu
#include <cuda_runtime.h>
#include <stdio.h>

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("{\"error\": \"%s\"}\n", cudaGetErrorString(e)); return 1; } } while (0)

constexpr int ITERS = 4096;

__global__ void k_ffma(float *out, float a, float b) {
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; i++) acc[i] = threadIdx.x * 1e-3f + i;
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = fmaf(acc[i], a, b);
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_ffma_imm(float *out) {
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; i++) acc[i] = threadIdx.x * 1e-3f + i;
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) acc[i] = fmaf(acc[i], 0.999f, 0.001f);
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_ffma2(float *out, float a, float b) {
  unsigned long long acc[8], av, bv;
  asm("mov.b64 %0, {%1, %1};" : "=l"(av) : "f"(a));
  asm("mov.b64 %0, {%1, %1};" : "=l"(bv) : "f"(b));
#pragma unroll
  for (int i = 0; i < 8; i++) {
    float lo = threadIdx.x * 1e-3f + i, hi = lo + 0.5f;
    asm("mov.b64 %0, {%1, %2};" : "=l"(acc[i]) : "f"(lo), "f"(hi));
  }
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[i]) : "l"(av), "l"(bv));
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i]));
    s += lo + hi;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <class F>
static double time_ms(F launch) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; i++) launch();
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int rep = 0; rep < 5; rep++) {
    cudaEventRecord(e0);
    launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp prop;
  CHECK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount;
  const int blocks = sms * 8, threads = 512;
  float *out;
  CHECK(cudaMalloc(&out, sizeof(float) * blocks * threads));
  const double flop = 2.0 * 16 * ITERS * (double)blocks * threads;
  double t1 = time_ms([&] { k_ffma<<<blocks, threads>>>(out, 0.999f, 0.001f); });
  double t2 = time_ms([&] { k_ffma_imm<<<blocks, threads>>>(out); });
  double t3 = time_ms([&] { k_ffma2<<<blocks, threads>>>(out, 0.999f, 0.001f); });
  CHECK(cudaDeviceSynchronize());
  int clk = 0;
  cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"max_clock_mhz\": %d, \"ffma_tflops\": %.2f, \"ffma_imm_tflops\": %.2f, "
         "\"ffma2_tflops\": %.2f, \"nominal_tflops\": %.2f}\n",
         prop.name, sms, clk / 1000, flop / t1 * 1e-9, flop / t2 * 1e-9, flop / t3 * 1e-9,
         sms * 128 * 2.0 * clk * 1e-9);
  return 0;
}
