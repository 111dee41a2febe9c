// tools/fp_pipes.cu -- throughput of FADD / FMUL / FFMA / mixes on this GPU (warp-instr per
// clock per SM), to see whether the FFT butterflies' FADDs run at the FFMA rate.
#include <cuda_runtime.h>
#include <stdio.h>
constexpr int ITERS = 4096;
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b,
                                                   unsigned long long c) {
  unsigned long long d;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
// MODE 0: FFMA2 alone; 1: FFMA2 + one integer add per FFMA2; 2: FFMA + one integer add per FFMA
template <int MODE>
__global__ void k2(float *out, float a, float b) {
  unsigned long long acc[16], pa, pb;
  asm("mov.b64 %0, {%1, %2};" : "=l"(pa) : "f"(a), "f"(a));
  asm("mov.b64 %0, {%1, %2};" : "=l"(pb) : "f"(b), "f"(b));
  float facc[16];
  int iacc[16];
#pragma unroll
  for (int i = 0; i < 16; i++) {
    asm("mov.b64 %0, {%1, %2};" : "=l"(acc[i]) : "f"(threadIdx.x * 1e-3f + i), "f"(1.f * i));
    facc[i] = threadIdx.x + i; iacc[i] = threadIdx.x * i;
  }
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) {
      if (MODE <= 1) acc[i] = fma2(acc[i], pa, pb);
      if (MODE == 2) facc[i] = fmaf(facc[i], a, b);
      if (MODE >= 1) asm volatile("add.s32 %0, %0, %1;" : "+r"(iacc[i]) : "r"(it));
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) {
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i]));
    s += lo + hi + facc[i] + (float)iacc[i];
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
double run2(float *out, int blocks, int threads) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 2; i++) k2<MODE><<<blocks, threads>>>(out, 0.999f, 0.001f);
  float best = 1e30f;
  for (int r = 0; r < 5; r++) {
    cudaEventRecord(e0); k2<MODE><<<blocks, threads>>>(out, 0.999f, 0.001f); cudaEventRecord(e1);
    cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  return best;
}
template <int MODE>
__global__ void k(float *out, float a, float b) {
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; i++) acc[i] = threadIdx.x * 1e-3f + i;
  for (int it = 0; it < ITERS; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) {
      if (MODE == 0) acc[i] = fmaf(acc[i], a, b);
      if (MODE == 1) acc[i] = __fadd_rn(acc[i], b);
      if (MODE == 2) acc[i] = __fmul_rn(acc[i], a);
      if (MODE == 3) acc[i] = (i & 1) ? __fadd_rn(acc[i], b) : fmaf(acc[i], a, b);
      if (MODE == 4) acc[i] = __fadd_rn(acc[i], acc[(i + 1) & 15]);      // two varying regs
      if (MODE == 5) acc[i] = fmaf(acc[(i + 1) & 15], a, acc[i]);         // 3 distinct regs
      if (MODE == 6) acc[i] = fmaf(acc[(i + 1) & 15], acc[(i + 2) & 15], acc[i]);
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
double run(float *out, int blocks, int threads) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 2; i++) k<MODE><<<blocks, threads>>>(out, 0.999f, 0.001f);
  float best = 1e30f;
  for (int r = 0; r < 5; r++) {
    cudaEventRecord(e0); k<MODE><<<blocks, threads>>>(out, 0.999f, 0.001f); cudaEventRecord(e1);
    cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  return best;
}
int main() {
  cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
  int sms = prop.multiProcessorCount, blocks = sms * 8, threads = 512;
  float *out; cudaMalloc(&out, sizeof(float) * blocks * threads);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  double winstr = 16.0 * ITERS * blocks * threads / 32.0;
  const char *names[] = {"ffma", "fadd", "fmul", "fadd+ffma", "fadd 2reg", "ffma 3reg(a const)", "ffma 3 varying"};
  double t[7] = {run<0>(out, blocks, threads), run<1>(out, blocks, threads), run<2>(out, blocks, threads),
                 run<3>(out, blocks, threads), run<4>(out, blocks, threads), run<5>(out, blocks, threads),
                 run<6>(out, blocks, threads)};
  for (int m = 0; m < 7; m++)
    printf("%-22s %.3f ms  %.2f warp-instr/clk/SM (at %d MHz)\n", names[m], t[m],
           winstr / (t[m] * 1e-3) / (clk * 1e3) / sms, clk / 1000);
  const char *n2[] = {"ffma2", "ffma2 + iadd", "ffma + iadd"};
  double t2[3] = {run2<0>(out, blocks, threads), run2<1>(out, blocks, threads), run2<2>(out, blocks, threads)};
  for (int m = 0; m < 3; m++)
    printf("%-22s %.3f ms  %.2f FP warp-instr/clk/SM\n", n2[m], t2[m],
           winstr / (t2[m] * 1e-3) / (clk * 1e3) / sms);
  return 0;
}
