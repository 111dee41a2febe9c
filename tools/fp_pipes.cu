// tools/fp_pipes.cu -- throughput of FADD / FMUL / FFMA / mixes on this GPU (warp-instr per
// clock per SM), to see whether the FFT butterflies' FADDs run at the FFMA rate.
#include <cuda_runtime.h>
#include <stdio.h>
constexpr int ITERS = 4096;
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
  return 0;
}
