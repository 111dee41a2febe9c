// tools/overlap_probe.cu -- do shared-memory phases and FP phases of different warps overlap?
// Each warp loops over: [32 x LDS.64] -> [FP block] -> [32 x STS.64], warps are independent
// (no CTA barrier).  Variants: FP only, LSU only, both.  One CTA of 512 threads per SM.
#include <cuda_runtime.h>
#include <stdio.h>
constexpr int ROUNDS = 200;
template <bool DO_LSU, bool DO_FP, int FPREP>
__global__ void __launch_bounds__(512, 1) probe(float *out) {
  extern __shared__ float2 sm[];
  const int t = threadIdx.x;
  float2 v[32];
#pragma unroll
  for (int i = 0; i < 32; i++) v[i] = make_float2(t * 1e-3f + i, i * 0.5f - t);
  float2 *base = sm + (t >> 5) * 1088 + (t & 31);      // warp-private 32x34 region
  if (DO_LSU) {
#pragma unroll
    for (int i = 0; i < 32; i++) base[i * 34] = v[i];
  }
  __syncwarp();
  for (int r = 0; r < ROUNDS; r++) {
    if (DO_LSU) {
#pragma unroll
      for (int i = 0; i < 32; i++) {
        float2 q = base[i * 34];
        v[i].x += q.x * 1e-9f; v[i].y += q.y * 1e-9f;     // 2 FP per element keep the loads live
      }
    }
    if (DO_FP) {
#pragma unroll
      for (int rep = 0; rep < FPREP; rep++) {
#pragma unroll
        for (int i = 0; i < 16; i++) {      // butterfly-like: 6 ops on a pair
          float2 a = v[i], b = v[i + 16];
          float nr = fmaf(-b.y, 0.38268343f, fmaf(b.x, 0.92387953f, a.x));
          float ni = fmaf(b.y, 0.92387953f, fmaf(b.x, 0.38268343f, a.y));
          v[i + 16] = make_float2(fmaf(2.0f, a.x, -nr), fmaf(2.0f, a.y, -ni));
          v[i] = make_float2(nr * 0.5f, ni * 0.5f);
        }
      }
    }
    if (DO_LSU) {
#pragma unroll
      for (int i = 0; i < 32; i++) base[i * 34] = v[i];
      __syncwarp();
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 32; i++) s += v[i].x + v[i].y;
  out[blockIdx.x * blockDim.x + t] = s;
}
template <bool L, bool F, int R>
float run(float *out, int blocks) {
  size_t smem = 16 * 1088 * sizeof(float2);
  cudaFuncSetAttribute(probe<L, F, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  probe<L, F, R><<<blocks, 512, smem>>>(out);
  cudaEventRecord(e0); probe<L, F, R><<<blocks, 512, smem>>>(out); cudaEventRecord(e1);
  cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main() {
  cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
  int blocks = prop.multiProcessorCount;
  float *out; cudaMalloc(&out, sizeof(float) * blocks * 512);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  auto cyc = [&](float ms) { return ms * 1e-3 * clk * 1e3 / ROUNDS; };
  float fp = run<false, true, 4>(out, blocks), ls = run<true, false, 4>(out, blocks), both = run<true, true, 4>(out, blocks);
  printf("per round (cycles): FP only (4x96+.. instr/thread) %.0f | LSU only %.0f | both %.0f\n", cyc(fp), cyc(ls), cyc(both));
  printf("expected FP issue cycles/round/SMSP = %d ; LSU wavefront cycles/round/SM = %d\n", 4 * 4 * 128, 16 * 64 * 2);
  float fp8 = run<false, true, 8>(out, blocks), both8 = run<true, true, 8>(out, blocks);
  printf("FPREP=8: FP only %.0f | both %.0f\n", cyc(fp8), cyc(both8));
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(e));
  return 0;
}
