// tools/tmem_probe.cu -- can Tensor Memory serve as a per-thread constant store?
//
// The 16384-point transform kernels rebuild 31 twiddle powers per radix-32 butterfly from
// one tabulated base (30 complex multiplications) four times per segment pair, because 512
// threads x 127 registers leave no room to keep them.  TMEM (256 KiB per SM, untouched by a
// kernel that issues no tcgen05.mma) is addressable per lane: with the 32x32b shape thread l
// of warp w reaches lane 32*(w%4)+l only, so it cannot exchange data between threads, but it
// can hold 128 private 32-bit words per thread of a 512-thread CTA.  This probe checks
//   1. tcgen05.st / tcgen05.ld round trip of 128 words per thread (512 columns allocated),
//   2. the read rate with all 16 warps loading (x16 / x32 / x64 shapes), in bytes per clock
//      per SM,
//   3. the latency of one dependent ld + wait,
//   4. whether loads steal issue slots from a concurrent FFMA stream.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tmem_probe tools/tmem_probe.cu
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void tmem_alloc(unsigned *dst, int cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst)), "r"(cols));
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
}
__device__ __forceinline__ void tmem_free(unsigned addr, int cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols));
}

#define R16(v, o) "=r"(v[o+0]), "=r"(v[o+1]), "=r"(v[o+2]), "=r"(v[o+3]), "=r"(v[o+4]), "=r"(v[o+5]), "=r"(v[o+6]), "=r"(v[o+7]), \
                  "=r"(v[o+8]), "=r"(v[o+9]), "=r"(v[o+10]), "=r"(v[o+11]), "=r"(v[o+12]), "=r"(v[o+13]), "=r"(v[o+14]), "=r"(v[o+15])
#define S16(v, o) "r"(v[o+0]), "r"(v[o+1]), "r"(v[o+2]), "r"(v[o+3]), "r"(v[o+4]), "r"(v[o+5]), "r"(v[o+6]), "r"(v[o+7]), \
                  "r"(v[o+8]), "r"(v[o+9]), "r"(v[o+10]), "r"(v[o+11]), "r"(v[o+12]), "r"(v[o+13]), "r"(v[o+14]), "r"(v[o+15])

__device__ __forceinline__ void ld16(unsigned (&v)[16], unsigned taddr) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : R16(v, 0) : "r"(taddr));
}
__device__ __forceinline__ void st16(unsigned taddr, const unsigned (&v)[16]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
               ::"r"(taddr), S16(v, 0));
}
__device__ __forceinline__ void ld32(unsigned (&v)[32], unsigned taddr) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
               "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
               : R16(v, 0), R16(v, 16) : "r"(taddr));
}
__device__ __forceinline__ void ld64(unsigned (&v)[64], unsigned taddr) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x64.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
               "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,"
               "%32,%33,%34,%35,%36,%37,%38,%39,%40,%41,%42,%43,%44,%45,%46,%47,"
               "%48,%49,%50,%51,%52,%53,%54,%55,%56,%57,%58,%59,%60,%61,%62,%63}, [%64];"
               : R16(v, 0), R16(v, 16), R16(v, 32), R16(v, 48) : "r"(taddr));
}
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ unsigned word_of(int gtid, int c) { return (unsigned)gtid * 2654435761u + (unsigned)c * 40503u + 17u; }

// thread (warp w, lane l) of a 512-thread CTA owns TMEM lane 32*(w%4)+l, columns 128*(w/4) .. +127
__device__ __forceinline__ unsigned my_taddr(unsigned base) {
  const int w = threadIdx.x >> 5;
  return base + ((unsigned)(32 * (w & 3)) << 16) + (unsigned)(128 * (w >> 2));
}

// MODE 0 round trip check; 1/2/3 read rate with x16/x32/x64; 4 latency (one warp); 5 FFMA alone; 6 FFMA + ld16
template <int MODE>
__global__ void __launch_bounds__(512) probe(unsigned *errors, long long *cycles, float *sink, int iters) {
  __shared__ unsigned tbase;
  const int w = threadIdx.x >> 5;
  if (w == 0) tmem_alloc(&tbase, 512);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const unsigned t0 = my_taddr(tbase);
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  // fill my 128 words
  for (int c0 = 0; c0 < 128; c0 += 16) {
    unsigned v[16];
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = word_of(gtid, c0 + i);
    st16(t0 + c0, v);
  }
  wait_st();
  __syncthreads();
  unsigned acc = 0;
  float facc[16];
#pragma unroll
  for (int i = 0; i < 16; i++) facc[i] = threadIdx.x * 1e-3f + i;
  long long c_begin = clock64();
  if (MODE == 0) {
    unsigned bad = 0;
    for (int c0 = 0; c0 < 128; c0 += 16) {
      unsigned v[16];
      ld16(v, t0 + c0);
      wait_ld();
#pragma unroll
      for (int i = 0; i < 16; i++) bad += v[i] != word_of(gtid, c0 + i);
    }
    if (bad) atomicAdd(errors, bad);
  } else if (MODE == 1) {
    for (int it = 0; it < iters; it++) {
#pragma unroll
      for (int c0 = 0; c0 < 128; c0 += 16) {
        unsigned v[16];
        ld16(v, t0 + c0);
        wait_ld();
#pragma unroll
        for (int i = 0; i < 16; i++) acc ^= v[i];
      }
    }
  } else if (MODE == 2) {
    for (int it = 0; it < iters; it++) {
#pragma unroll
      for (int c0 = 0; c0 < 128; c0 += 32) {
        unsigned v[32];
        ld32(v, t0 + c0);
        wait_ld();
#pragma unroll
        for (int i = 0; i < 32; i++) acc ^= v[i];
      }
    }
  } else if (MODE == 3) {
    for (int it = 0; it < iters; it++) {
#pragma unroll
      for (int c0 = 0; c0 < 128; c0 += 64) {
        unsigned v[64];
        ld64(v, t0 + c0);
        wait_ld();
#pragma unroll
        for (int i = 0; i < 64; i++) acc ^= v[i];
      }
    }
  } else if (MODE == 4) {
    if (w == 0) {
      for (int it = 0; it < iters; it++) {
        unsigned v[16];
        ld16(v, t0 + ((acc & 7u) << 4));        // address depends on the previous load
        wait_ld();
        acc ^= v[0] & 0u;
        acc += v[1] & 1u;
      }
    }
  } else {
    // 128 FFMAs per thread per iteration, with (6) or without (5) one x16 load per iteration
    for (int it = 0; it < iters; it++) {
      unsigned v[16];
      if (MODE == 6) ld16(v, t0 + ((it & 7) << 4));
#pragma unroll
      for (int r = 0; r < 8; r++) {
#pragma unroll
        for (int i = 0; i < 16; i++) facc[i] = fmaf(facc[i], 0.999f, 0.001f);
      }
      if (MODE == 6) {
        wait_ld();
#pragma unroll
        for (int i = 0; i < 16; i++) acc ^= v[i];
      }
    }
  }
  long long c_end = clock64();
  __syncthreads();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += facc[i];
  sink[gtid] = s + (float)acc;
  if (threadIdx.x == 0) cycles[blockIdx.x] = c_end - c_begin;
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (w == 0) tmem_free(tbase, 512);
}

template <int MODE>
void run(const char *name, int iters, unsigned *d_err, long long *d_cyc, float *d_sink, double bytes_per_iter_per_sm) {
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  probe<MODE><<<148, 512>>>(d_err, d_cyc, d_sink, iters);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  probe<MODE><<<148, 512>>>(d_err, d_cyc, d_sink, iters);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  long long cyc[148];
  CK(cudaMemcpy(cyc, d_cyc, sizeof(cyc), cudaMemcpyDeviceToHost));
  long long mx = 0;
  for (int i = 0; i < 148; i++) if (cyc[i] > mx) mx = cyc[i];
  printf("%-34s iters %6d  %.3f ms  thread-0 cycles (max over CTAs) %lld  = %.1f cycles/iter", name, iters, ms, mx,
         (double)mx / iters);
  if (bytes_per_iter_per_sm > 0) printf("  -> %.1f bytes/clk/SM", bytes_per_iter_per_sm * iters / (double)mx);
  printf("\n");
}

int main() {
  unsigned *d_err; long long *d_cyc; float *d_sink;
  CK(cudaMalloc(&d_err, 4)); CK(cudaMemset(d_err, 0, 4));
  CK(cudaMalloc(&d_cyc, 148 * 8)); CK(cudaMalloc(&d_sink, 148 * 512 * 4));
  run<0>("round trip 128 words/thread", 1, d_err, d_cyc, d_sink, 0);
  unsigned err = 0;
  CK(cudaMemcpy(&err, d_err, 4, cudaMemcpyDeviceToHost));
  printf("round-trip mismatches: %u (of %d words)\n", err, 148 * 512 * 128);
  const double all = 512.0 * 128 * 4;           // bytes read per iteration by the 16 warps of an SM
  run<1>("read 128 words/thread as 8 x x16", 2000, d_err, d_cyc, d_sink, all);
  run<2>("read 128 words/thread as 4 x x32", 2000, d_err, d_cyc, d_sink, all);
  run<3>("read 128 words/thread as 2 x x64", 2000, d_err, d_cyc, d_sink, all);
  run<4>("dependent x16 load + wait, one warp", 2000, d_err, d_cyc, d_sink, 0);
  run<5>("128 FFMA/thread/iter, 16 warps", 2000, d_err, d_cyc, d_sink, 0);
  run<6>("same + one x16 load per iteration", 2000, d_err, d_cyc, d_sink, 0);
  printf(err ? "FAIL\n" : "OK\n");
  return err ? 1 : 0;
}
