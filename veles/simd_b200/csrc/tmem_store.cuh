// tmem_store.cuh -- Tensor Memory as a per-thread constant store (device code only).
//
// The 16384-point transform needs 31 inter-pass twiddle factors per radix-32 butterfly, two
// sets per thread (first/last pass and the two middle passes).  512 threads x 127 registers
// leave no room to keep them, so the plain kernels rebuild them from one tabulated base with
// 30 complex multiplications, four times per segment pair: 30 % of all FP32 instructions
// (profiles/ncu_c4_fused_r01.txt).  These kernels issue no tcgen05.mma, so the SM's 256 KiB
// of Tensor Memory is idle; it is addressable per lane only (with the 32x32b shape thread l
// of warp w reaches lane 32*(w%4)+l), which rules it out for exchanging data between threads
// but makes it a private scratchpad of 128 words per thread of a 512-thread CTA: exactly
// the two sets of 32 complex factors.  A persistent CTA fills it once and then reads 62
// words per pass at ~470 B/clk/SM (tools/tmem_probe.cu, profiles/tmem_probe_r02.txt).
#pragma once

namespace tmem {

__device__ __forceinline__ unsigned smem_addr(const void *p) {
  return (unsigned)__cvta_generic_to_shared(p);
}

// One warp allocates all 512 columns (the CTA is alone on its SM) and publishes the base
// address through shared memory; the same warp frees them before the CTA retires.
__device__ __forceinline__ void alloc_all(unsigned *base_in_smem) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                   smem_addr(base_in_smem)),
               "r"(512));
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
}
__device__ __forceinline__ void free_all(unsigned base) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(512));
}
__device__ __forceinline__ void fence_before_sync() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void fence_after_sync() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// Address of this thread's 128 private words: lane quadrant of its warp, column block of its
// warp group (warp-uniform, as the instructions require).
__device__ __forceinline__ unsigned thread_base(unsigned base, int thread) {
  const int w = thread >> 5;
  return base + ((unsigned)(32 * (w & 3)) << 16) + (unsigned)(128 * (w >> 2));
}

// Eight complex factors = 16 words = one x16 access.
struct Chunk {
  unsigned r[16];
  __device__ __forceinline__ float2 get(int i) const {
    return make_float2(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1]));
  }
  __device__ __forceinline__ void set(int i, float2 v) {
    r[2 * i] = __float_as_uint(v.x);
    r[2 * i + 1] = __float_as_uint(v.y);
  }
};

#define TM_OUT16(c) "=r"(c.r[0]), "=r"(c.r[1]), "=r"(c.r[2]), "=r"(c.r[3]), "=r"(c.r[4]), "=r"(c.r[5]), \
                    "=r"(c.r[6]), "=r"(c.r[7]), "=r"(c.r[8]), "=r"(c.r[9]), "=r"(c.r[10]), "=r"(c.r[11]), \
                    "=r"(c.r[12]), "=r"(c.r[13]), "=r"(c.r[14]), "=r"(c.r[15])
#define TM_IN16(c) "r"(c.r[0]), "r"(c.r[1]), "r"(c.r[2]), "r"(c.r[3]), "r"(c.r[4]), "r"(c.r[5]), \
                   "r"(c.r[6]), "r"(c.r[7]), "r"(c.r[8]), "r"(c.r[9]), "r"(c.r[10]), "r"(c.r[11]), \
                   "r"(c.r[12]), "r"(c.r[13]), "r"(c.r[14]), "r"(c.r[15])
#define TM_IO16(c) "+r"(c.r[0]), "+r"(c.r[1]), "+r"(c.r[2]), "+r"(c.r[3]), "+r"(c.r[4]), "+r"(c.r[5]), \
                   "+r"(c.r[6]), "+r"(c.r[7]), "+r"(c.r[8]), "+r"(c.r[9]), "+r"(c.r[10]), "+r"(c.r[11]), \
                   "+r"(c.r[12]), "+r"(c.r[13]), "+r"(c.r[14]), "+r"(c.r[15])

// The TMEM instructions are .sync.aligned: every thread of the warp must execute them
// together.  Per-thread guards in the surrounding code (bounds-checked loads and stores of
// signal edges) leave a warp's threads at different program points, and nothing but an
// explicit __syncwarp() brings them back together -- without it the warp hangs.
__device__ __forceinline__ void load(Chunk &c, unsigned addr) {
  __syncwarp();
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : TM_OUT16(c)
      : "r"(addr));
}
__device__ __forceinline__ void store(unsigned addr, const Chunk &c) {
  __syncwarp();
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(addr), TM_IN16(c));
}
// The loads are asynchronous: their destination registers may only be read after this wait.
// Passing the registers through the statement ties every later use to it.
__device__ __forceinline__ void wait_loads(Chunk &a, Chunk &b) {
  __syncwarp();
  asm volatile("tcgen05.wait::ld.sync.aligned;" : TM_IO16(a), TM_IO16(b)::"memory");
}
__device__ __forceinline__ void wait_stores() {
  __syncwarp();
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

}  // namespace tmem
