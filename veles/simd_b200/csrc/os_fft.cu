// os_fft.cu -- fused overlap-save / FFT convolution kernels.
//
// Replaces convolve_overlap_save (/root/reference/src/convolve.c:156-229) and convolve_fft
// (/root/reference/src/convolve.c:289-326).  The reference walks the segments serially and
// per segment runs: gather -> FFTF forward -> complex multiply with the filter spectrum ->
// FFTF inverse -> scale -> copy.  Here one CTA does all of that for a PAIR of segments
// (packed as real + i*imag of one complex transform, see fft_core.cuh) without the spectrum
// ever leaving the SM: global -> registers -> (shared-memory passes) -> registers -> global.
//
// Kernels:
//   os_spectrum_kernel<P>   H = FFT_N(zero-pad(h or reversed h)) / N, written in the register
//                           order of the fused kernel's centre pass (src/convolve.c:167-176:
//                           the reference also rebuilds H on every call because h is a call
//                           argument).  One CTA per impulse-response partition.
//   os_fused_kernel<P, G>   one CTA = G independent segment pairs of P::N complex points.
//
// Segment length is NOT the reference's L = 4*2^floor(log2 h) (src/convolve.c:116-121); it
// is chosen per impulse-response length from the plans below (b200_os_pick_n).  Impulse
// responses longer than one transform can hold are split into partitions of part_taps taps;
// each partition is one launch accumulating into the result (uniformly partitioned
// overlap-save), which also serves the whole-signal "FFT" algorithm for long inputs.
#include "b200_device.cuh"
#include "fft_core.cuh"
#include <math.h>
#include <mutex>
#include <vector>

namespace {

using namespace fc;

struct OsParams {
  const float *x;
  i64 x_len, x_stride;
  float *out;
  i64 out_len, out_stride;
  const float2 *spectrum;   // this partition's filter spectrum
  const float2 *tw;
  i64 in_shift;             // X index of segment 0's first sample
  int c_off;                // first useful circular sample
  int step;                 // outputs per segment
  i64 nseg;                 // segments per signal
  i64 npairs;               // segment pairs per signal
  i64 total;                // npairs * batch
  int accumulate;
  i64 prefetch_distance;    // work items between a CTA and its successor on the same SM
};

struct SpecParams {
  const float *h;
  i64 h_len;
  int reverse;
  int part_taps;
  float2 *spectrum;         // [parts][N]
  const float2 *tw;
  float scale;
};

// Barrier over the threads that share one transform.  Transforms of up to 32 threads live
// inside a warp (warp-level barrier: groups in one CTA never wait for each other); larger
// groups of a multi-group CTA use a named barrier of their own; a single group is the CTA.
template <class P, int G>
__device__ __forceinline__ void group_sync(int grp) {
  if constexpr (P::T <= 32) {
    __syncwarp();
  } else if constexpr (G > 1) {
    asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "n"(P::T) : "memory");
  } else {
    __syncthreads();
  }
}

// item -> (signal, unit inside the signal); 32-bit division when the numbers allow it (the
// 64-bit one costs more than a radix-4 butterfly network on the small plans)
__device__ __forceinline__ void split_item(i64 item, i64 per_signal, i64 &sig, i64 &unit) {
  if (((item | per_signal) >> 31) == 0) {
    const unsigned q = (unsigned)item / (unsigned)per_signal;
    sig = q;
    unit = (unsigned)item - q * (unsigned)per_signal;
  } else {
    sig = item / per_signal;
    unit = item - sig * per_signal;
  }
}

template <class P>
__global__ void __launch_bounds__(P::T) os_spectrum_kernel(SpecParams p) {
  extern __shared__ float4 smem4[];
  float2 *sm = reinterpret_cast<float2 *>(smem4);
  const int t = threadIdx.x;
  const int part = blockIdx.x;
  // programmatic dependent launch: the fused kernel behind us may start now; it only needs
  // the spectrum at its centre pass and waits there (griddepcontrol.wait)
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  // correlation-form taps of this partition: g[j], j in [j0, j0+taps); its convolution-form
  // impulse response is gf[m] = g[j0+taps-1-m]  (g[j] = reverse ? h[j] : h[h_len-1-j]).
  const i64 j0 = (i64)part * p.part_taps;
  const i64 left = p.h_len - j0;
  const int taps = (int)(left < p.part_taps ? left : p.part_taps);
  i64 start;
  int dir;
  if (p.reverse) { start = j0 + taps - 1; dir = -1; }
  else { start = p.h_len - j0 - taps; dir = 1; }
  TwBases<P> tb;
  load_twiddle_bases<P>(t, p.tw, tb);
  spec_first<P>(t, p.h, start, dir, taps, tb, sm);
  __syncthreads();
  if constexpr (P::NPASS == 3) {
    fwd_mid<P>(t, tb, sm);
    __syncthreads();
  }
  centre_store_spectrum<P>(t, p.spectrum + (size_t)part * P::N, p.scale, sm);
}

template <class P, int G>
__global__ void __launch_bounds__(P::T *G) os_fused_kernel(OsParams p) {
  extern __shared__ float4 smem4[];
  const int grp = threadIdx.x / P::T;
  const int t = threadIdx.x % P::T;
  float2 *sm = reinterpret_cast<float2 *>(smem4) + grp * P::SMEM_FLOAT2;

  const i64 item = (i64)blockIdx.x * G + grp;
  const bool live = item < p.total;     // whole group idles (but still hits the barriers)
  SegPair w;
  TwBases<P> tb;
  HRegs<P> hr;
  if (t == 0) {
    // Pull the input of the CTA that will follow this one on the SM into L2 while this CTA
    // computes (CTAs are dispatched in index order, one wave = p.prefetch_distance items).
    const i64 nxt = item + p.prefetch_distance;
    if (p.prefetch_distance > 0 && nxt < p.total) {
      i64 nb, nsa;
      split_item(nxt, p.npairs, nb, nsa);
      nsa *= 2;
      i64 lo = nsa * p.step + p.in_shift;
      i64 hi = lo + p.step + P::N;
      if (lo < 0) lo = 0;
      if (hi > p.x_len) hi = p.x_len;
      const float *base = p.x + nb * p.x_stride;
      unsigned long long a0 = ((unsigned long long)(base + lo) + 15ull) & ~15ull;
      unsigned long long a1 = (unsigned long long)(base + hi) & ~15ull;
      if (a1 > a0) {
        unsigned bytes = (unsigned)(a1 - a0);
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a0), "r"(bytes) : "memory");
      }
    }
  }
  if (live) {
    load_twiddle_bases<P>(t, p.tw, tb);
    i64 b, sa;
    split_item(item, p.npairs, b, sa);
    sa *= 2;
    w.x = p.x + b * p.x_stride;
    w.x_len = p.x_len;
    w.inA = sa * p.step + p.in_shift;
    w.inB = w.inA + p.step;
    w.haveB = sa + 1 < p.nseg;
    w.out = p.out + b * p.out_stride;
    w.outA = sa * p.step;
    w.outB = w.outA + p.step;
    const i64 remA = p.out_len - w.outA, remB = p.out_len - w.outB;
    w.cntA = (int)(remA < p.step ? remA : p.step);
    w.cntB = w.haveB ? (int)(remB < p.step ? remB : p.step) : 0;
    w.c_off = p.c_off;
    w.accumulate = p.accumulate != 0;
    fwd_first<P>(t, w, tb, sm);
  }
  group_sync<P, G>(grp);
  if constexpr (P::NPASS == 3) {
    // Between the first and the last pass every half-warp works on sub-blocks nobody else
    // touches (PassGeom::locate_mid / locate_centre), so warp-level barriers suffice and the
    // warps of a CTA run these three phases out of step with each other.
    if (live) fwd_mid<P>(t, tb, sm);
    // the spectrum kernel may still be running (programmatic dependent launch): everything
    // above overlapped with it, the spectrum itself is needed from here on
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (live) load_spectrum<P>(t, p.spectrum, hr);
    __syncwarp();
    if (live) centre<P>(t, hr, sm);
    __syncwarp();
    if (live) inv_mid<P>(t, tb, sm);
  } else {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (live) load_spectrum<P>(t, p.spectrum, hr);
    if (live) centre<P>(t, hr, sm);
  }
  group_sync<P, G>(grp);
  if (live) inv_last<P>(t, w, tb, sm);
}

// ---- real-packed mode (fft_core.cuh, "mode B"): one real segment of 2N samples per CTA ------
template <class P>
__global__ void __launch_bounds__(P::T) os_spectrum_real_kernel(SpecParams p) {
  extern __shared__ float4 smem4[];
  float2 *sm = reinterpret_cast<float2 *>(smem4);
  const int t = threadIdx.x;
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  // single partition: convolution-form impulse response gf[m] = reverse ? h[taps-1-m] : h[m]
  const int taps = (int)p.h_len;
  const i64 start = p.reverse ? taps - 1 : 0;
  const int dir = p.reverse ? -1 : 1;
  TwBases<P> tb;
  load_twiddle_bases<P>(t, p.tw, tb);
  spec_first_real<P>(t, p.h, start, dir, taps, tb, sm);
  __syncthreads();
  if constexpr (P::NPASS == 3) {
    fwd_mid<P>(t, tb, sm);
    __syncthreads();
  }
  centre_real_fwd<P>(t, sm);
  __syncthreads();
  centre_real_coeffs<P>(t, reinterpret_cast<float4 *>(p.spectrum), p.scale, sm);
}

template <class P, int G>
__global__ void __launch_bounds__(P::T *G) os_fused_real_kernel(OsParams p) {
  extern __shared__ float4 smem4[];
  const int grp = threadIdx.x / P::T;
  const int t = threadIdx.x % P::T;
  float2 *sm = reinterpret_cast<float2 *>(smem4) + grp * P::SMEM_FLOAT2;

  const i64 item = (i64)blockIdx.x * G + grp;
  const bool live = item < p.total;
  SegReal w;
  TwBases<P> tb;
  float2 u[P::E];
  if (t == 0) {
    const i64 nxt = item + p.prefetch_distance;
    if (p.prefetch_distance > 0 && nxt < p.total) {
      i64 nb, nsg;
      split_item(nxt, p.nseg, nb, nsg);
      i64 lo = nsg * p.step + p.in_shift;
      i64 hi = lo + 2 * P::N;
      if (lo < 0) lo = 0;
      if (hi > p.x_len) hi = p.x_len;
      const float *base = p.x + nb * p.x_stride;
      unsigned long long a0 = ((unsigned long long)(base + lo) + 15ull) & ~15ull;
      unsigned long long a1 = (unsigned long long)(base + hi) & ~15ull;
      if (a1 > a0) {
        unsigned bytes = (unsigned)(a1 - a0);
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a0), "r"(bytes) : "memory");
      }
    }
  }
  if (live) {
    load_twiddle_bases<P>(t, p.tw, tb);
    i64 b, sg;
    split_item(item, p.nseg, b, sg);
    w.x = p.x + b * p.x_stride;
    w.x_len = p.x_len;
    w.in0 = sg * p.step + p.in_shift;
    w.out = p.out + b * p.out_stride;
    w.out0 = sg * p.step;
    const i64 rem = p.out_len - w.out0;
    w.cnt = (int)(rem < p.step ? rem : p.step);
    w.c_off = p.c_off;
    w.accumulate = p.accumulate != 0;
    fwd_first_real<P>(t, w, tb, sm);
  }
  group_sync<P, G>(grp);
  if constexpr (P::NPASS == 3) {
    if (live) fwd_mid<P>(t, tb, sm);
    __syncwarp();                       // middle passes stay inside the half-warp
  }
  if (live) centre_real_fwd<P>(t, sm);
  asm volatile("griddepcontrol.wait;" ::: "memory");
  group_sync<P, G>(grp);                // mirror bins live in other threads of the group
  if (live) centre_real_mix<P>(t, reinterpret_cast<const float4 *>(p.spectrum), sm, u);
  group_sync<P, G>(grp);                // every mirror has been read: safe to overwrite
  if (live) centre_real_store<P>(t, sm, u);
  if constexpr (P::NPASS == 3) {
    __syncwarp();
    if (live) inv_mid<P>(t, tb, sm);
  }
  group_sync<P, G>(grp);
  if (live) inv_last_real<P>(t, w, tb, sm);
}

// ---- persistent variant of the paired kernel for the one-CTA-per-SM plans ------------------
// One CTA per SM loops over segment pairs.  While a pair is being transformed, a bulk
// asynchronous copy (cp.async.bulk / TMA, completion on an mbarrier) brings the next pair's
// input range into the shared memory left over beside the transform buffer, so the next
// pair's first pass starts from shared memory instead of waiting for HBM, and the current
// pair's result stores drain while that first pass computes.  Needs 16-byte aligned segment
// starts (the host picks step and leading-sample count accordingly) -- otherwise, and for
// pairs that stick out of the signal, the plain kernel's guarded loads are used.
constexpr int kParkFloats = 20992;      // 82 KiB beside the 136 KiB transform buffer (+ 8 KiB twiddle bases)

__device__ __forceinline__ unsigned smem_u32(const void *p) {
  return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_load(void *dst, const void *src, unsigned bytes,
                                          unsigned long long *bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
  unsigned ok;
  do {
    asm volatile(
        "{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!ok);
}

template <class P>
__global__ void __launch_bounds__(P::T) os_fused_persist_kernel(OsParams p) {
  extern __shared__ float4 smem4[];
  float2 *sm = reinterpret_cast<float2 *>(smem4);
  float *park = reinterpret_cast<float *>(sm + P::SMEM_FLOAT2);
  TwBases<P> *tb_sm = reinterpret_cast<TwBases<P> *>(park + kParkFloats);   // one per thread
  unsigned long long *bar = reinterpret_cast<unsigned long long *>(tb_sm + P::T);
  const int t = threadIdx.x;

  // parked range of a pair: x[inA, inA + park_len), clipped to what the pair reads
  const int need = P::N + p.step;                       // A followed by B's tail
  const int park_len = need < kParkFloats ? need : kParkFloats;

  auto pair_of = [&](i64 item, SegPair &w) {
    const i64 b = item / p.npairs;
    const i64 sa = (item % p.npairs) * 2;
    w.x = p.x + b * p.x_stride;
    w.x_len = p.x_len;
    w.inA = sa * p.step + p.in_shift;
    w.inB = w.inA + p.step;
    w.haveB = sa + 1 < p.nseg;
    w.out = p.out + b * p.out_stride;
    w.outA = sa * p.step;
    w.outB = w.outA + p.step;
    const i64 remA = p.out_len - w.outA, remB = p.out_len - w.outB;
    w.cntA = (int)(remA < p.step ? remA : p.step);
    w.cntB = w.haveB ? (int)(remB < p.step ? remB : p.step) : 0;
    w.c_off = p.c_off;
    w.accumulate = p.accumulate != 0;
  };
  // a pair can be parked when everything it reads lies inside the signal
  auto parkable = [&](const SegPair &w) {
    return w.haveB && w.inA >= 0 && w.inB + P::N <= w.x_len;
  };
  auto issue_park = [&](const SegPair &w) {             // one thread
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    const unsigned bytes = (unsigned)park_len * 4u;
    mbar_expect_tx(bar, bytes);
    const char *src = reinterpret_cast<const char *>(w.x + w.inA);
    char *dst = reinterpret_cast<char *>(park);
    for (unsigned off = 0; off < bytes; off += 32768u) {
      const unsigned n = bytes - off < 32768u ? bytes - off : 32768u;
      bulk_load(dst + off, src + off, n, bar);
    }
    // the rest of segment B does not fit beside the transform buffer: at least pull it into
    // L2 now, so the first pass finds it there instead of in HBM
    if (need > park_len) {
      const unsigned rest = (unsigned)(need - park_len) * 4u;
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src + bytes), "r"(rest) : "memory");
    }
  };

  if (t == 0) mbar_init(bar, 1);
  {
    TwBases<P> tb0;
    load_twiddle_bases<P>(t, p.tw, tb0);
    tb_sm[t] = tb0;             // read back by the same thread only
  }
  __syncthreads();
  HRegs<P> hr;

  // Loop state is kept small on purpose (the passes need ~124 registers): the pair
  // descriptor is recomputed where it is needed instead of being carried across phases.
  int item = blockIdx.x;
  bool parked = false;
  if (item < p.total) {
    SegPair w0;
    pair_of(item, w0);
    parked = parkable(w0);
    if (parked && t == 0) issue_park(w0);
  }
  unsigned parity = 0;
  bool store_pending = false;     // a bulk store is still reading staged results from `sm`
  asm volatile("griddepcontrol.wait;" ::: "memory");     // the spectrum kernel is short
  for (; item < p.total; item += gridDim.x) {
    {
      SegPair w;
      pair_of(item, w);
      const TwBases<P> tb = tb_sm[t];
      if (parked) {
        mbar_wait(bar, parity);
        parity ^= 1u;
        fwd_first_parked<P>(t, w, tb, sm, park, park_len, p.step, store_pending);
      } else {
        if (store_pending) wait_staged_results_read(t);
        fwd_first<P>(t, w, tb, sm);
      }
      store_pending = false;
    }
    __syncthreads();            // first-pass results visible; the parked range is consumed
    // start fetching the next pair's input while this pair is transformed
    {
      const i64 nxt = (i64)item + gridDim.x;
      parked = false;
      if (nxt < p.total) {
        SegPair wn;
        pair_of(nxt, wn);
        parked = parkable(wn);
        if (parked && t == 0) issue_park(wn);
      }
    }
    if constexpr (P::NPASS == 3) {
      { const TwBases<P> tb = tb_sm[t]; fwd_mid<P>(t, tb, sm); }
      load_spectrum<P>(t, p.spectrum, hr);
      __syncwarp();
      centre<P>(t, hr, sm);
      __syncwarp();
      { const TwBases<P> tb = tb_sm[t]; inv_mid<P>(t, tb, sm); }
    } else {
      load_spectrum<P>(t, p.spectrum, hr);
      centre<P>(t, hr, sm);
    }
    __syncthreads();
    {
      SegPair w;
      pair_of(item, w);
      float2 v[P::E];
      { const TwBases<P> tb = tb_sm[t]; inv_last_regs<P>(t, tb, sm, v); }
      const bool full = !w.accumulate && w.haveB && w.cntA == p.step && w.cntB == p.step;
      if (full) {
        // Results leave through shared memory and one bulk copy per segment (TMA store): the
        // warps do not wait for the stores to drain, the next pair's first pass computes
        // meanwhile.  stage[phase + tt] keeps shared and global addresses congruent mod 16 B.
        float *st = reinterpret_cast<float *>(sm);
        const int phase = (int)((((unsigned long long)(w.out + w.outA)) >> 2) & 3ull);
        const int span = (p.step + 8 + 3) & ~3;
        float *stA = st, *stB = st + span;
        __syncthreads();                  // every thread has read its last-pass inputs
        stage_outputs<P>(t, v, w.c_off, p.step, phase, stA, stB);
        __syncthreads();
        const int head = (4 - phase) & 3;
        const int mid = (p.step - head) & ~3;
        const int tail = p.step - head - mid;
        float *ga = w.out + w.outA, *gb = w.out + w.outB;
        if (t < head) { ga[t] = stA[phase + t]; gb[t] = stB[phase + t]; }
        if (t >= 32 && t < 32 + tail) {
          const int k = head + mid + (t - 32);
          ga[k] = stA[phase + k]; gb[k] = stB[phase + k];
        }
        if (t == 0) {
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          const unsigned bytes = (unsigned)mid * 4u;
          asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(ga + head),
                       "r"(smem_u32(stA + phase + head)), "r"(bytes) : "memory");
          asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gb + head),
                       "r"(smem_u32(stB + phase + head)), "r"(bytes) : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        store_pending = true;
      } else {
        store_outputs<P>(t, w, v);
      }
    }
  }
  // staged results must be out of shared memory before the CTA retires
  if (t == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// ---- plan table ---------------------------------------------------------------------------
// N         radices      E   threads/transform   transforms per CTA
using P256 = Plan<256, 16, 16, 1, 16>;      // 16                8
using P512 = Plan<512, 16, 2, 16, 16>;      // 32                4
using P1024 = Plan<1024, 16, 4, 16, 16>;    // 64                2
using P2048 = Plan<2048, 16, 8, 16, 16>;    // 128               1
using P4096 = Plan<4096, 16, 16, 16, 16>;   // 256               1
using P8192 = Plan<8192, 32, 32, 16, 16>;   // 256               1
using P16384 = Plan<16384, 32, 32, 32, 16>; // 512               1

template <class P, int G, bool REAL>
void launch_fused(const OsParams &p, cudaStream_t stream, bool after_spectrum) {
  auto kernel = REAL ? os_fused_real_kernel<P, G> : os_fused_kernel<P, G>;
  static bool configured[16] = {false};
  size_t smem = sizeof(float2) * P::SMEM_FLOAT2 * G;
  int dev = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 16 && !configured[dev]) {
    B200_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[dev] = true;
  }
  i64 blocks = (p.total + G - 1) / G;
  B200_REQUIRE(blocks < 2147483647LL, "overlap-save grid too large");
  b200_prof_begin(B200_PROF_FUSED, stream);
  OsParams q = p;
  {
    // resident CTAs per wave = SMs x CTAs per SM (by shared memory; registers allow as many)
    static int wave[16] = {0};
    if (dev < 16 && wave[dev] == 0) {
      int per_sm = 0, sms = 0;
      B200_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, P::T * G, smem));
      B200_CUDA_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
      wave[dev] = per_sm * sms > 0 ? per_sm * sms : 1;
    }
    q.prefetch_distance = (i64)(dev < 16 ? wave[dev] : 148) * G;
  }
  if (after_spectrum) {
    // launched right behind the spectrum kernel: let it start before that kernel finishes
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)blocks);
    cfg.blockDim = dim3(P::T * G);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kernel, q));
  } else {
    kernel<<<(unsigned)blocks, P::T * G, smem, stream>>>(q);
  }
  b200_prof_end(B200_PROF_FUSED, stream);
  B200_CUDA_CHECK(cudaGetLastError());
  b200_count_launch(1);
}

template <class P, bool REAL>
void launch_spectrum(const SpecParams &p, int parts, cudaStream_t stream) {
  auto kernel = REAL ? os_spectrum_real_kernel<P> : os_spectrum_kernel<P>;
  static bool configured[16] = {false};
  size_t smem = sizeof(float2) * P::SMEM_FLOAT2;
  int dev = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 16 && !configured[dev]) {
    B200_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[dev] = true;
  }
  b200_prof_begin(B200_PROF_SPECTRUM, stream);
  kernel<<<parts, P::T, smem, stream>>>(p);
  b200_prof_end(B200_PROF_SPECTRUM, stream);
  B200_CUDA_CHECK(cudaGetLastError());
  b200_count_launch(1);
}

// Persistent launch of the 16384-point paired plan (see os_fused_persist_kernel).
bool launch_persist_16384(OsParams op, const DevJob &job, i64 j0, int taps, cudaStream_t stream,
                          bool after_spectrum) {
  using PLAN = P16384;
  if (getenv("VELES_B200_NO_PERSIST")) return false;
  if ((((unsigned long long)job.x) & 15ull) != 0 || (job.x_stride & 3) != 0) return false;
  // leading extra samples so that every segment starts on a 16-byte boundary
  const i64 shift = j0 - job.pad;
  const int e = (int)(((shift % 4) + 4) % 4);
  op.in_shift = shift - e;
  op.c_off = taps - 1 + e;
  op.step = (PLAN::N - op.c_off) & ~3;
  if (op.step < PLAN::N / 4) return false;
  op.nseg = (job.out_len + op.step - 1) / op.step;
  op.npairs = (op.nseg + 1) / 2;
  op.total = op.npairs * job.batch;
  op.prefetch_distance = 0;
  static bool configured[16] = {false};
  static int sms[16] = {0};
  if (op.total >= 2147483647LL) return false;
  const size_t smem = sizeof(float2) * PLAN::SMEM_FLOAT2 + sizeof(float) * kParkFloats +
                      sizeof(fc::TwBases<PLAN>) * PLAN::T + 16;
  int dev = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev >= 16) return false;
  if (!configured[dev]) {
    B200_CUDA_CHECK(cudaFuncSetAttribute(os_fused_persist_kernel<PLAN>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    B200_CUDA_CHECK(cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev));
    configured[dev] = true;
  }
  i64 blocks = op.total < sms[dev] ? op.total : sms[dev];
  b200_prof_begin(B200_PROF_FUSED, stream);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)blocks);
  cfg.blockDim = dim3(PLAN::T);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = after_spectrum ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, os_fused_persist_kernel<PLAN>, op));
  b200_prof_end(B200_PROF_FUSED, stream);
  b200_count_launch(1);
  return true;
}

struct TwiddleCache {
  std::mutex mu;
  const float2 *tab[16][16] = {};   // [device][log2 n]
};
TwiddleCache g_tw;

}  // namespace

// Twiddle table W_n^j = exp(-2*pi*i*j/n), j < n/2, computed in double on the host once per
// device and transform length, kept for the life of the process (<= 64 KB each).
const float2 *b200_twiddles(int n_fft) {
  int dev = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  B200_REQUIRE(dev < 16, "device index out of range");
  int lg = 0;
  while ((1 << lg) < n_fft) lg++;
  std::lock_guard<std::mutex> lock(g_tw.mu);
  if (g_tw.tab[dev][lg]) return g_tw.tab[dev][lg];
  int cnt = n_fft / 2 > 0 ? n_fft / 2 : 1;
  std::vector<float2> host(cnt);
  const double pi2 = 6.283185307179586476925286766559;
  for (int j = 0; j < cnt; j++) {
    double a = -pi2 * j / n_fft;
    host[j] = make_float2((float)cos(a), (float)sin(a));
  }
  float2 *dptr = nullptr;
  B200_CUDA_CHECK(cudaMalloc(&dptr, sizeof(float2) * cnt));
  B200_CUDA_CHECK(cudaMemcpy(dptr, host.data(), sizeof(float2) * cnt, cudaMemcpyHostToDevice));
  g_tw.tab[dev][lg] = dptr;
  return dptr;
}

void b200_os_launch_with(const float2 *tw, float2 *spectrum, int n_fft, int part_taps,
                         const DevJob &job, bool single_segment, cudaStream_t stream) {
  if (job.batch <= 0 || job.out_len <= 0) return;
  const int parts = (int)((job.h_len + part_taps - 1) / part_taps);

  SpecParams sp;
  sp.h = job.h; sp.h_len = job.h_len; sp.reverse = job.reverse;
  sp.part_taps = part_taps; sp.spectrum = spectrum; sp.tw = tw;
  sp.scale = 1.0f / (float)n_fft;

  OsParams op;
  op.x = job.x; op.x_len = job.x_len; op.x_stride = job.x_stride;
  op.out = job.out; op.out_len = job.out_len; op.out_stride = job.out_stride;
  op.tw = tw;

#define B200_DISPATCH(PLAN, GROUPS)                                                    \
  case PLAN::N: {                                                                      \
    launch_spectrum<PLAN, false>(sp, parts, stream);                                   \
    for (int q = 0; q < parts; q++) {                                                  \
      const i64 j0 = (i64)q * part_taps;                                               \
      const i64 left = job.h_len - j0;                                                 \
      const int taps = (int)(left < part_taps ? left : part_taps);                     \
      op.spectrum = spectrum + (size_t)q * PLAN::N;                                    \
      op.accumulate = q > 0;                                                           \
      if (single_segment) {                                                            \
        /* whole-signal transform: leading zeros implicit, no wrapped outputs needed */ \
        op.in_shift = 0; op.c_off = 0; op.step = PLAN::N;                              \
        B200_REQUIRE(parts == 1 && job.pad == job.h_len - 1 &&                         \
                     job.out_len <= PLAN::N && job.x_len + job.h_len - 1 <= PLAN::N,   \
                     "single-segment FFT does not fit the transform");                 \
      } else {                                                                         \
        op.in_shift = j0 - job.pad; op.c_off = taps - 1; op.step = PLAN::N - taps + 1; \
      }                                                                                \
      op.nseg = (job.out_len + op.step - 1) / op.step;                                 \
      op.npairs = (op.nseg + 1) / 2;                                                   \
      op.total = op.npairs * job.batch;                                                \
      if (PLAN::N == 16384 && !single_segment && parts == 1 &&                         \
          launch_persist_16384(op, job, j0, taps, stream, q == 0))                     \
        continue;                                                                      \
      launch_fused<PLAN, GROUPS, false>(op, stream, q == 0);                           \
    }                                                                                  \
  } break;

  switch (n_fft) {
    B200_DISPATCH(P256, 8)
    B200_DISPATCH(P512, 4)
    B200_DISPATCH(P1024, 2)
    B200_DISPATCH(P2048, 1)
    B200_DISPATCH(P4096, 1)
    B200_DISPATCH(P8192, 1)
    B200_DISPATCH(P16384, 1)
    default:
      B200_REQUIRE(false, "unsupported transform length");
  }
#undef B200_DISPATCH
}

// Real-packed mode: one real segment of 2*n_fft samples per CTA, all taps in one partition
// (h_len <= n_fft).  `coeffs` holds n_fft float4 {alpha, beta}.
void b200_os_launch_real(const float2 *tw, float4 *coeffs, int n_fft, const DevJob &job,
                         cudaStream_t stream) {
  if (job.batch <= 0 || job.out_len <= 0) return;
  B200_REQUIRE(job.h_len <= n_fft, "real-packed mode needs all taps in one transform");
  SpecParams sp;
  sp.h = job.h; sp.h_len = job.h_len; sp.reverse = job.reverse;
  sp.part_taps = (int)job.h_len; sp.spectrum = reinterpret_cast<float2 *>(coeffs); sp.tw = tw;
  sp.scale = 1.0f / (float)n_fft;

  OsParams op;
  op.x = job.x; op.x_len = job.x_len; op.x_stride = job.x_stride;
  op.out = job.out; op.out_len = job.out_len; op.out_stride = job.out_stride;
  op.tw = tw;
  op.spectrum = reinterpret_cast<const float2 *>(coeffs);
  op.accumulate = 0;
  // One extra leading sample when h_len is even makes the first useful circular sample even,
  // so every segment's results are whole {2p, 2p+1} pairs (8-byte stores).  For a full
  // convolution (pad = h_len-1) the segment starts are then even too (8-byte loads); for a
  // shard with another pad they may be odd, which only costs two 4-byte loads per pair.
  const int e = (int)((job.h_len - 1) & 1);
  op.in_shift = -job.pad - e;
  op.c_off = (int)job.h_len - 1 + e;
  op.step = (2 * n_fft - op.c_off) & ~1;
  op.nseg = (job.out_len + op.step - 1) / op.step;
  op.npairs = op.nseg;
  op.total = op.nseg * job.batch;

#define B200_DISPATCH_REAL(PLAN, GROUPS)                       \
  case PLAN::N:                                                \
    launch_spectrum<PLAN, true>(sp, 1, stream);                \
    launch_fused<PLAN, GROUPS, true>(op, stream, true);        \
    break;
  switch (n_fft) {
    B200_DISPATCH_REAL(P256, 8)
    B200_DISPATCH_REAL(P512, 4)
    B200_DISPATCH_REAL(P1024, 2)
    B200_DISPATCH_REAL(P2048, 1)
    B200_DISPATCH_REAL(P4096, 1)
    B200_DISPATCH_REAL(P8192, 1)
    B200_DISPATCH_REAL(P16384, 1)
    default:
      B200_REQUIRE(false, "unsupported transform length");
  }
#undef B200_DISPATCH_REAL
}

void b200_os_launch(const OsDevPlan &plan, const DevJob &job, bool single_segment,
                    cudaStream_t stream) {
  if (plan.real_packed && !single_segment) {
    b200_os_launch_real(plan.twiddles, reinterpret_cast<float4 *>(plan.spectrum), plan.n_fft, job,
                        stream);
    return;
  }
  const int parts = (int)((job.h_len + plan.part_taps - 1) / plan.part_taps);
  B200_REQUIRE(parts <= plan.max_parts, "impulse response longer than the plan was sized for");
  b200_os_launch_with(plan.twiddles, plan.spectrum, plan.n_fft, plan.part_taps, job,
                      single_segment, stream);
}

// ---- allocation-free entry point ---------------------------------------------------------
// simd_b200_filter_noalloc (declared in include/simd/b200.h) runs the same two kernels as
// convolve_overlap_save, but every byte of device memory they need besides the caller's own
// x, h and result lives in the two file-scope arrays below, which the CUDA runtime places
// on each GPU when this library is loaded there and keeps until the process exits:
//   g_noalloc_spectrum   the fused kernel's scratch (the filter spectrum of one call)
//   g_noalloc_twiddles   read-only twiddle tables for every supported transform length,
//                        packed back to back (table for length n starts at n/2 - 128)
// In host code these two names are only handles for the runtime, not device addresses: the
// address a kernel can use is obtained with cudaGetSymbolAddress, once per device, while
// that device is current (noalloc_setup below).
#define B200_NOALLOC_MAX_N 16384          /* complex bins -> 128 KiB of scratch per GPU */
#define B200_NOALLOC_MAX_TAPS (B200_NOALLOC_MAX_N / 2)
#define B200_NOALLOC_TW_TOTAL (B200_NOALLOC_MAX_N - 128)   /* sum of n/2 for n = 256..16384 */

static __device__ __align__(256) float2 g_noalloc_spectrum[B200_NOALLOC_MAX_N];
static __device__ __align__(256) float2 g_noalloc_twiddles[B200_NOALLOC_TW_TOTAL];

namespace {

// Fills every packed twiddle table: entry j of the table for length n is exp(-2*pi*i*j/n),
// evaluated in double precision on the device (no host buffer, no copy).
__global__ void noalloc_twiddle_kernel(float2 *tables) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B200_NOALLOC_TW_TOTAL) return;
  int n = 256;
  while (i >= n - 128) n *= 2;            // table of length n occupies [n/2 - 128, n - 128)
  const int j = i - (n / 2 - 128);
  double s, c;
  sincospi(-2.0 * (double)j / (double)n, &s, &c);
  tables[i] = make_float2((float)c, (float)s);
}

struct NoAllocDevice {
  bool ready = false;
  float2 *spectrum = nullptr;
  float2 *twiddles = nullptr;
};
NoAllocDevice g_noalloc_dev[16];
std::mutex g_noalloc_mu;

// Once per device, with that device current.  Returns cudaSuccess or the failing call's code.
cudaError_t noalloc_setup(int dev, cudaStream_t stream) {
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  NoAllocDevice &d = g_noalloc_dev[dev];
  if (d.ready) return cudaSuccess;
  cudaError_t err = cudaGetSymbolAddress((void **)&d.spectrum, g_noalloc_spectrum);
  if (err != cudaSuccess) return err;
  err = cudaGetSymbolAddress((void **)&d.twiddles, g_noalloc_twiddles);
  if (err != cudaSuccess) return err;
  noalloc_twiddle_kernel<<<(B200_NOALLOC_TW_TOTAL + 255) / 256, 256, 0, stream>>>(d.twiddles);
  err = cudaGetLastError();
  if (err != cudaSuccess) return err;
  b200_count_launch(1);
  d.ready = true;     // later launches on other streams are ordered by the caller (see header)
  return cudaStreamSynchronize(stream);
}

bool noalloc_device_accessible(const void *p) {
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged;
}

}  // namespace

// See include/simd/b200.h for the contract.  Returns 0 on success.
extern "C" __attribute__((visibility("default"))) int simd_b200_filter_noalloc(
    const float *x, size_t x_length, const float *h, size_t h_length, int reverse,
    float *result, void *cuda_stream) {
  if (!x || !h || !result || x_length == 0 || h_length == 0) {
    fprintf(stderr, "libSimd(b200): simd_b200_filter_noalloc: null pointer or zero length\n");
    return 2;
  }
  if (h_length > B200_NOALLOC_MAX_TAPS) {
    fprintf(stderr,
            "libSimd(b200): simd_b200_filter_noalloc: h_length %zu exceeds the %d taps that the "
            "built-in %d-bin scratch array can hold; use convolve_overlap_save for longer "
            "impulse responses\n",
            h_length, B200_NOALLOC_MAX_TAPS, B200_NOALLOC_MAX_N);
    return 1;
  }
  if (x_length + h_length - 1 >= ((size_t)1 << 40)) {
    fprintf(stderr, "libSimd(b200): simd_b200_filter_noalloc: signal too long\n");
    return 1;
  }
  if (!noalloc_device_accessible(x) || !noalloc_device_accessible(h) ||
      !noalloc_device_accessible(result)) {
    fprintf(stderr,
            "libSimd(b200): simd_b200_filter_noalloc: x, h and result must be device memory "
            "(staging host buffers would need an allocation)\n");
    return 2;
  }
  int dev = -1;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess || dev < 0 || dev >= 16) {
    fprintf(stderr, "libSimd(b200): simd_b200_filter_noalloc: no usable CUDA device\n");
    return 3;
  }
  cudaStream_t stream = (cudaStream_t)cuda_stream;
  err = noalloc_setup(dev, stream);
  if (err != cudaSuccess) {
    fprintf(stderr, "libSimd(b200): simd_b200_filter_noalloc: set-up failed: %s\n",
            cudaGetErrorString(err));
    return 3;
  }
  // transform length: the library's usual choice, raised until all taps fit one spectrum
  int n = b200_os_pick_n((long long)h_length, (long long)(x_length + h_length - 1));
  while ((size_t)(n / 2) < h_length) n *= 2;
  DevJob job;
  job.x = x; job.x_len = (i64)x_length; job.x_stride = (i64)x_length;
  job.h = h; job.h_len = (i64)h_length;
  job.out = result; job.out_len = (i64)(x_length + h_length - 1); job.out_stride = job.out_len;
  job.pad = (i64)h_length - 1; job.batch = 1; job.reverse = reverse ? 1 : 0;
  const NoAllocDevice &d = g_noalloc_dev[dev];
  // every bin of the spectrum array is rewritten by the spectrum kernel of this call, so
  // nothing left over from the previous call is ever read
  b200_os_launch_with(d.twiddles + (n / 2 - 128), d.spectrum, n, (int)h_length, job, false, stream);
  return 0;
}
