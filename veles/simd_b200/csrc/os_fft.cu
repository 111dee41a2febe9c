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
#include "tmem_store.cuh"
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
  unsigned per_signal_rcp;  // floor(2^32 / work items per signal) when everything fits 31 bits, else 0
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

// item -> (signal, unit inside the signal).  The division is by a launch constant: when the
// numbers fit 31 bits the host passes rcp = floor(2^32 / per_signal) (0xffffffff for 1) and the
// quotient is umulhi(item, rcp), at most one too small -- five instructions instead of the
// two dozen of a 32-bit division, and the 64-bit one costs more than a radix-4 butterfly network.
__device__ __forceinline__ void split_item(i64 item, i64 per_signal, unsigned rcp, i64 &sig, i64 &unit) {
  if (rcp != 0) {
    const unsigned n = (unsigned)item, d = (unsigned)per_signal;
    unsigned q = __umulhi(n, rcp);
    unsigned r = n - q * d;
    if (r >= d) { q++; r -= d; }
    sig = q;
    unit = r;
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
  if (threadIdx.x < G) {
    // Pull the input of the CTA that will follow this one on the SM into L2 while this CTA
    // computes (CTAs are dispatched in index order, one wave = p.prefetch_distance items).
    // The first G threads do it, one group's item each, so that one warp pays for it.
    const i64 nxt = (i64)blockIdx.x * G + threadIdx.x + p.prefetch_distance;
    if (p.prefetch_distance > 0 && nxt < p.total) {
      i64 nb, nsa;
      split_item(nxt, p.npairs, p.per_signal_rcp, nb, nsa);
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
    split_item(item, p.npairs, p.per_signal_rcp, b, sa);
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
  if (threadIdx.x < G) {
    const i64 nxt = (i64)blockIdx.x * G + threadIdx.x + p.prefetch_distance;
    if (p.prefetch_distance > 0 && nxt < p.total) {
      i64 nb, nsg;
      split_item(nxt, p.nseg, p.per_signal_rcp, nb, nsg);
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
    split_item(item, p.nseg, p.per_signal_rcp, b, sg);
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
  if (live) centre_real_fwd_keep<P>(t, sm, u);
  asm volatile("griddepcontrol.wait;" ::: "memory");
  group_sync<P, G>(grp);                // mirror bins live in other threads of the group
  if (live) centre_real_mix_keep<P>(t, reinterpret_cast<const float4 *>(p.spectrum), sm, u);
  group_sync<P, G>(grp);                // every mirror has been read: safe to overwrite
  if (live) centre_real_store<P>(t, sm, u);
  if constexpr (P::NPASS == 3) {
    __syncwarp();
    if (live) inv_mid<P>(t, tb, sm);
  }
  group_sync<P, G>(grp);
  if (live) inv_last_real<P>(t, w, tb, sm);
}

// ---- mbarrier / bulk-copy helpers of the persistent kernel ---------------------------------
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

// ---- passes of the persistent kernel with the twiddle factors kept in Tensor Memory -------
// Same arithmetic as fwd_first / fwd_mid / inv_mid / inv_last of fft_core.cuh, except that
// the 31 factors of a butterfly come out of the thread's private TMEM words (tmem_store.cuh)
// instead of being rebuilt from their base: column 2k, 2k+1 of set 0 holds W_N^(t*k) (first /
// last pass), of set 1 (64 columns further) W_NB2^((t & 15)*k) (middle passes).  Four x16
// loads per pass, issued two at a time ahead of their use.

// v[k] *= factor k for k in [8c, 8c+8), k > 0 (forward passes: factors after the butterfly)
template <int C>
__device__ __forceinline__ void tm_apply_fwd(float2 *v, const tmem::Chunk &c) {
  static_for<0, 8>([&](auto ii) {
    constexpr int i = decltype(ii)::value;
    constexpr int k = 8 * C + i;
    if constexpr (k > 0) v[k] = cmul(v[k], c.get(i));
  });
}
// v[bitrev(k)] = e[k] * conj(factor k) (inverse passes: factors before the butterfly)
template <int C, int R>
__device__ __forceinline__ void tm_apply_inv(float2 *v, const float2 *e, const tmem::Chunk &c) {
  static_for<0, 8>([&](auto ii) {
    constexpr int i = decltype(ii)::value;
    constexpr int k = 8 * C + i;
    if constexpr (k > 0) v[bitrev(k, ilog2(R))] = cmul_conj(e[k], c.get(i));
    else v[0] = e[0];
  });
}

__device__ __forceinline__ void tm_twiddle_fwd(float2 *v, unsigned set_addr, tmem::Chunk &a, tmem::Chunk &b) {
  // a, b hold chunks 0 and 1 (loads issued by the caller before the butterfly)
  tmem::wait_loads(a, b);
  tm_apply_fwd<0>(v, a);
  tmem::load(a, set_addr + 32);
  tm_apply_fwd<1>(v, b);
  tmem::load(b, set_addr + 48);
  tmem::wait_loads(a, b);
  tm_apply_fwd<2>(v, a);
  tm_apply_fwd<3>(v, b);
}

// Twiddle factors applied chunk by chunk, each chunk stored as soon as it is done, so that the 32
// stores of a thread spread over the multiplications instead of leaving in one burst.
template <class G>
__device__ __forceinline__ void tm_twiddle_store_fwd(float2 *v, float2 *smb, unsigned set_addr, tmem::Chunk &a,
                                                     tmem::Chunk &b) {
  auto store8 = [&](auto cc) {
    constexpr int c = decltype(cc)::value;
    static_for<0, 8>([&](auto ii) {
      constexpr int k = 8 * c + decltype(ii)::value;
      smb[G::off(k)] = v[k];
    });
  };
  tmem::wait_loads(a, b);
  tm_apply_fwd<0>(v, a);
  store8(std::integral_constant<int, 0>{});
  tmem::load(a, set_addr + 32);
  tm_apply_fwd<1>(v, b);
  store8(std::integral_constant<int, 1>{});
  tmem::load(b, set_addr + 48);
  tmem::wait_loads(a, b);
  tm_apply_fwd<2>(v, a);
  store8(std::integral_constant<int, 2>{});
  tm_apply_fwd<3>(v, b);
  store8(std::integral_constant<int, 3>{});
}

// Forward first pass.  x[inA, inA + park_len) -- zeros where that leaves the signal -- was
// brought into shared memory (`park`) while the previous pair was being processed: all of
// segment A and the head of segment B (which starts `step` samples later); the rows of B
// beyond it come from global memory (the caller parks only pairs whose segment B ends inside
// the signal).  `before_stores` runs between the arithmetic and the stores into the transform
// buffer (the previous pair's staged results may still be leaving it).
template <class P, class F>
__device__ __forceinline__ void fwd_first_parked_tm(int t, const SegPair &w, float2 *sm, const float *park,
                                                    int park_len, int step, unsigned tw_addr,
                                                    F &&before_stores) {
  constexpr int R = P::R1, NB = P::N;
  static_assert(R == 32 && P::E == R, "one radix-32 butterfly per thread");
  using G = PassGeom<P, NB, R>;
  int base, npp, bf;
  G::locate(t, 0, base, npp, bf);
  auto *smb = sm + pad_pos(base);
  float2 v[R];
  tmem::Chunk a, b;
  tmem::load(a, tw_addr);
  tmem::load(b, tw_addr + 16);
  const float *xb = w.x + (w.inB + base);
  const float *pa = park + base;
  static_for<0, R>([&](auto rr) {
    constexpr int r = decltype(rr)::value;
    const float va = pa[r * G::S];
    float vb;
    if (step + (r + 1) * G::S <= park_len) vb = pa[step + r * G::S];   // CTA-uniform
    else vb = xb[r * G::S];
    v[bitrev(r, ilog2(R))] = make_float2(va, vb);
  });
  dft_inreg<R, false>(v);
  before_stores();          // ahead of the twiddle multiplications, so that the stores mix with them
  tm_twiddle_store_fwd<G>(v, smb, tw_addr, a, b);
}

// pairs that cannot be parked (nothing of them inside the signal is bulk-copyable)
template <class P, class F>
__device__ __forceinline__ void fwd_first_tm(int t, const SegPair &w, float2 *sm, unsigned tw_addr,
                                             F &&before_stores) {
  constexpr int R = P::R1, NB = P::N;
  using G = PassGeom<P, NB, R>;
  int base, npp, bf;
  G::locate(t, 0, base, npp, bf);
  auto *smb = sm + pad_pos(base);
  float2 v[R];
  tmem::Chunk a, b;
  tmem::load(a, tw_addr);
  tmem::load(b, tw_addr + 16);
  if (w.haveB && w.inA >= 0 && w.inB + P::N <= w.x_len) {      // CTA-uniform: no bounds checks
    const float *xa = w.x + (w.inA + base), *xb = w.x + (w.inB + base);
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      v[bitrev(r, ilog2(R))] = make_float2(xa[r * G::S], xb[r * G::S]);
    });
  } else {
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      const int p = base + r * G::S;
      const float va = load_x(w.x, w.x_len, w.inA + p);
      const float vb = w.haveB ? load_x(w.x, w.x_len, w.inB + p) : 0.0f;
      v[bitrev(r, ilog2(R))] = make_float2(va, vb);
    });
  }
  dft_inreg<R, false>(v);
  before_stores();
  tm_twiddle_store_fwd<G>(v, smb, tw_addr, a, b);
}

template <class P>
__device__ __forceinline__ void fwd_mid_tm(int t, float2 *sm, unsigned tw_addr) {
  constexpr int R = P::R2, NB = P::NB2;
  static_assert(R == 32 && P::E == R, "one radix-32 butterfly per thread");
  using G = PassGeom<P, NB, R>;
  int base, npp, bf;
  G::locate_mid(t, 0, base, npp, bf);
  auto *smb = sm + pad_pos(base);
  float2 v[R];
  tmem::Chunk a, b;
  tmem::load(a, tw_addr + 64);
  tmem::load(b, tw_addr + 64 + 16);
  static_for<0, R>([&](auto rr) {
    constexpr int r = decltype(rr)::value;
    v[bitrev(r, ilog2(R))] = smb[G::off(r)];
  });
  dft_inreg<R, false>(v);
  tm_twiddle_fwd(v, tw_addr + 64, a, b);
  static_for<0, R>([&](auto kk) {
    constexpr int k = decltype(kk)::value;
    smb[G::off(k)] = v[k];
  });
}

// shared -> conj twiddle -> v[] in butterfly input order (both inverse strided passes)
template <class P, class G, int R>
__device__ __forceinline__ void tm_load_twiddled_inv(const float2 *smb, float2 *v, unsigned set_addr) {
  tmem::Chunk a, b;
  tmem::load(a, set_addr);
  tmem::load(b, set_addr + 16);
  float2 e[R];
  static_for<0, R>([&](auto kk) {
    constexpr int k = decltype(kk)::value;
    e[k] = smb[G::off(k)];
  });
  tmem::wait_loads(a, b);
  tm_apply_inv<0, R>(v, e, a);
  tmem::load(a, set_addr + 32);
  tm_apply_inv<1, R>(v, e, b);
  tmem::load(b, set_addr + 48);
  tmem::wait_loads(a, b);
  tm_apply_inv<2, R>(v, e, a);
  tm_apply_inv<3, R>(v, e, b);
}

template <class P>
__device__ __forceinline__ void inv_mid_tm(int t, float2 *sm, unsigned tw_addr) {
  constexpr int R = P::R2, NB = P::NB2;
  using G = PassGeom<P, NB, R>;
  int base, npp, bf;
  G::locate_mid(t, 0, base, npp, bf);
  auto *smb = sm + pad_pos(base);
  float2 v[R];
  tm_load_twiddled_inv<P, G, R>(smb, v, tw_addr + 64);
  dft_inreg<R, true>(v);
  static_for<0, R>([&](auto rr) {
    constexpr int r = decltype(rr)::value;
    smb[G::off(r)] = v[r];
  });
}

// `after_loads` runs once this thread's inputs are out of the transform buffer, before the
// butterfly network
template <class P, class F>
__device__ __forceinline__ void inv_last_regs_tm(int t, const float2 *sm, float2 *v, unsigned tw_addr,
                                                 F &&after_loads) {
  constexpr int R = P::R1, NB = P::N;
  using G = PassGeom<P, NB, R>;
  int base, npp, bf;
  G::locate(t, 0, base, npp, bf);
  tm_load_twiddled_inv<P, G, R>(sm + pad_pos(base), v, tw_addr);
  after_loads();
  dft_inreg<R, true>(v);
}

// Once per CTA: every thread writes its two sets of 32 factors, taken straight from the
// tabulated W_N^j (j < N/2, computed in double on the host) -- no product trees, so each
// factor is correctly rounded.
template <class P>
__device__ __forceinline__ void tm_fill_twiddles(int t, const float2 *tw, unsigned tw_addr) {
  auto w_pow = [&](int idx) {           // W_N^idx, idx in [0, N)
    const float2 v = tw[idx & (P::N / 2 - 1)];
    return (idx & (P::N / 2)) ? make_float2(-v.x, -v.y) : v;
  };
  const int n1 = t;                      // first / last pass: butterfly t, stride N/32
  const int n2 = (t & 15) * P::R1;       // middle passes: W_NB2^(t & 15) = W_N^(R1 * (t & 15))
  for (int c = 0; c < 4; c++) {
    tmem::Chunk a, b;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const int k = 8 * c + i;
      a.set(i, w_pow((n1 * k) & (P::N - 1)));
      b.set(i, w_pow((n2 * k) & (P::N - 1)));
    }
    tmem::store(tw_addr + 16 * c, a);
    tmem::store(tw_addr + 64 + 16 * c, b);
  }
  tmem::wait_stores();
}

// ---- persistent kernel for the one-CTA-per-SM plan (N = 16384, paired segments) -----------
// One CTA per SM loops over segment pairs.
//  * Twiddle factors live in Tensor Memory (filled once per CTA, see above).
//  * While a pair is being transformed, a bulk asynchronous copy (cp.async.bulk / TMA,
//    completion on an mbarrier) brings the next pair's input range into the shared memory
//    left beside the transform buffer, so the next pair's first pass starts from shared
//    memory instead of waiting for HBM, and the current pair's result stores drain while
//    that first pass computes.  Needs 16-byte aligned segment starts (the host picks step
//    and leading-sample count accordingly).  The first pair of a signal starts before it:
//    that part of the parked range is zero-filled by the threads, the part inside is
//    bulk-copied as for every other pair.  (A last pair whose segment B runs past the
//    signal's end takes the bounds-checked loads.)
//  * Results leave through shared memory and one bulk store per segment.
constexpr int kParkFloats = 23040;      // 90 KiB beside the 136 KiB transform buffer

template <class P>
__global__ void __launch_bounds__(P::T) os_fused_persist_kernel(OsParams p) {
  extern __shared__ float4 smem4[];
  float2 *sm = reinterpret_cast<float2 *>(smem4);
  float *park = reinterpret_cast<float *>(sm + P::SMEM_FLOAT2);
  unsigned long long *bar = reinterpret_cast<unsigned long long *>(park + kParkFloats);
  unsigned *tmem_base = reinterpret_cast<unsigned *>(bar + 1);
  const int t = threadIdx.x;

  // parked range of a pair: X(inA + q), q in [0, park_len), clipped to what the pair reads
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
  // The part of the parked range that lies inside the signal, [lo, lo + n) in park indices
  // (both multiples of 4 floats: segment starts are 16-byte aligned); n <= 0: nothing to copy.
  auto park_span = [&](const SegPair &w, int &lo, int &n) {
    lo = w.inA < 0 ? (int)(-w.inA < (i64)park_len ? -w.inA : park_len) : 0;
    const i64 end = w.x_len - w.inA;                    // first park index past the signal
    const int hi = end < (i64)park_len ? (int)(end > 0 ? end : 0) : park_len;
    n = hi > lo ? (hi - lo) & ~3 : 0;
  };
  // all threads, after the barrier that ends the reads of the previous parked range
  auto start_park = [&](const SegPair &w) -> bool {
    int lo, n;
    park_span(w, lo, n);
    if (n <= 0) return false;                           // CTA-uniform
    // rows of B behind the parked range are read from global memory without bounds checks
    // (a per-row check in the hot first pass measured 2.4 % on C4)
    if (need > park_len && w.inB + P::N > w.x_len) return false;
    if (lo > 0 || lo + n < park_len) {
      // a signal edge: zeros in front of the signal, zeros (and up to 3 samples the 16-byte
      // granularity left over) behind it
      for (int q = t; q < lo; q += P::T) park[q] = 0.0f;
      for (int q = lo + n + t; q < park_len; q += P::T) park[q] = load_x(w.x, w.x_len, w.inA + q);
    }
    if (t == 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      const unsigned bytes = (unsigned)n * 4u;
      mbar_expect_tx(bar, bytes);
      const char *src = reinterpret_cast<const char *>(w.x + (w.inA + lo));
      char *dst = reinterpret_cast<char *>(park + lo);
      for (unsigned off = 0; off < bytes; off += 32768u) {
        const unsigned m = bytes - off < 32768u ? bytes - off : 32768u;
        bulk_load(dst + off, src + off, m, bar);
      }
      // the rest of segment B does not fit beside the transform buffer: at least pull it into
      // L2 now, so the first pass finds it there instead of in HBM
      const i64 rest_lo = w.inA + park_len;
      i64 rest_hi = w.inA + need;
      if (rest_hi > w.x_len) rest_hi = w.x_len;
      if (rest_hi > rest_lo) {
        const unsigned rest = (unsigned)(rest_hi - rest_lo) * 4u & ~15u;
        if (rest) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(w.x + rest_lo), "r"(rest) : "memory");
      }
    }
    return true;
  };

  if (t == 0) mbar_init(bar, 1);
  if (t < 32) tmem::alloc_all(tmem_base);
  tmem::fence_before_sync();
  __syncthreads();
  tmem::fence_after_sync();
  const unsigned tw_addr = tmem::thread_base(*tmem_base, t);
  tm_fill_twiddles<P>(t, p.tw, tw_addr);
  HRegs<P> hr;

  // Loop state is kept small on purpose (the passes need ~124 registers): the pair
  // descriptor is recomputed where it is needed instead of being carried across phases.
  int item = blockIdx.x;
  bool parked = false;
  if (item < p.total) {
    SegPair w0;
    pair_of(item, w0);
    parked = start_park(w0);
  }
  __syncthreads();                // zeros the threads wrote into an edge pair's parked range
                                  // (later pairs: the barriers of the previous pair do this)
  unsigned parity = 0;
  bool store_pending = false;     // a bulk store is still reading staged results from `sm`
  // the spectrum kernel is short; waiting for it here rather than at the first centre pass keeps
  // the statement out of the loop (inside it measured 2.7 % on C4)
  asm volatile("griddepcontrol.wait;" ::: "memory");
  for (; item < p.total; item += gridDim.x) {
    {
      SegPair w;
      pair_of(item, w);
      // before the first pass overwrites the transform buffer, the bulk store of the previous
      // pair must have read its staged results out of it
      auto drained = [&]() {
        if (store_pending) wait_staged_results_read(t);     // CTA-uniform
      };
      if (parked) {
        mbar_wait(bar, parity);
        parity ^= 1u;
        fwd_first_parked_tm<P>(t, w, sm, park, park_len, p.step, tw_addr, drained);
      } else {
        fwd_first_tm<P>(t, w, sm, tw_addr, drained);
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
        parked = start_park(wn);
      }
    }
    fwd_mid_tm<P>(t, sm, tw_addr);
    load_spectrum<P>(t, p.spectrum, hr);
    __syncwarp();
    centre<P>(t, hr, sm);
    __syncwarp();
    inv_mid_tm<P>(t, sm, tw_addr);
    __syncthreads();
    {
      SegPair w;
      pair_of(item, w);
      float2 v[P::E];
      // the barrier "every thread has read its last-pass inputs" sits between those reads and
      // the last butterfly network: the staging stores below then follow the arithmetic that
      // produces them instead of a barrier that releases all 16 warps' stores at once
      inv_last_regs_tm<P>(t, sm, v, tw_addr, [&]() { if (!p.accumulate) __syncthreads(); });
      if (!w.accumulate) {
        // Results leave through shared memory and one bulk copy per segment (TMA store): the
        // warps do not wait for the stores to drain, the next pair's first pass computes
        // meanwhile.  stage[phase + tt] keeps shared and global addresses congruent mod 16 B
        // (both segments: their outputs start a multiple of 4 floats apart).
        float *st = reinterpret_cast<float *>(sm);
        const int phase = (int)((((unsigned long long)(w.out + w.outA)) >> 2) & 3ull);
        const int span = (p.step + 8 + 3) & ~3;
        float *stA = st, *stB = st + span;
        stage_outputs<P>(t, v, w.c_off, p.step, phase, stA, stB);
        __syncthreads();
        // per segment: `head` scalar samples up to the next 16-byte boundary, a bulk store of
        // whole 16-byte units, up to 3 scalar samples behind it
        const int head = (4 - phase) & 3;
        float *ga = w.out + w.outA, *gb = w.out + w.outB;
        const int midA = w.cntA > head ? (w.cntA - head) & ~3 : 0;
        const int midB = w.cntB > head ? (w.cntB - head) & ~3 : 0;
        auto scalar_edges = [&](int u, float *g, const float *stg, int cnt, int mid) {
          const int k = u < 4 ? u : head + mid + (u - 4);
          if ((u < 4 ? k < head : true) && k < cnt) g[k] = stg[phase + k];
        };
        if (t < 8) scalar_edges(t, ga, stA, w.cntA, midA);
        else if (t >= 32 && t < 40) scalar_edges(t - 32, gb, stB, w.cntB, midB);
        if (t == 0) {
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          if (midA > 0)
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(ga + head),
                         "r"(smem_u32(stA + phase + head)), "r"((unsigned)midA * 4u) : "memory");
          if (midB > 0)
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gb + head),
                         "r"(smem_u32(stB + phase + head)), "r"((unsigned)midB * 4u) : "memory");
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
  tmem::fence_before_sync();
  __syncthreads();
  if (t < 32) tmem::free_all(*tmem_base);
}

// ---- persistent variant of the real-packed kernel (N = 16384: one real segment of 32768
// samples per transform; config C5's h = 8192) ------------------------------------------------
// One CTA per SM loops over segments with its twiddle factors in Tensor Memory (filled once per
// CTA), the next segment's input prefetched into L2.  Everything else is the plain kernel's:
// 8-byte global loads of the input rows, {alpha, beta} from L2, 8-byte global stores of the
// results.  This kernel is bound by the shared-memory instruction queue (`mio_throttle` is its
// largest stall), so what helped the paired kernel hurts here -- measured on one GPU, C5:
//   plain kernel 4.670 ms | + Tensor-Memory twiddles, persistent 4.256 | + results staged in
//   shared memory and bulk-stored 4.504 | + the first 5 of 16 {alpha, beta} rows resident in
//   shared memory 4.438 | both 4.583.  Against the final kernel (4.03 ms): the first 22 of 32
//   input rows of the next segment parked in shared memory by a bulk copy (the paired kernel's
//   +5 %) 4.47 ms -- 22 more shared-memory reads per thread cost more than the global loads they
//   replace, whose latency the L2 prefetch already hides.
template <class P>
__device__ __forceinline__ void fwd_first_real_tm(int t, const SegReal &w, float2 *sm, unsigned tw_addr,
                                                  bool buffer_in_use) {
  constexpr int R = P::R1, NB = P::N;
  static_assert(R == 32 && P::E == R, "one radix-32 butterfly per thread");
  using G = PassGeom<P, NB, R>;
  int base, npp, bf;
  G::locate(t, 0, base, npp, bf);
  auto *smb = sm + pad_pos(base);
  float2 v[R];
  tmem::Chunk a, b;
  tmem::load(a, tw_addr);
  tmem::load(b, tw_addr + 16);
  const bool inside = w.in0 >= 0 && w.in0 + 2 * P::N <= w.x_len;                   // CTA-uniform
  const bool fast = inside && (((unsigned long long)(w.x + w.in0)) & 7ull) == 0;   // 8-byte loads
  if (fast) {
    const float2 *xa = reinterpret_cast<const float2 *>(w.x + w.in0) + base;
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      v[bitrev(r, ilog2(R))] = xa[r * G::S];
    });
  } else if (inside) {
    const float *xa = w.x + (w.in0 + 2 * (i64)base);
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      v[bitrev(r, ilog2(R))] = make_float2(xa[2 * r * G::S], xa[2 * r * G::S + 1]);
    });
  } else {
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      const i64 i = w.in0 + 2 * (i64)(base + r * G::S);
      v[bitrev(r, ilog2(R))] = make_float2(load_x(w.x, w.x_len, i), load_x(w.x, w.x_len, i + 1));
    });
  }
  dft_inreg<R, false>(v);
  tm_twiddle_fwd(v, tw_addr, a, b);
  // the previous segment's last pass may still be reading the transform buffer in other threads:
  // wait here, after this segment's loads and first butterflies, not before them (CTA-uniform)
  if (buffer_in_use) __syncthreads();
  static_for<0, R>([&](auto kk) {
    constexpr int k = decltype(kk)::value;
    smb[G::off(k)] = v[k];
  });
}

// last pass -> registers: register r holds circular samples 2p and 2p+1, p = base + r*S
template <class P>
__device__ __forceinline__ void inv_last_real_regs_tm(int t, const float2 *sm, float2 *v, unsigned tw_addr) {
  constexpr int R = P::R1, NB = P::N;
  using G = PassGeom<P, NB, R>;
  int base, npp, bf;
  G::locate(t, 0, base, npp, bf);
  tm_load_twiddled_inv<P, G, R>(sm + pad_pos(base), v, tw_addr);
  dft_inreg<R, true>(v);
}

template <class P>
__global__ void __launch_bounds__(P::T) os_fused_real_persist_kernel(OsParams p) {
  extern __shared__ float4 smem4[];
  float2 *sm = reinterpret_cast<float2 *>(smem4);
  unsigned *tmem_base = reinterpret_cast<unsigned *>(sm + P::SMEM_FLOAT2);
  const int t = threadIdx.x;
  constexpr int R = P::R1;
  using G = PassGeom<P, P::N, R>;

  if (t < 32) tmem::alloc_all(tmem_base);
  tmem::fence_before_sync();
  __syncthreads();
  tmem::fence_after_sync();
  const unsigned tw_addr = tmem::thread_base(*tmem_base, t);
  tm_fill_twiddles<P>(t, p.tw, tw_addr);

  auto seg_of = [&](i64 item, SegReal &w) {
    const i64 b = item / p.nseg;
    const i64 sg = item % p.nseg;
    w.x = p.x + b * p.x_stride;
    w.x_len = p.x_len;
    w.in0 = sg * p.step + p.in_shift;
    w.out = p.out + b * p.out_stride;
    w.out0 = sg * p.step;
    const i64 rem = p.out_len - w.out0;
    w.cnt = (int)(rem < p.step ? rem : p.step);
    w.c_off = p.c_off;
    w.accumulate = false;
  };
  auto prefetch = [&](i64 item) {                 // one thread: next segment's input into L2
    if (item >= p.total) return;
    SegReal w;
    seg_of(item, w);
    i64 lo = w.in0 < 0 ? 0 : w.in0;
    i64 hi = w.in0 + 2 * P::N;
    if (hi > w.x_len) hi = w.x_len;
    const unsigned long long a0 = ((unsigned long long)(w.x + lo) + 15ull) & ~15ull;
    const unsigned long long a1 = (unsigned long long)(w.x + hi) & ~15ull;
    if (a1 > a0)
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a0), "r"((unsigned)(a1 - a0)) : "memory");
  };

  // the coefficient kernel runs ahead of us (programmatic dependent launch)
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const float4 *AB = reinterpret_cast<const float4 *>(p.spectrum);

  for (i64 item = blockIdx.x; item < p.total; item += gridDim.x) {
    SegReal w;
    seg_of(item, w);
    if (t == 0) prefetch(item + gridDim.x);
    fwd_first_real_tm<P>(t, w, sm, tw_addr, item != blockIdx.x);
    __syncthreads();
    fwd_mid_tm<P>(t, sm, tw_addr);
    __syncwarp();                       // middle passes stay inside the half-warp
    float2 u[P::E];
    centre_real_fwd_keep<P>(t, sm, u);
    float4 pre[kRealPre];
    centre_real_preload<P>(t, AB, pre);  // in flight across the barrier
    __syncthreads();                    // mirror bins live in other threads
    centre_real_mix_keep<P, true>(t, AB, sm, u, pre);
    __syncthreads();                    // every mirror has been read: safe to overwrite
    centre_real_store<P>(t, sm, u);
    __syncwarp();
    inv_mid_tm<P>(t, sm, tw_addr);
    __syncthreads();
    inv_last_real_regs_tm<P>(t, sm, u, tw_addr);
    // register r holds circular samples 2p and 2p+1, p = base + r*S: c_off is even, so they are
    // a pair of outputs -- one 8-byte store when the result row allows it
    int base, npp, bf;
    G::locate(t, 0, base, npp, bf);
    float *o = w.out + (w.out0 + 2 * base - w.c_off);       // dereferenced only where 0 <= index < cnt
    const bool pairs = (w.cnt & 1) == 0 && (((unsigned long long)(w.out + w.out0)) & 7ull) == 0;
    if (pairs) {                        // CTA-uniform
      static_for<0, R>([&](auto rr) {
        constexpr int r = decltype(rr)::value;
        const unsigned tt = (unsigned)(2 * (base + r * G::S) - w.c_off);      // negative -> huge
        if (tt < (unsigned)w.cnt) *reinterpret_cast<float2 *>(o + 2 * r * G::S) = u[r];
      });
    } else {
      static_for<0, R>([&](auto rr) {
        constexpr int r = decltype(rr)::value;
        const unsigned tt = (unsigned)(2 * (base + r * G::S) - w.c_off);
        if (tt < (unsigned)w.cnt) o[2 * r * G::S] = u[r].x;
        if (tt + 1u < (unsigned)w.cnt) o[2 * r * G::S + 1] = u[r].y;
      });
    }
  }
  tmem::fence_before_sync();
  __syncthreads();
  if (t < 32) tmem::free_all(*tmem_base);
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
  {
    const i64 per_signal = REAL ? p.nseg : p.npairs;
    q.per_signal_rcp = 0;
    if (per_signal > 0 && per_signal < (1LL << 31) && p.total + q.prefetch_distance < (1LL << 31))
      q.per_signal_rcp = per_signal == 1 ? 0xffffffffu : (unsigned)((1ULL << 32) / (unsigned long long)per_signal);
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
  const size_t smem = sizeof(float2) * PLAN::SMEM_FLOAT2 + sizeof(float) * kParkFloats + 16;
  auto kernel = os_fused_persist_kernel<PLAN>;
  int dev = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev >= 16) return false;
  if (!configured[dev]) {
    B200_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
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
  B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kernel, op));
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
  // a copy from pageable memory may return before the DMA has landed, and the kernels that
  // read the table run on non-blocking streams: wait once, here
  B200_CUDA_CHECK(cudaDeviceSynchronize());
  g_tw.tab[dev][lg] = dptr;
  return dptr;
}

// Up to 15 extra leading samples per segment make the first useful circular sample a multiple of
// 16 -- taken only when the number of segment pairs stays the same, i.e. when it is free.  The
// last pass holds circular samples n + 16r (or a coarser stride) in register r of the thread at
// position n: with c_off a multiple of 16 the test "is this register an output" is the same for
// every thread of a warp, so the store sequence is entered once instead of once per distinct
// first register (C3, c_off 63 -> 64: two passes through the stores became one), and segment
// inputs and outputs start on 64-byte boundaries of their rows.
static void align_segments(OsParams &op, i64 out_len) {
  const int e = (16 - (op.c_off & 15)) & 15;
  const int step2 = op.step - e;
  if (e == 0 || step2 <= 0) return;
  const i64 pairs1 = ((out_len + op.step - 1) / op.step + 1) / 2;
  const i64 pairs2 = ((out_len + step2 - 1) / step2 + 1) / 2;
  if (pairs2 != pairs1) return;
  op.in_shift -= e;
  op.c_off += e;
  op.step = step2;
}

void b200_os_launch_with(const float2 *tw, float2 *spectrum, int n_fft, int part_taps,
                         const DevJob &job, bool single_segment, cudaStream_t stream) {
  if (job.batch <= 0 || job.out_len <= 0) return;
  const int parts = (int)((job.h_len + part_taps - 1) / part_taps);

  SpecParams sp;
  sp.h = job.h; sp.h_len = job.h_len; sp.reverse = job.reverse;
  sp.part_taps = part_taps; sp.spectrum = spectrum; sp.tw = tw;
  sp.scale = 1.0f / (float)n_fft;

  OsParams op = {};
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
        if (PLAN::N < 16384) align_segments(op, job.out_len);                          \
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

  OsParams op = {};
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

  if (n_fft == 16384 && !getenv("VELES_B200_NO_PERSIST")) {
    using PLAN = P16384;
    static bool configured[16] = {false};
    static int sms[16] = {0};
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    const size_t smem = sizeof(float2) * PLAN::SMEM_FLOAT2 + 16;
    if (dev < 16 && op.total < 2147483647LL) {
      if (!configured[dev]) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(os_fused_real_persist_kernel<PLAN>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        B200_CUDA_CHECK(cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev));
        configured[dev] = true;
      }
      launch_spectrum<PLAN, true>(sp, 1, stream);
      const i64 blocks = op.total < sms[dev] ? op.total : sms[dev];
      b200_prof_begin(B200_PROF_FUSED, stream);
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3((unsigned)blocks);
      cfg.blockDim = dim3(PLAN::T);
      cfg.dynamicSmemBytes = smem;
      cfg.stream = stream;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      B200_CUDA_CHECK(cudaLaunchKernelEx(&cfg, os_fused_real_persist_kernel<PLAN>, op));
      b200_prof_end(B200_PROF_FUSED, stream);
      b200_count_launch(1);
      return;
    }
  }
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
  err = cudaStreamSynchronize(stream);
  if (err != cudaSuccess) return err;       // not ready: the next call tries again
  d.ready = true;     // later launches on other streams are ordered by the caller (see header)
  return cudaSuccess;
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
  b200_prof_suppress(1);      // simd_b200_profile() would create events
  b200_os_launch_with(d.twiddles + (n / 2 - 128), d.spectrum, n, (int)h_length, job, false, stream);
  b200_prof_suppress(0);
  return 0;
}
