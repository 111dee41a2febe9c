// fft_core.cuh -- in-register / shared-memory FFT building blocks for the fused
// overlap-save kernel (os_fft.cu).  Replaces the FFTF calls of the reference
// (/root/reference/src/convolve.c:131-143,175,198,216) with a hand-written transform whose
// spectrum never leaves the SM.
//
// Everything here is __host__ __device__ and takes the thread index as an argument so that
// tests/emul/emul_os.cpp can execute the very same index / twiddle logic on the CPU
// (one loop iteration per CUDA thread, one loop nest per barrier interval).
//
// Transform structure for a complex length N = R1*R2*R3 (R2 may be 1):
//   forward  : decimation in frequency.  A pass over sub-blocks of length NB with radix R
//              and stride S = NB/R does an R-point DFT over the strided index, applies the
//              twiddle W_NB^(n''*k) and writes the results back to the positions they were
//              read from.  After the last pass (S = 1) bin k1 + R1*k2 + R1*R2*k3 sits at
//              position k1*N/R1 + k2*R3 + k3 ("digit-reversed").
//   inverse  : the same passes in the opposite order with conjugated twiddles applied before
//              the butterfly (decimation in time): digit-reversed in, natural order out.
//   Forward-last and inverse-first work on the same R3 contiguous elements, so the pointwise
//   product with the filter spectrum happens in registers between them, and the filter
//   spectrum is stored in exactly the register order of that pass (no reordering pass
//   anywhere; 1/N is folded into it).
//
// Two real segments are transformed at once as z = a + i*b: filtering is linear with a real
// impulse response, so IFFT(FFT(z) * H) = (a (*) h) + i*(b (*) h) and no Hermitian
// untangling step is needed (the reference runs one real FFT pair per segment,
// src/convolve.c:198-216).
//
// Shared-memory layout: float2 per complex sample, logical position p stored at
// p + (p >> 4) (one pad slot per 16) which keeps every pass's half-warp accesses on 16
// distinct 8-byte bank pairs.
#pragma once

#include <type_traits>
#include <math.h>

#if defined(__CUDACC__)
#define FC_HD __host__ __device__ __forceinline__
#else
#define FC_HD inline __attribute__((always_inline))
#ifndef FC_HOST_FLOAT2
#define FC_HOST_FLOAT2
struct float2 { float x, y; };
static inline float2 make_float2(float a, float b) { float2 r; r.x = a; r.y = b; return r; }
struct float4 { float x, y, z, w; };
static inline float4 make_float4(float a, float b, float c, float d) { float4 r; r.x = a; r.y = b; r.z = c; r.w = d; return r; }
#endif
#endif

namespace fc {

typedef long long i64;

// Before the first pass of a pair overwrites the transform buffer: the bulk store that is
// still reading the previous pair's staged results out of it must have finished reading.
FC_HD void wait_staged_results_read(int t) {
#if defined(__CUDA_ARCH__)
  if (t == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  __syncthreads();
#else
  (void)t;
#endif
}

template <int I, int END, class F>
FC_HD void static_for(F &&f) {
  if constexpr (I < END) {
    f(std::integral_constant<int, I>{});
    static_for<I + 1, END>(f);
  }
}

FC_HD constexpr int pad_pos(int p) { return p + (p >> 4); }
constexpr int padded_len(int n) { return n + (n >> 4); }

// cos(2*pi*k/32) for k in [0,8]; other octants by symmetry.
FC_HD constexpr float cos32_tab(int k) {
  return k == 0 ? 1.0f : k == 1 ? 0.98078528040323044913f : k == 2 ? 0.92387953251128675613f
       : k == 3 ? 0.83146961230254523708f : k == 4 ? 0.70710678118654752440f
       : k == 5 ? 0.55557023301960222474f : k == 6 ? 0.38268343236508977173f
       : k == 7 ? 0.19509032201612826785f : 0.0f;
}
FC_HD constexpr float cos32(int k) { return k <= 8 ? cos32_tab(k) : -cos32_tab(16 - k); }  // k in [0,16]
FC_HD constexpr float sin32(int k) { return k <= 8 ? cos32_tab(8 - k) : cos32_tab(k - 8); }

FC_HD constexpr int bitrev(int v, int bits) {
  int r = 0;
  for (int i = 0; i < bits; i++) r |= ((v >> i) & 1) << (bits - 1 - i);
  return r;
}
FC_HD constexpr int ilog2(int v) { int l = 0; while ((1 << l) < v) l++; return l; }
FC_HD constexpr int hibit(int v) { int h = 1; while (h * 2 <= v) h *= 2; return h; }

FC_HD float2 cmul(float2 a, float2 b) {
  return make_float2(fmaf(a.x, b.x, -(a.y * b.y)), fmaf(a.x, b.y, a.y * b.x));
}
FC_HD float2 cmul_conj(float2 a, float2 b) {   // a * conj(b)
  return make_float2(fmaf(a.x, b.x, a.y * b.y), fmaf(a.y, b.x, -(a.x * b.y)));
}

// One radix-2 DIT butterfly with compile-time twiddle w = W_32^K32 (forward:
// exp(-i*2*pi*K32/32); inverse: its conjugate).
//   a' = a + w*b ;  b' = a - w*b = 2a - a'      (6 FMA-class operations)
template <int K32, bool INV>
FC_HD void bfly(float2 &a, float2 &b) {
  if constexpr (K32 == 0) {
    float2 t = b;
    b = make_float2(a.x - t.x, a.y - t.y);
    a = make_float2(a.x + t.x, a.y + t.y);
  } else if constexpr (K32 == 8) {
    // forward w = -i : w*b = (b.y, -b.x);  inverse w = +i : w*b = (-b.y, b.x)
    float2 t = INV ? make_float2(-b.y, b.x) : make_float2(b.y, -b.x);
    b = make_float2(a.x - t.x, a.y - t.y);
    a = make_float2(a.x + t.x, a.y + t.y);
  } else {
    constexpr float wr = cos32(K32);
    constexpr float wi = INV ? sin32(K32) : -sin32(K32);
    // explicit fmaf: the compiler otherwise rewrites 2*a as a+a and splits the last FMA
    const float nr = fmaf(-b.y, wi, fmaf(b.x, wr, a.x));
    const float ni = fmaf(b.y, wr, fmaf(b.x, wi, a.y));
    b = make_float2(fmaf(2.0f, a.x, -nr), fmaf(2.0f, a.y, -ni));
    a = make_float2(nr, ni);
  }
}

// R-point DFT in registers, R in {2,4,8,16,32}.  Decimation in time: input element r must
// sit in v[bitrev(r)], output bin k is produced in v[k].  INV -> un-normalised inverse.
template <int R, bool INV>
FC_HD void dft_inreg(float2 *v) {
  static_for<1, ilog2(R) + 1>([&](auto s) {
    constexpr int LEN = 1 << decltype(s)::value;
    static_for<0, R / 2>([&](auto i) {
      constexpr int I = decltype(i)::value;
      constexpr int blk = (I / (LEN / 2)) * LEN;
      constexpr int j = I % (LEN / 2);
      bfly<j * (32 / LEN), INV>(v[blk + j], v[blk + j + LEN / 2]);
    });
  });
}

// Powers of a per-thread twiddle base w: w^k for k in [1,R).  Two-level scheme
// (w^(4a+b) = w^(4a) * w^b) so that only R/4+2 powers are live at once and every power is at
// most log2(R) multiplications away from the tabulated base.
template <int R>
struct Powers {
  float2 lo[4];
  float2 hi[R / 4 > 0 ? R / 4 : 1];
  FC_HD void init(float2 w) {
    lo[1] = w;
    if constexpr (R > 2) {
      lo[2] = cmul(w, w);
      lo[3] = cmul(lo[2], w);
    }
    if constexpr (R > 4) {
      hi[1] = cmul(lo[2], lo[2]);
      static_for<2, R / 4>([&](auto a) {
        constexpr int A = decltype(a)::value;
        constexpr int H = hibit(A);
        if constexpr (H == A) hi[A] = cmul(hi[A / 2], hi[A / 2]);
        else hi[A] = cmul(hi[H], hi[A - H]);
      });
    }
  }
  template <int K>
  FC_HD float2 get() const {
    constexpr int a = K >> 2, b = K & 3;
    if constexpr (a == 0) return lo[b];
    else if constexpr (b == 0) return hi[a];
    else return cmul(hi[a], lo[b]);
  }
};

// ------------------------------------------------------------------------------------------
// Plan: N complex points held by T = N/E threads, E points per thread in registers.
template <int N_, int E_, int R1_, int R2_, int R3_>
struct Plan {
  static constexpr int N = N_, E = E_, R1 = R1_, R2 = R2_, R3 = R3_;
  static constexpr int T = N / E;
  static constexpr int NPASS = R2 > 1 ? 3 : 2;
  static constexpr int NB2 = N / R1;          // block length seen by the 2nd pass
  static constexpr int NBF3 = N / R3;         // butterflies in the last (contiguous) pass
  static constexpr int SMEM_FLOAT2 = padded_len(N);
  static_assert(R1 * R2 * R3 == N, "radices must multiply to N");
  static_assert(E % R1 == 0 && E % R2 == 0 && E % R3 == 0, "E must be a multiple of each radix");
  static_assert(NPASS == 3 || R3 == NB2, "two-pass plan: R3 = N/R1");
  static_assert(NPASS == 2 || (R3 == 16 && R2 % 16 == 0 && T % 16 == 0 && (N / R1) == R2 * 16),
                "three-pass plans: 16-point centre, half-warp-closed middle passes");
};

// Where butterfly number j of thread t lives for a pass with block length NB and radix R.
template <class P, int NB, int R>
struct PassGeom {
  static constexpr int S = NB / R;            // stride between a butterfly's elements
  static_assert(S == 1 || S % 16 == 0, "strides are 1 or multiples of the pad period");
  FC_HD static void locate(int t, int j, int &base, int &npp, int &bf) {
    bf = t + j * P::T;
    int blk = bf / S;
    npp = bf % S;
    base = blk * NB + npp;
  }
  // Middle-pass geometry of 3-pass plans, arranged so that everything a half-warp touches
  // between the first and the last pass stays inside that half-warp: after the first pass
  // the R1 sub-blocks of NB2 = R2*16 points are independent transforms, and half-warp g
  // (threads 16g..16g+15) owns sub-blocks g*HB .. g*HB+HB-1 (HB = E/R2) through the second
  // forward pass, the centre and the first inverse pass.  Only warp-level synchronisation
  // is needed between those three phases, so warps drift apart and one warp's shared-memory
  // traffic overlaps another's arithmetic.
  FC_HD static void locate_mid(int t, int j, int &base, int &npp, int &bf) {
    // pass over sub-blocks of NB = NB2 with radix R = R2, S = 16: butterfly j of thread t
    const int blk = (t >> 4) * (P::E / P::R2) + j;
    npp = t & 15;
    bf = blk * S + npp;
    base = blk * NB + npp;
  }
  FC_HD static void locate_centre(int t, int j, int &base, int &npp, int &bf) {
    // contiguous 16-point butterflies: sub-block has R2 of them, R2/16 per thread
    constexpr int PER = P::R2 / 16;
    const int blk = (t >> 4) * (P::E / P::R2) + j / PER;
    bf = blk * P::R2 + (t & 15) + 16 * (j % PER);
    npp = 0;
    base = bf * NB;
  }
  // Padded shared-memory offset of element r relative to pad_pos(base): a compile-time
  // constant, because r*S is a multiple of 16 (S % 16 == 0) or stays inside base's 16-slot
  // row (S == 1, base % 16 == 0, r < 16).
  FC_HD static constexpr int off(int r) { return S == 1 ? r : r * S + (r * S) / 16; }
};

// Description of one unit of work: a pair of overlap-save segments of one signal.
struct SegPair {
  const float *x;      // signal base
  i64 x_len;           // valid samples; X(i) = 0 outside [0, x_len)
  i64 inA;             // X index of segment A's sample 0 (may be negative)
  i64 inB;             // same for segment B
  bool haveB;
  float *out;          // result base of this signal
  i64 outA, outB;      // out index receiving c[c_off] of each segment
  int cntA, cntB;      // number of valid outputs of each segment (<= step), 0 = none
  int c_off;           // first useful circular-convolution sample (h_len - 1 in OS)
  bool accumulate;     // add to out instead of overwriting (partitioned impulse response)
};

FC_HD float load_x(const float *x, i64 x_len, i64 i) {
  return (i >= 0 && i < x_len) ? x[i] : 0.0f;
}

// Per-thread twiddle bases of the strided passes, loaded once per kernel: the first and
// last pass share W_N^(n''), the two middle passes share W_NB2^(n'').
template <class P>
struct TwBases {
  float2 w1[P::E / P::R1];
  float2 w2[P::E / P::R2];
};

template <class P>
FC_HD void load_twiddle_bases(int t, const float2 *tw, TwBases<P> &tb) {
  static_for<0, P::E / P::R1>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, npp, bf;
    PassGeom<P, P::N, P::R1>::locate(t, j, base, npp, bf);
    tb.w1[j] = tw[npp];
  });
  if constexpr (P::NPASS == 3) {
    static_for<0, P::E / P::R2>([&](auto jj) {
      constexpr int j = decltype(jj)::value;
      int base, npp, bf;
      PassGeom<P, P::NB2, P::R2>::locate_mid(t, j, base, npp, bf);
      tb.w2[j] = tw[npp * P::R1];
    });
  }
}

// Filter spectrum in the register order of the centre pass.  Layout of Hperm: the bin held in
// register k of butterfly j of thread t is Hperm[(k*T + t)*JB + j], JB = E/R3 butterflies per
// thread, so a thread fetches its JB bins of one k with a single 8*JB-byte load and a warp's
// loads are contiguous.
template <class P>
struct HRegs {
  float2 h[P::E];      // [j*R3 + k]
};

template <class P>
FC_HD void load_spectrum(int t, const float2 *Hperm, HRegs<P> &hr) {
  constexpr int R = P::R3, JB = P::E / R;
  static_assert(JB == 1 || JB == 2, "one or two centre butterflies per thread");
  static_for<0, R>([&](auto kk) {
    constexpr int k = decltype(kk)::value;
    if constexpr (JB == 2) {
#if defined(__CUDA_ARCH__)
      const float4 q = *reinterpret_cast<const float4 *>(Hperm + (size_t)(k * P::T + t) * 2);
      hr.h[k] = make_float2(q.x, q.y);
      hr.h[R + k] = make_float2(q.z, q.w);
#else
      hr.h[k] = Hperm[(size_t)(k * P::T + t) * 2];
      hr.h[R + k] = Hperm[(size_t)(k * P::T + t) * 2 + 1];
#endif
    } else {
      hr.h[k] = Hperm[k * P::T + t];
    }
  });
}

// forward first pass: global -> registers -> R1-point DFT -> twiddle -> shared memory
template <class P>
FC_HD void fwd_first(int t, const SegPair &w, const TwBases<P> &tb, float2 *sm) {
  constexpr int R = P::R1, NB = P::N;
  using G = PassGeom<P, NB, R>;
  // CTA-uniform: both segments lie entirely inside the signal (no bounds checks needed)
  const bool fast = w.haveB && w.inA >= 0 && w.inB + P::N <= w.x_len;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, npp, bf;
    G::locate(t, j, base, npp, bf);
    auto *smb = sm + pad_pos(base);
    float2 v[R];
    if (fast) {
      const float *xa = w.x + (w.inA + base);
      const float *xb = w.x + (w.inB + base);
      static_for<0, R>([&](auto rr) {
        constexpr int r = decltype(rr)::value;
        v[bitrev(r, ilog2(R))] = make_float2(xa[r * G::S], xb[r * G::S]);
      });
    } else {
      static_for<0, R>([&](auto rr) {
        constexpr int r = decltype(rr)::value;
        const int p = base + r * G::S;
        const float a = load_x(w.x, w.x_len, w.inA + p);
        const float b = w.haveB ? load_x(w.x, w.x_len, w.inB + p) : 0.0f;
        v[bitrev(r, ilog2(R))] = make_float2(a, b);
      });
    }
    dft_inreg<R, false>(v);
    if constexpr (G::S > 1) {
      Powers<R> pw;
      pw.init(tb.w1[j]);
      static_for<1, R>([&](auto kk) {
        constexpr int k = decltype(kk)::value;
        v[k] = cmul(v[k], pw.template get<k>());
      });
    }
    static_for<0, R>([&](auto kk) {
      constexpr int k = decltype(kk)::value;
      smb[G::off(k)] = v[k];
    });
  });
}

// forward middle pass (3-pass plans only): shared -> DFT_R2 -> twiddle -> shared (in place)
template <class P>
FC_HD void fwd_mid(int t, const TwBases<P> &tb, float2 *sm) {
  constexpr int R = P::R2, NB = P::NB2;
  using G = PassGeom<P, NB, R>;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, npp, bf;
    G::locate_mid(t, j, base, npp, bf);
    auto *smb = sm + pad_pos(base);
    float2 v[R];
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      v[bitrev(r, ilog2(R))] = smb[G::off(r)];
    });
    dft_inreg<R, false>(v);
    Powers<R> pw;
    pw.init(tb.w2[j]);
    static_for<1, R>([&](auto kk) {
      constexpr int k = decltype(kk)::value;
      v[k] = cmul(v[k], pw.template get<k>());
    });
    static_for<0, R>([&](auto kk) {
      constexpr int k = decltype(kk)::value;
      smb[G::off(k)] = v[k];
    });
  });
}

// centre: last forward pass, pointwise product with the filter spectrum (already in
// registers, see load_spectrum), first inverse pass.
template <class P>
FC_HD void centre(int t, const HRegs<P> &hr, float2 *sm) {
  constexpr int R = P::R3;
  static_assert(R == 16, "contiguous pass assumes 16-slot padded rows");
  using G = PassGeom<P, R, R>;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, npp, bf;
    if constexpr (P::NPASS == 3) G::locate_centre(t, j, base, npp, bf); else G::locate(t, j, base, npp, bf);
    auto *smb = sm + pad_pos(base);
    float2 v[R], u[R];
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      v[bitrev(r, ilog2(R))] = smb[G::off(r)];
    });
    dft_inreg<R, false>(v);
    static_for<0, R>([&](auto kk) {
      constexpr int k = decltype(kk)::value;
      u[bitrev(k, ilog2(R))] = cmul(v[k], hr.h[j * R + k]);
    });
    dft_inreg<R, true>(u);
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      smb[G::off(r)] = u[r];
    });
  });
}

// spectrum variant of the centre: last forward pass, then write bins * scale to Hperm.
template <class P>
FC_HD void centre_store_spectrum(int t, float2 *Hperm, float scale, const float2 *sm) {
  constexpr int R = P::R3;
  using G = PassGeom<P, R, R>;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, npp, bf;
    if constexpr (P::NPASS == 3) G::locate_centre(t, j, base, npp, bf); else G::locate(t, j, base, npp, bf);
    auto *smb = sm + pad_pos(base);
    float2 v[R];
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      v[bitrev(r, ilog2(R))] = smb[G::off(r)];
    });
    dft_inreg<R, false>(v);
    static_for<0, R>([&](auto kk) {
      constexpr int k = decltype(kk)::value;
      Hperm[(size_t)(k * P::T + t) * (P::E / R) + j] = make_float2(v[k].x * scale, v[k].y * scale);
    });
  });
}

// inverse middle pass (3-pass plans only): shared -> conj twiddle -> IDFT_R2 -> shared
template <class P>
FC_HD void inv_mid(int t, const TwBases<P> &tb, float2 *sm) {
  constexpr int R = P::R2, NB = P::NB2;
  using G = PassGeom<P, NB, R>;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, npp, bf;
    G::locate_mid(t, j, base, npp, bf);
    auto *smb = sm + pad_pos(base);
    float2 v[R];
    Powers<R> pw;
    pw.init(tb.w2[j]);
    static_for<0, R>([&](auto kk) {
      constexpr int k = decltype(kk)::value;
      float2 e = smb[G::off(k)];
      if constexpr (k > 0) e = cmul_conj(e, pw.template get<k>());
      v[bitrev(k, ilog2(R))] = e;
    });
    dft_inreg<R, true>(v);
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      smb[G::off(r)] = v[r];
    });
  });
}

// inverse last pass: shared -> conj twiddle -> IDFT_R1 -> global (real part -> segment A,
// imaginary part -> segment B), keeping only circular samples >= c_off.
template <class P>
FC_HD void inv_last(int t, const SegPair &w, const TwBases<P> &tb, const float2 *sm) {
  constexpr int R = P::R1, NB = P::N;
  using G = PassGeom<P, NB, R>;
  // CTA-uniform: both segments deliver all N - c_off samples and overwrite
  const bool fast = !w.accumulate && w.cntA == P::N - w.c_off && w.cntB == w.cntA;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, npp, bf;
    G::locate(t, j, base, npp, bf);
    auto *smb = sm + pad_pos(base);
    float2 v[R];
    if constexpr (G::S > 1) {
      Powers<R> pw;
      pw.init(tb.w1[j]);
      static_for<0, R>([&](auto kk) {
        constexpr int k = decltype(kk)::value;
        float2 e = smb[G::off(k)];
        if constexpr (k > 0) e = cmul_conj(e, pw.template get<k>());
        v[bitrev(k, ilog2(R))] = e;
      });
    } else {
      static_for<0, R>([&](auto kk) {
        constexpr int k = decltype(kk)::value;
        v[bitrev(k, ilog2(R))] = smb[G::off(k)];
      });
    }
    dft_inreg<R, true>(v);
    // natural-order circular index of register r is base + r*S; output index tt = it - c_off
    const int t0 = base - w.c_off;
    float *oa = w.out + (w.outA + t0);     // only dereferenced where 0 <= t0 + r*S < cnt
    float *ob = w.out + (w.outB + t0);
    if (fast) {
      static_for<0, R>([&](auto rr) {
        constexpr int r = decltype(rr)::value;
        if (t0 + r * G::S >= 0) {
          oa[r * G::S] = v[r].x;
          ob[r * G::S] = v[r].y;
        }
      });
    } else {
      static_for<0, R>([&](auto rr) {
        constexpr int r = decltype(rr)::value;
        const unsigned tt = (unsigned)(t0 + r * G::S);     // negative -> huge
        if (tt < (unsigned)w.cntA) oa[r * G::S] = w.accumulate ? oa[r * G::S] + v[r].x : v[r].x;
        if (tt < (unsigned)w.cntB) ob[r * G::S] = w.accumulate ? ob[r * G::S] + v[r].y : v[r].y;
      });
    }
  });
}

// Results of a full segment pair into a shared-memory staging area (two float arrays, the
// element for output index tt at stage[phase + tt]) from where a bulk copy takes them to HBM.
template <class P>
FC_HD void stage_outputs(int t, const float2 *v, int c_off, int step, int phase, float *stA,
                         float *stB) {
  constexpr int R = P::R1, NB = P::N;
  using G = PassGeom<P, NB, R>;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, npp, bf;
    G::locate(t, j, base, npp, bf);
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      const unsigned tt = (unsigned)(base + r * G::S - c_off);      // negative -> huge
      if (tt < (unsigned)step) {
        stA[phase + tt] = v[j * R + r].x;
        stB[phase + tt] = v[j * R + r].y;
      }
    });
  });
}

// Guarded direct stores of the same register contents (edge pairs, accumulation).
template <class P>
FC_HD void store_outputs(int t, const SegPair &w, const float2 *v) {
  constexpr int R = P::R1, NB = P::N;
  using G = PassGeom<P, NB, R>;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, npp, bf;
    G::locate(t, j, base, npp, bf);
    const int t0 = base - w.c_off;
    float *oa = w.out + (w.outA + t0);
    float *ob = w.out + (w.outB + t0);
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      const unsigned tt = (unsigned)(t0 + r * G::S);
      if (tt < (unsigned)w.cntA) oa[r * G::S] = w.accumulate ? oa[r * G::S] + v[j * R + r].x : v[j * R + r].x;
      if (tt < (unsigned)w.cntB) ob[r * G::S] = w.accumulate ? ob[r * G::S] + v[j * R + r].y : v[j * R + r].y;
    });
  });
}

// Filter-spectrum first pass: zero-padded real impulse response gf[m] = h[start + dir*m]
// (m < taps) -> registers -> DFT -> twiddle -> shared.
template <class P>
FC_HD void spec_first(int t, const float *h, i64 start, int dir, int taps,
                      const TwBases<P> &tb, float2 *sm) {
  constexpr int R = P::R1, NB = P::N;
  using G = PassGeom<P, NB, R>;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, npp, bf;
    G::locate(t, j, base, npp, bf);
    auto *smb = sm + pad_pos(base);
    float2 v[R];
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      int p = base + r * G::S;
      float a = p < taps ? h[start + (i64)dir * p] : 0.0f;
      v[bitrev(r, ilog2(R))] = make_float2(a, 0.0f);
    });
    dft_inreg<R, false>(v);
    if constexpr (G::S > 1) {
      Powers<R> pw;
      pw.init(tb.w1[j]);
      static_for<1, R>([&](auto kk) {
        constexpr int k = decltype(kk)::value;
        v[k] = cmul(v[k], pw.template get<k>());
      });
    }
    static_for<0, R>([&](auto kk) {
      constexpr int k = decltype(kk)::value;
      smb[G::off(k)] = v[k];
    });
  });
}

// ==========================================================================================
// Real-packed ("mode B") variant: ONE real segment of L = 2N samples per N-point complex
// transform, z[p] = a[2p] + i*a[2p+1].  Filtering a real sequence with a real impulse
// response is a real-linear map of z, so in the frequency domain it is widely linear:
//     Z'[k] = alpha[k] * Z[k] + beta[k] * conj(Z[(N-k) mod N])
// with  E = Zh[k] + conj(Zh[N-k]),  T = -i * exp(-i*pi*k/N) * (Zh[k] - conj(Zh[N-k])),
//       alpha = (E - T*sin(pi*k/N)) / 2,   beta = i * T * cos(pi*k/N) / 2,
// Zh = FFT_N of the impulse response packed the same way (checked numerically against a
// length-L real FFT convolution; derivation in DESIGN.md).  Compared with the paired mode
// this yields 2N - taps + 1 outputs per transform instead of 2(N - taps + 1): it pays when
// the impulse response is long relative to N, at the price of one more shared-memory
// exchange (each bin needs its mirror bin, which lives in another thread).
// ==========================================================================================

struct SegReal {
  const float *x;      // signal base
  i64 x_len;
  i64 in0;             // X index of the segment's sample 0 (may be negative)
  float *out;
  i64 out0;            // out index receiving circular sample c_off
  int cnt;             // valid outputs of this segment
  int c_off;
  bool accumulate;
};

// Position of the mirror bin N-k of the bin held in register k3 of centre butterfly
// (k1 = sub-block, k2 = butterfly inside it).  Bin index k = k1 + R1*k2 + R1*R2*k3.
template <class P>
FC_HD int mirror_pos(int k1, int k2, int k3) {
  constexpr int R2e = P::NPASS == 3 ? P::R2 : 1;
  int m1, m2, m3;
  if (k1 > 0) { m1 = P::R1 - k1; m2 = R2e - 1 - k2; m3 = 15 - k3; }
  else if (k2 > 0) { m1 = 0; m2 = R2e - k2; m3 = 15 - k3; }
  else { m1 = 0; m2 = 0; m3 = (16 - k3) & 15; }
  return (m1 * R2e + m2) * 16 + m3;
}

template <class P>
FC_HD void centre_locate(int t, int j, int &base, int &bf, int &k1, int &k2) {
  using G = PassGeom<P, P::R3, P::R3>;
  int npp;
  if constexpr (P::NPASS == 3) {
    G::locate_centre(t, j, base, npp, bf);
    k1 = bf / P::R2;
    k2 = bf % P::R2;
  } else {
    G::locate(t, j, base, npp, bf);
    k1 = bf;
    k2 = 0;
  }
}

// first pass, real-packed input
template <class P>
FC_HD void fwd_first_real(int t, const SegReal &w, const TwBases<P> &tb, float2 *sm) {
  constexpr int R = P::R1, NB = P::N;
  using G = PassGeom<P, NB, R>;
  const bool inside = w.in0 >= 0 && w.in0 + 2 * P::N <= w.x_len;
  // 8-byte loads need an even sample offset from an 8-byte aligned base
  const bool fast = inside && (((unsigned long long)(w.x + w.in0)) & 7ull) == 0;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, npp, bf;
    G::locate(t, j, base, npp, bf);
    auto *smb = sm + pad_pos(base);
    float2 v[R];
    if (fast) {
      const float2 *xa = reinterpret_cast<const float2 *>(w.x + w.in0) + base;
      static_for<0, R>([&](auto rr) {
        constexpr int r = decltype(rr)::value;
        v[bitrev(r, ilog2(R))] = xa[r * G::S];
      });
    } else if (inside) {
      // inside the signal but not 8-byte aligned: two 4-byte loads, no bounds checks
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
    if constexpr (G::S > 1) {
      Powers<R> pw;
      pw.init(tb.w1[j]);
      static_for<1, R>([&](auto kk) {
        constexpr int k = decltype(kk)::value;
        v[k] = cmul(v[k], pw.template get<k>());
      });
    }
    static_for<0, R>([&](auto kk) {
      constexpr int k = decltype(kk)::value;
      smb[G::off(k)] = v[k];
    });
  });
}

// centre, step 1: last forward pass, bins written back in place (mirror bins are needed next)
template <class P>
FC_HD void centre_real_fwd(int t, float2 *sm) {
  constexpr int R = P::R3;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, bf, k1, k2;
    centre_locate<P>(t, j, base, bf, k1, k2);
    auto *smb = sm + pad_pos(base);
    float2 v[R];
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      v[bitrev(r, ilog2(R))] = smb[r];
    });
    dft_inreg<R, false>(v);
    static_for<0, R>([&](auto kk) {
      constexpr int k = decltype(kk)::value;
      smb[k] = v[k];
    });
  });
}

// centre, step 2 (after a CTA barrier): Z' = alpha*Z + beta*conj(Z_mirror), then the first
// inverse pass in registers.  Results stay in u[] until every thread has read its mirrors.
// AB layout: float4 {alpha, beta} of register k of butterfly j of thread t at (k*T + t)*JB + j.
template <class P>
FC_HD void centre_real_mix(int t, const float4 *AB, const float2 *sm, float2 *u) {
  constexpr int R = P::R3, JB = P::E / R;
  static_for<0, JB>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, bf, k1, k2;
    centre_locate<P>(t, j, base, bf, k1, k2);
    const float2 *smb = sm + pad_pos(base);
    static_for<0, R>([&](auto kk) {
      constexpr int k = decltype(kk)::value;
      const float2 z = smb[k];
      const float2 zm = sm[pad_pos(mirror_pos<P>(k1, k2, k))];
      const float4 ab = AB[(size_t)(k * P::T + t) * JB + j];
      float2 r = cmul(z, make_float2(ab.x, ab.y));
      // + beta * conj(zm)
      r.x = fmaf(ab.z, zm.x, fmaf(ab.w, zm.y, r.x));
      r.y = fmaf(ab.w, zm.x, fmaf(-ab.z, zm.y, r.y));
      u[j * R + bitrev(k, ilog2(R))] = r;
    });
    dft_inreg<R, true>(u + j * R);
  });
}

// centre of the real-packed transform with this thread's own bins kept in registers across the
// barrier (the plain kernel writes them to shared memory for the mirror reads of other threads
// AND reads its own back afterwards: 32 shared-memory loads per thread that this form saves).
// step 1: last forward pass; bins to shared memory (others read them as mirrors) and to z[]
template <class P>
FC_HD void centre_real_fwd_keep(int t, float2 *sm, float2 *z) {
  constexpr int R = P::R3;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, bf, k1, k2;
    centre_locate<P>(t, j, base, bf, k1, k2);
    auto *smb = sm + pad_pos(base);
    float2 *v = z + j * R;
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      v[bitrev(r, ilog2(R))] = smb[r];
    });
    dft_inreg<R, false>(v);
    static_for<0, R>([&](auto kk) {
      constexpr int k = decltype(kk)::value;
      smb[k] = v[k];
    });
  });
}
// step 2 (after a CTA barrier): Z' = alpha*Z + beta*conj(Z_mirror) in place, then the first
// inverse pass; z[] holds the results in natural order afterwards
// The coefficients of the first kPre registers of butterfly 0 can be fetched BEFORE the barrier
// (they do not depend on it) and handed in through `pre`: their L2 latency then hides behind
// the barrier wait instead of opening the phase.
constexpr int kRealPre = 8;
template <class P>
FC_HD void centre_real_preload(int t, const float4 *AB, float4 *pre) {
  constexpr int JB = P::E / P::R3;
  static_for<0, kRealPre>([&](auto kk) {
    constexpr int k = decltype(kk)::value;
    pre[k] = AB[(size_t)(k * P::T + t) * JB];
  });
}
template <class P, bool PRE = false>
FC_HD void centre_real_mix_keep(int t, const float4 *AB, const float2 *sm, float2 *z, const float4 *pre = nullptr) {
  constexpr int R = P::R3, JB = P::E / R;
  static_for<0, JB>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, bf, k1, k2;
    centre_locate<P>(t, j, base, bf, k1, k2);
    float2 u[R];
    static_for<0, R>([&](auto kk) {
      constexpr int k = decltype(kk)::value;
      const float2 own = z[j * R + k];
      const float2 zm = sm[pad_pos(mirror_pos<P>(k1, k2, k))];
      float4 ab;
      if constexpr (PRE && j == 0 && k < kRealPre) ab = pre[k];
      else ab = AB[(size_t)(k * P::T + t) * JB + j];
      float2 r = cmul(own, make_float2(ab.x, ab.y));
      r.x = fmaf(ab.z, zm.x, fmaf(ab.w, zm.y, r.x));          // + beta * conj(zm)
      r.y = fmaf(ab.w, zm.x, fmaf(-ab.z, zm.y, r.y));
      u[bitrev(k, ilog2(R))] = r;
    });
    dft_inreg<R, true>(u);
    static_for<0, R>([&](auto kk) {
      constexpr int k = decltype(kk)::value;
      z[j * R + k] = u[k];
    });
  });
}

// centre, step 3 (after a second CTA barrier): write the first inverse pass back
template <class P>
FC_HD void centre_real_store(int t, float2 *sm, const float2 *u) {
  constexpr int R = P::R3;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, bf, k1, k2;
    centre_locate<P>(t, j, base, bf, k1, k2);
    auto *smb = sm + pad_pos(base);
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      smb[r] = u[j * R + r];
    });
  });
}

// last pass, real-packed output: register r holds circular samples 2p and 2p+1, p = base+r*S
template <class P>
FC_HD void inv_last_real(int t, const SegReal &w, const TwBases<P> &tb, const float2 *sm) {
  constexpr int R = P::R1, NB = P::N;
  using G = PassGeom<P, NB, R>;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, npp, bf;
    G::locate(t, j, base, npp, bf);
    auto *smb = sm + pad_pos(base);
    float2 v[R];
    if constexpr (G::S > 1) {
      Powers<R> pw;
      pw.init(tb.w1[j]);
      static_for<0, R>([&](auto kk) {
        constexpr int k = decltype(kk)::value;
        float2 e = smb[G::off(k)];
        if constexpr (k > 0) e = cmul_conj(e, pw.template get<k>());
        v[bitrev(k, ilog2(R))] = e;
      });
    } else {
      static_for<0, R>([&](auto kk) {
        constexpr int k = decltype(kk)::value;
        v[bitrev(k, ilog2(R))] = smb[G::off(k)];
      });
    }
    dft_inreg<R, true>(v);
    const int t0 = 2 * base - w.c_off;          // output index of circular sample 2*base
    float *o = w.out + (w.out0 + t0);           // dereferenced only where 0 <= index < cnt
    // CTA-uniform: whole pairs, 8-byte aligned, overwrite
    const bool fast = !w.accumulate && (w.c_off & 1) == 0 && (w.cnt & 1) == 0 &&
                      (((unsigned long long)(w.out + w.out0)) & 7ull) == 0;
    if (fast) {
      static_for<0, R>([&](auto rr) {
        constexpr int r = decltype(rr)::value;
        const unsigned tt = (unsigned)(t0 + 2 * r * G::S);
        if (tt < (unsigned)w.cnt) *reinterpret_cast<float2 *>(o + 2 * r * G::S) = v[r];
      });
    } else {
      static_for<0, R>([&](auto rr) {
        constexpr int r = decltype(rr)::value;
        const unsigned tt = (unsigned)(t0 + 2 * r * G::S);
        float *q = o + 2 * r * G::S;
        if (tt < (unsigned)w.cnt) q[0] = w.accumulate ? q[0] + v[r].x : v[r].x;
        if (tt + 1u < (unsigned)w.cnt) q[1] = w.accumulate ? q[1] + v[r].y : v[r].y;
      });
    }
  });
}

// spectrum side: first pass over the packed, zero-padded impulse response gf[m] = h[start+dir*m]
template <class P>
FC_HD void spec_first_real(int t, const float *h, i64 start, int dir, int taps,
                           const TwBases<P> &tb, float2 *sm) {
  constexpr int R = P::R1, NB = P::N;
  using G = PassGeom<P, NB, R>;
  static_for<0, P::E / R>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, npp, bf;
    G::locate(t, j, base, npp, bf);
    auto *smb = sm + pad_pos(base);
    float2 v[R];
    static_for<0, R>([&](auto rr) {
      constexpr int r = decltype(rr)::value;
      const int m = 2 * (base + r * G::S);
      const float a = m < taps ? h[start + (i64)dir * m] : 0.0f;
      const float b = m + 1 < taps ? h[start + (i64)dir * (m + 1)] : 0.0f;
      v[bitrev(r, ilog2(R))] = make_float2(a, b);
    });
    dft_inreg<R, false>(v);
    if constexpr (G::S > 1) {
      Powers<R> pw;
      pw.init(tb.w1[j]);
      static_for<1, R>([&](auto kk) {
        constexpr int k = decltype(kk)::value;
        v[k] = cmul(v[k], pw.template get<k>());
      });
    }
    static_for<0, R>([&](auto kk) {
      constexpr int k = decltype(kk)::value;
      smb[G::off(k)] = v[k];
    });
  });
}

// spectrum side, after centre_real_fwd + CTA barrier: alpha / beta of every bin
template <class P>
FC_HD void centre_real_coeffs(int t, float4 *AB, float scale, const float2 *sm) {
  constexpr int R = P::R3, JB = P::E / R;
  constexpr int R2e = P::NPASS == 3 ? P::R2 : 1;
  static_for<0, JB>([&](auto jj) {
    constexpr int j = decltype(jj)::value;
    int base, bf, k1, k2;
    centre_locate<P>(t, j, base, bf, k1, k2);
    const float2 *smb = sm + pad_pos(base);
    static_for<0, R>([&](auto kk) {
      constexpr int k3 = decltype(kk)::value;
      const float2 z = smb[k3];
      const float2 zm = sm[pad_pos(mirror_pos<P>(k1, k2, k3))];
      const int kbin = k1 + P::R1 * (k2 + R2e * k3);
      float sn, cs;
#if defined(__CUDA_ARCH__)
      sincospif((float)kbin / (float)P::N, &sn, &cs);
#else
      sn = (float)sin(3.14159265358979323846 * kbin / P::N);
      cs = (float)cos(3.14159265358979323846 * kbin / P::N);
#endif
      const float2 e = make_float2(z.x + zm.x, z.y - zm.y);      // Z + conj(Zm)
      const float2 d = make_float2(z.x - zm.x, z.y + zm.y);      // Z - conj(Zm)
      // T = -i * (cs - i*sn) * d = (-sn - i*cs) * d
      const float2 tt = cmul(make_float2(-sn, -cs), d);
      const float2 al = make_float2(0.5f * scale * (e.x - tt.x * sn), 0.5f * scale * (e.y - tt.y * sn));
      // beta = i * T * cs / 2
      const float2 be = make_float2(-0.5f * scale * tt.y * cs, 0.5f * scale * tt.x * cs);
      AB[(size_t)(k3 * P::T + t) * JB + j] = make_float4(al.x, al.y, be.x, be.y);
    });
  });
}

}  // namespace fc
