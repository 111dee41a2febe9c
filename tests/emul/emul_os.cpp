// tests/emul/emul_os.cpp -- CPU emulation of the fused overlap-save kernel's index and
// twiddle logic (veles/simd_b200/csrc/fft_core.cuh compiled as host code): one loop
// iteration per CUDA thread, one loop nest per barrier interval.  Checks every plan against
// a double-precision direct convolution and reports shared-memory bank conflicts of each
// pass.  Built and run by tests/test_emulation.py (no GPU needed).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../veles/simd_b200/csrc/fft_core.cuh"

using namespace fc;

template <class P>
static void make_twiddles(std::vector<float2> &tw) {
  tw.resize(P::N / 2 > 0 ? P::N / 2 : 1);
  for (int j = 0; j < (int)tw.size(); j++) {
    double a = -2.0 * M_PI * j / P::N;
    tw[j] = make_float2((float)cos(a), (float)sin(a));
  }
}

template <class P, int NB, int R, int MODE>
static int conflicts_of_pass() {
  using G = PassGeom<P, NB, R>;
  int worst = 1;
  for (int j = 0; j < P::E / R; j++)
    for (int r = 0; r < R; r++)
      for (int t0 = 0; t0 < P::T; t0 += 16) {
        int cnt[16] = {0};
        for (int t = t0; t < t0 + 16 && t < P::T; t++) {
          int base, npp, bf;
          if (MODE == 1) G::locate_mid(t, j, base, npp, bf);
          else if (MODE == 2 && P::NPASS == 3) G::locate_centre(t, j, base, npp, bf);
          else G::locate(t, j, base, npp, bf);
          cnt[pad_pos(base + r * G::S) % 16]++;
        }
        for (int b = 0; b < 16; b++) if (cnt[b] > worst) worst = cnt[b];
      }
  return worst;
}

template <class P>
static int run_plan(const char *name, int h_len, int x_len, int reverse) {
  std::vector<float2> tw;
  make_twiddles<P>(tw);
  std::vector<float> x(x_len), h(h_len);
  unsigned s = 12345u + h_len;
  auto rnd = [&]() { s = s * 1664525u + 1013904223u; return ((s >> 8) & 0xFFFF) / 32768.0f - 1.0f; };
  for (auto &v : x) v = rnd();
  for (auto &v : h) v = rnd();

  // filter spectrum
  std::vector<float2> sm(P::SMEM_FLOAT2), Hperm(P::N);
  i64 start = reverse ? h_len - 1 : 0;
  int dir = reverse ? -1 : 1;
  std::vector<TwBases<P>> tb(P::T);
  for (int t = 0; t < P::T; t++) load_twiddle_bases<P>(t, tw.data(), tb[t]);
  for (int t = 0; t < P::T; t++) spec_first<P>(t, h.data(), start, dir, h_len, tb[t], sm.data());
  if (P::NPASS == 3) for (int t = 0; t < P::T; t++) fwd_mid<P>(t, tb[t], sm.data());
  for (int t = 0; t < P::T; t++) centre_store_spectrum<P>(t, Hperm.data(), 1.0f / P::N, sm.data());

  int out_len = x_len + h_len - 1;
  int pad = h_len - 1;
  int step = P::N - h_len + 1;
  int nseg = (out_len + step - 1) / step;
  std::vector<float> out(out_len, NAN);
  for (int sa = 0; sa < nseg; sa += 2) {
    SegPair w;
    w.x = x.data(); w.x_len = x_len;
    w.inA = (i64)sa * step - pad; w.inB = w.inA + step;
    w.haveB = sa + 1 < nseg;
    w.out = out.data();
    w.outA = (i64)sa * step; w.outB = w.outA + step;
    i64 remA = out_len - w.outA, remB = out_len - w.outB;
    w.cntA = (int)(remA < step ? remA : step);
    w.cntB = w.haveB ? (int)(remB < step ? remB : step) : 0;
    w.c_off = h_len - 1;
    w.accumulate = false;
    for (int t = 0; t < P::T; t++) fwd_first<P>(t, w, tb[t], sm.data());
    if constexpr (P::NPASS == 3) {
      // the three middle phases need only warp-level ordering: run them warp by warp, each
      // warp all the way through before the next one starts (and in reverse warp order)
      for (int w0 = (P::T - 1) / 32 * 32; w0 >= 0; w0 -= 32) {
        const int w1 = w0 + 32 < P::T ? w0 + 32 : P::T;
        for (int t = w0; t < w1; t++) fwd_mid<P>(t, tb[t], sm.data());
        for (int t = w0; t < w1; t++) {
          HRegs<P> hr;
          load_spectrum<P>(t, Hperm.data(), hr);
          centre<P>(t, hr, sm.data());
        }
        for (int t = w0; t < w1; t++) inv_mid<P>(t, tb[t], sm.data());
      }
    } else {
      for (int t = 0; t < P::T; t++) {
        HRegs<P> hr;
        load_spectrum<P>(t, Hperm.data(), hr);
        centre<P>(t, hr, sm.data());
      }
    }
    for (int t = 0; t < P::T; t++) inv_last<P>(t, w, tb[t], sm.data());
  }
  // double-precision check
  double num = 0, den = 0, maxabs = 0;
  for (int n = 0; n < out_len; n++) {
    double acc = 0;
    for (int m = 0; m < h_len; m++) {
      int i = n - m;
      if (i >= 0 && i < x_len) acc += (double)(reverse ? h[h_len - 1 - m] : h[m]) * x[i];
    }
    double d = out[n] - acc;
    if (!(out[n] == out[n])) { printf("%s: NaN/unwritten at %d\n", name, n); return 1; }
    num += d * d; den += acc * acc;
    if (fabs(d) > maxabs) maxabs = fabs(d);
  }
  double rel = sqrt(num / den);
  int c1 = conflicts_of_pass<P, P::N, P::R1, 0>();
  int c2 = 1;
  if constexpr (P::NPASS == 3) c2 = conflicts_of_pass<P, P::NB2, P::R2, 1>();
  int c3 = conflicts_of_pass<P, P::R3, P::R3, 2>();
  printf("%-22s N=%5d h=%5d x=%6d rev=%d nseg=%3d rel_l2=%.3e maxabs=%.3e conflicts=%d/%d/%d\n",
         name, P::N, h_len, x_len, reverse, nseg, rel, maxabs, c1, c2, c3);
  return (rel < 2e-6 && c1 == 1 && c2 == 1 && c3 == 1) ? 0 : 1;
}

// Real-packed mode: one real segment of 2N samples per transform.
template <class P>
static int run_plan_real(const char *name, int h_len, int x_len, int reverse, int pad_extra) {
  std::vector<float2> tw;
  make_twiddles<P>(tw);
  std::vector<float> x(x_len), h(h_len);
  unsigned s = 777u + h_len;
  auto rnd = [&]() { s = s * 1664525u + 1013904223u; return ((s >> 8) & 0xFFFF) / 32768.0f - 1.0f; };
  for (auto &v : x) v = rnd();
  for (auto &v : h) v = rnd();
  std::vector<TwBases<P>> tb(P::T);
  for (int t = 0; t < P::T; t++) load_twiddle_bases<P>(t, tw.data(), tb[t]);
  std::vector<float2> sm(P::SMEM_FLOAT2);
  std::vector<float4> AB(P::N);
  i64 start = reverse ? h_len - 1 : 0;
  int dir = reverse ? -1 : 1;
  for (int t = 0; t < P::T; t++) spec_first_real<P>(t, h.data(), start, dir, h_len, tb[t], sm.data());
  if constexpr (P::NPASS == 3) for (int t = 0; t < P::T; t++) fwd_mid<P>(t, tb[t], sm.data());
  for (int t = 0; t < P::T; t++) centre_real_fwd<P>(t, sm.data());
  for (int t = 0; t < P::T; t++) centre_real_coeffs<P>(t, AB.data(), 1.0f / P::N, sm.data());

  const int L = 2 * P::N;
  const int out_len = x_len + h_len - 1;
  const int pad = h_len - 1;
  const int e = pad_extra;                 // extra leading samples (alignment shift)
  const int c_off = h_len - 1 + e;
  const int step = (L - c_off) & ~1;
  const int nseg = (out_len + step - 1) / step;
  std::vector<float> out(out_len, NAN);
  std::vector<float2> u((size_t)P::T * P::E);
  for (int sg = 0; sg < nseg; sg++) {
    SegReal w;
    w.x = x.data(); w.x_len = x_len;
    w.in0 = (i64)sg * step - pad - e;
    w.out = out.data(); w.out0 = (i64)sg * step;
    i64 rem = out_len - w.out0;
    w.cnt = (int)(rem < step ? rem : step);
    w.c_off = c_off; w.accumulate = false;
    for (int t = 0; t < P::T; t++) fwd_first_real<P>(t, w, tb[t], sm.data());
    if constexpr (P::NPASS == 3) for (int t = 0; t < P::T; t++) fwd_mid<P>(t, tb[t], sm.data());
    for (int t = 0; t < P::T; t++) centre_real_fwd_keep<P>(t, sm.data(), u.data() + (size_t)t * P::E);
    for (int t = 0; t < P::T; t++) centre_real_mix_keep<P>(t, AB.data(), sm.data(), u.data() + (size_t)t * P::E);
    for (int t = 0; t < P::T; t++) centre_real_store<P>(t, sm.data(), u.data() + (size_t)t * P::E);
    if constexpr (P::NPASS == 3) for (int t = 0; t < P::T; t++) inv_mid<P>(t, tb[t], sm.data());
    for (int t = 0; t < P::T; t++) inv_last_real<P>(t, w, tb[t], sm.data());
  }
  double num = 0, den = 0;
  for (int n = 0; n < out_len; n++) {
    double acc = 0;
    for (int m = 0; m < h_len; m++) {
      int i = n - m;
      if (i >= 0 && i < x_len) acc += (double)(reverse ? h[h_len - 1 - m] : h[m]) * x[i];
    }
    if (!(out[n] == out[n])) { printf("%s real: NaN/unwritten at %d\n", name, n); return 1; }
    double d = out[n] - acc;
    num += d * d; den += acc * acc;
  }
  double rel = sqrt(num / den);
  printf("%-22s real-packed N=%5d h=%5d x=%6d rev=%d e=%d nseg=%3d rel_l2=%.3e\n", name, P::N, h_len,
         x_len, reverse, e, nseg, rel);
  return rel < 2e-6 ? 0 : 1;
}

int main() {
  int bad = 0;
  bad += run_plan<Plan<256, 16, 16, 1, 16>>("256=16x16", 50, 1024, 0);
  bad += run_plan<Plan<256, 16, 16, 1, 16>>("256=16x16", 64, 777, 1);
  bad += run_plan<Plan<512, 16, 2, 16, 16>>("512=2x16x16", 128, 3000, 0);
  bad += run_plan<Plan<1024, 16, 4, 16, 16>>("1024=4x16x16", 200, 5000, 1);
  bad += run_plan<Plan<2048, 16, 8, 16, 16>>("2048=8x16x16", 512, 9000, 0);
  bad += run_plan<Plan<4096, 16, 16, 16, 16>>("4096=16x16x16", 1000, 20000, 1);
  bad += run_plan<Plan<8192, 32, 32, 16, 16>>("8192=32x16x16", 2048, 30000, 0);
  bad += run_plan<Plan<16384, 32, 32, 32, 16>>("16384=32x32x16", 4096, 60000, 0);
  bad += run_plan<Plan<16384, 32, 32, 32, 16>>("16384=32x32x16", 8192, 40000, 1);
  bad += run_plan_real<Plan<256, 16, 16, 1, 16>>("256=16x16", 100, 3000, 0, 1);
  bad += run_plan_real<Plan<512, 16, 2, 16, 16>>("512=2x16x16", 300, 4000, 1, 0);
  bad += run_plan_real<Plan<2048, 16, 8, 16, 16>>("2048=8x16x16", 1500, 20000, 0, 0);
  bad += run_plan_real<Plan<8192, 32, 32, 16, 16>>("8192=32x16x16", 6000, 50000, 1, 1);
  bad += run_plan_real<Plan<16384, 32, 32, 32, 16>>("16384=32x32x16", 8192, 100000, 0, 1);
  printf(bad ? "EMUL FAIL (%d)\n" : "EMUL OK\n", bad);
  return bad ? 1 : 0;
}
