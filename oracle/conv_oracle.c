/* oracle/conv_oracle.c -- TEST INFRASTRUCTURE, not product code.
 *
 * CPU restatement ("port") of the veles.simd 1D convolution / cross-correlation path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference legs may
 * load this; the product library (libSimd.so) never links or calls it.
 *
 * Every function cites the reference lines it follows (paths relative to /root/reference).
 * Parity status: PINNED -- checked against (i) the reference's two known-answer tests
 * (tests/convolve.cc:53-71, tests/correlate.cc:53-71), (ii) the reference's own sources
 * compiled here into oracle/_ref/libSimd_ref.so (brute-force paths bit-exact with simd=0;
 * overlap-save / FFT paths to rel-L2 <= 1e-6, both run over the same FFT stand-in because
 * FFTF itself is absent), (iii) the committed fixtures in tests/golden/ that were generated
 * from oracle/_ref by oracle/make_golden.py.
 *
 * Deliberate deviations from the reference:
 *   Q1  overlap-save first block: samples past x_length are treated as zero instead of
 *       being read from beyond the caller's buffer (src/convolve.c:196).
 *   Q2  the FFT method zero-pads X and H on every call (the reference zeroes only at
 *       init, src/convolve.c:254-257, so its handle is single-use).
 */
#include <assert.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "fft_pow2.h"

#define ORACLE_API __attribute__((visibility("default")))

enum { ORACLE_ALG_BRUTE = 0, ORACLE_ALG_FFT = 1, ORACLE_ALG_OS = 2 };

/* src/convolve.c:40-101, scalar branch (:95-97): y[n] = sum_{m=beg}^{end-1} h[m]*x[n-m],
 * float accumulation in ascending m.  Bit-exact with convolve_simd(0, ...). */
ORACLE_API void oracle_convolve_brute(const float *x, size_t x_length,
                                      const float *h, size_t h_length, float *result) {
  assert(x && h && result && x_length > 0 && h_length > 0);
  for (long n = 0; n < (long)(x_length + h_length - 1); n++) {
    float sum = 0.f;
    long beg = n < (long)x_length ? 0 : n - (long)x_length + 1;   /* :51 */
    long end = n + 1;                                             /* :52-55 */
    if (end > (long)h_length) end = (long)h_length;
    for (long m = beg; m < end; m++) sum += h[m] * x[n - m];      /* :95-97 */
    result[n] = sum;                                              /* :99 */
  }
}

/* src/correlate.c:74-126, scalar branch (:120-122):
 * result[h_length-1-n] = sum_m x[m]*h[n+m], n from h_length-1 down to -(x_length-1). */
ORACLE_API void oracle_cross_correlate_brute(const float *x, size_t x_length,
                                             const float *h, size_t h_length,
                                             float *result) {
  assert(x && h && result && x_length > 0 && h_length > 0);
  for (long n = (long)h_length - 1; n > -(long)x_length; n--) {   /* :78 */
    float sum = 0.f;
    long beg = n <= 0 ? -n : 0;                                   /* :80 */
    long end = -n + (long)h_length;                               /* :81-84 */
    if (end > (long)x_length) end = (long)x_length;
    for (long m = beg; m < end; m++) sum += x[m] * h[n + m];      /* :120-122 */
    result[-n + (long)h_length - 1] = sum;                        /* :124 */
  }
}

/* Double-precision anchor (not in the reference): exact-to-rounding linear convolution,
 * optionally with time-reversed h (correlation, identity pinned by tests/correlate.cc). */
ORACLE_API void oracle_convolve_f64(const float *x, size_t x_length, const float *h,
                                    size_t h_length, int reverse, double *result) {
  for (long n = 0; n < (long)(x_length + h_length - 1); n++) {
    double sum = 0.0;
    long beg = n < (long)x_length ? 0 : n - (long)x_length + 1;
    long end = n + 1 > (long)h_length ? (long)h_length : n + 1;
    for (long m = beg; m < end; m++) {
      double hv = reverse ? h[h_length - 1 - m] : h[m];
      sum += hv * (double)x[n - m];
    }
    result[n] = sum;
  }
}

/* src/convolve.c:116-121: L = 4 * 2^floor(log2 h)  ("log" starts at 2). */
ORACLE_API int oracle_overlap_save_length(size_t h_length) {
  int L = (int)h_length, lg = 2;
  while (L >>= 1) lg++;
  return 1 << lg;
}

/* src/convolve.c:237-244: next power of two >= x+h-1. */
ORACLE_API int oracle_fft_length(size_t x_length, size_t h_length) {
  int M = (int)(x_length + h_length - 1);
  if ((M & (M - 1)) != 0) {
    int lg = 1;
    while (M >>= 1) lg++;
    M = 1 << lg;
  }
  return M;
}

/* src/convolve.c:349-363 (x86 branch): handle-time selector of the reference. */
ORACLE_API int oracle_select_algorithm(size_t x_length, size_t h_length) {
  if (x_length > h_length * 2) {
    return x_length > 200 ? ORACLE_ALG_OS : ORACLE_ALG_BRUTE;
  }
  return x_length > 350 ? ORACLE_ALG_FFT : ORACLE_ALG_BRUTE;
}

/* interleaved complex multiply a*=b over nbins bins: inc/simd/arithmetic.h:142-150 */
static void spectrum_multiply(float *a, const float *b, int nbins) {
  for (int k = 0; k < nbins; k++) {
    float re = a[2 * k] * b[2 * k] - a[2 * k + 1] * b[2 * k + 1];
    float im = a[2 * k] * b[2 * k + 1] + a[2 * k + 1] * b[2 * k];
    a[2 * k] = re; a[2 * k + 1] = im;
  }
}

/* src/convolve.c:156-229.  reverse != 0 -> h is time-reversed first (:167-168), which is
 * how cross_correlate_overlap_save works (src/correlate.c:54-66). */
ORACLE_API void oracle_convolve_overlap_save(const float *x, size_t x_length,
                                             const float *h, size_t h_length,
                                             int reverse, float *result) {
  assert(x && h && result);
  assert(h_length < x_length / 2);                 /* :105 */
  size_t M = h_length;
  int L = oracle_overlap_save_length(h_length);
  OracleFFT *fft = oracle_fft_create(L);
  float *H = (float *)calloc(L + 2, sizeof(float));
  float *buf = (float *)calloc(L + 2, sizeof(float));
  for (size_t i = 0; i < M; i++) buf[i] = reverse ? h[M - 1 - i] : h[i];   /* :167-172 */
  oracle_rfft_forward(fft, buf, H);                                        /* :175-176 */
  size_t step = (size_t)L - (M - 1);                                       /* :178 */
  size_t out_len = x_length + M - 1;
  for (size_t i = 0; i < out_len; i += step) {                             /* :181 */
    /* gather X[i-(M-1) .. i-(M-1)+L) of the zero-extended signal (:186-196) */
    for (int k = 0; k < L; k++) {
      long idx = (long)i - (long)(M - 1) + k;
      buf[k] = (idx >= 0 && idx < (long)x_length) ? x[idx] : 0.f;          /* Q1 */
    }
    oracle_rfft_forward(fft, buf, buf);                                    /* :198 */
    spectrum_multiply(buf, H, L / 2 + 1);                                  /* :201-213 */
    oracle_rfft_backward(fft, buf, buf);                                   /* :216 */
    float scale = 1.0f / L;                                                /* :218-219 */
    size_t cnt = (i + step < out_len + 1) ? step : out_len - i;            /* :221-227 */
    if (cnt > out_len - i) cnt = out_len - i;
    for (size_t k = 0; k < cnt; k++) result[i + k] = buf[M - 1 + k] * scale;
  }
  free(H); free(buf);
  oracle_fft_destroy(fft);
}

/* src/convolve.c:289-326 with a fresh zero-padded X/H per call (Q2). */
ORACLE_API void oracle_convolve_fft(const float *x, size_t x_length, const float *h,
                                    size_t h_length, int reverse, float *result) {
  assert(x && h && result && x_length > 0 && h_length > 0);
  int M = oracle_fft_length(x_length, h_length);
  if (M < 2) M = 2;
  OracleFFT *fft = oracle_fft_create(M);
  float *X = (float *)calloc(M + 2, sizeof(float));
  float *H = (float *)calloc(M + 2, sizeof(float));
  memcpy(X, x, x_length * sizeof(float));                                  /* :301 */
  for (size_t i = 0; i < h_length; i++) H[i] = reverse ? h[h_length - 1 - i] : h[i];
  oracle_rfft_forward(fft, X, X);                                          /* :309 */
  oracle_rfft_forward(fft, H, H);
  spectrum_multiply(X, H, M / 2 + 1);                                      /* :311-320 */
  oracle_rfft_backward(fft, X, X);                                         /* :323 */
  float scale = 1.0f / M;                                                  /* :325 */
  for (size_t i = 0; i < x_length + h_length - 1; i++) result[i] = X[i] * scale;
  free(X); free(H);
  oracle_fft_destroy(fft);
}

/* src/convolve.c:381-395 after :328-366 -- the auto-selecting handle path. */
ORACLE_API void oracle_convolve(const float *x, size_t x_length, const float *h,
                                size_t h_length, float *result) {
  switch (oracle_select_algorithm(x_length, h_length)) {
    case ORACLE_ALG_FFT: oracle_convolve_fft(x, x_length, h, h_length, 0, result); break;
    case ORACLE_ALG_OS: oracle_convolve_overlap_save(x, x_length, h, h_length, 0, result); break;
    default: oracle_convolve_brute(x, x_length, h, h_length, result); break;
  }
}

/* src/correlate.c:128-156 -- FFT/OS with reverse=1, brute force through its own kernel. */
ORACLE_API void oracle_cross_correlate(const float *x, size_t x_length, const float *h,
                                       size_t h_length, float *result) {
  switch (oracle_select_algorithm(x_length, h_length)) {
    case ORACLE_ALG_FFT: oracle_convolve_fft(x, x_length, h, h_length, 1, result); break;
    case ORACLE_ALG_OS: oracle_convolve_overlap_save(x, x_length, h, h_length, 1, result); break;
    default: oracle_cross_correlate_brute(x, x_length, h, h_length, result); break;
  }
}

/* Batch = the caller's loop over convolve() in the reference (no batched entry point
 * exists there); one independent "handle" per signal, signals spread over `threads`
 * host threads for the timed CPU baseline.  Returns the thread count actually used. */
ORACLE_API int oracle_convolve_batch(const float *x, size_t x_length, size_t x_stride,
                                     const float *h, size_t h_length, float *result,
                                     size_t result_stride, size_t batch, int reverse,
                                     int threads) {
  int used = 1;
#ifdef _OPENMP
  if (threads <= 0) threads = omp_get_max_threads();
  used = threads;
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads)
#endif
  for (long b = 0; b < (long)batch; b++) {
    if (reverse) oracle_cross_correlate(x + b * x_stride, x_length, h, h_length, result + b * result_stride);
    else oracle_convolve(x + b * x_stride, x_length, h, h_length, result + b * result_stride);
  }
  return used;
}
