/* oracle/wavelet_oracle.c -- TEST INFRASTRUCTURE, not product code.
 *
 * CPU restatement of the reference's scalar wavelet loops, used by tests/ and by bench.py's
 * cpu_baseline leg as the checker for the GPU kernels in veles/simd_b200/csrc/dwt.cu.  Nothing
 * in the shipped library calls into this file.
 *
 * The filter table row is an argument (the `order` lowpass taps of the wavelet); the tests
 * pass the row the product uses and pin that row separately against the reference's own
 * tables.  The arithmetic follows the reference operation for operation (float accumulator,
 * taps in ascending order, multiply then add; this file is compiled with -ffp-contract=off
 * like the ISO-C build of the reference), so with an identical row the output equals
 * wavelet_apply_na / stationary_wavelet_apply_na bit for bit.
 */
#include <stddef.h>
#include <stdlib.h>

#define ORACLE_API __attribute__((visibility("default")))

enum { EXT_PERIODIC = 0, EXT_MIRROR = 1, EXT_CONSTANT = 2, EXT_ZERO = 3 };

/* /root/reference/src/wavelet.c:248-269 (initialize_extension): the value `e` samples past
 * the end of the signal */
static float beyond_end(int ext, const float *src, size_t length, size_t e) {
  switch (ext) {
    case EXT_PERIODIC: return src[e % length];
    case EXT_MIRROR:   return src[length - 1 - (e % length)];
    case EXT_CONSTANT: return src[length - 1];
    default:           return 0.f;
  }
}

/* /root/reference/src/wavelet.c:185-206 (initialize_highpass_lowpass): the highpass filter is
 * the lowpass one reversed with alternating sign */
static void filter_pair(const float *row, int order, float *highpass, float *lowpass) {
  for (int i = 0; i < order; i++) {
    lowpass[i] = row[i];
    highpass[order - 1 - i] = (i & 1) ? row[i] : -row[i];
  }
}

/* /root/reference/src/wavelet.c:271-323 (wavelet_apply_na); its three branches (length !=
 * order, order == 8, otherwise) perform the same sums */
ORACLE_API void oracle_wavelet_apply(const float *row, int order, int ext, const float *src,
                                     size_t length, float *desthi, float *destlo) {
  float *hp = malloc(sizeof(float) * order), *lp = malloc(sizeof(float) * order);
  filter_pair(row, order, hp, lp);
  for (size_t i = 0, di = 0; i < length; i += 2, di++) {
    float reshi = 0.f, reslo = 0.f;
    for (int j = 0; j < order; j++) {
      size_t n = i + (size_t)j;
      float v = n < length ? src[n] : beyond_end(ext, src, length, n - length);
      reshi += hp[j] * v;
      reslo += lp[j] * v;
    }
    desthi[di] = reshi;
    destlo[di] = reslo;
  }
  free(hp);
  free(lp);
}

/* /root/reference/src/wavelet.c:325-382 (stationary_wavelet_apply_na) with the dilated filter
 * pair of :208-246: taps sit at multiples of stride = 2^(level-1), the loop visits only those */
ORACLE_API void oracle_stationary_wavelet_apply(const float *row, int order, int level, int ext,
                                                const float *src, size_t length, float *desthi,
                                                float *destlo) {
  const size_t stride = (size_t)1 << (level - 1);
  float *hp = malloc(sizeof(float) * order), *lp = malloc(sizeof(float) * order);
  filter_pair(row, order, hp, lp);
  for (size_t i = 0; i < length; i++) {
    float reshi = 0.f, reslo = 0.f;
    for (int k = 0; k < order; k++) {
      size_t n = i + (size_t)k * stride;
      float v = n < length ? src[n] : beyond_end(ext, src, length, n - length);
      reshi += hp[k] * v;
      reslo += lp[k] * v;
    }
    desthi[i] = reshi;
    destlo[i] = reslo;
  }
  free(hp);
  free(lp);
}
