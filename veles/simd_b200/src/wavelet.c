/* wavelet.c -- plain-C host layer of the wavelet module (B200 build): table lookup, argument
 * checks and the host-memory helpers of /root/reference/src/wavelet.c; the transform itself
 * runs in csrc/dwt.cu.  Every entry point, including the *_na ones, goes to the GPU:
 * there is no CPU path. */
#define LIBSIMD_IMPLEMENTATION
#include <simd/wavelet.h>
#include <simd/memory.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "wavelet_tables.h"

#define B200_WAVELET_MAX_ORDER 76

/* csrc/dwt.cu.  out[i] = sum_{k < order} f[k] * S(step*i + k*dilation), step = 2 when
 * decimate (then dilation = 1), else 1. */
void b200_wavelet_run(const float *src, long long length, float *desthi, float *destlo,
                      const float *highpass, const float *lowpass, int order, int dilation,
                      int decimate, int ext);

static void die(const char *what) {
  fprintf(stderr, "libSimd(b200): wavelet: %s\n", what);
  abort();
}

/* smallest power of two >= length (the reference's aligned_length, src/wavelet.c:54-62) */
static size_t pow2_ceil(size_t length) {
  size_t p = 1;
  while (p < length) p <<= 1;
  return p;
}

/* the table row of (type, order), or NULL (row selection: src/wavelet.c:190-206) */
static const float *table_row(WaveletType type, int order) {
  if (order < 2) return NULL;
  switch (type) {
    case WAVELET_TYPE_DAUBECHIES:
      return (order <= 76 && order % 2 == 0) ? b200_daubechies[order / 2 - 1] : NULL;
    case WAVELET_TYPE_SYMLET:
      return (order <= 76 && order % 2 == 0) ? b200_symlets[order / 2 - 1] : NULL;
    case WAVELET_TYPE_COIFLET:
      return (order <= 30 && order % 6 == 0) ? b200_coiflets[order / 6 - 1] : NULL;
    default:
      return NULL;
  }
}

/* reference: src/wavelet.c:83-98 (an unsigned compare and a divisibility test: 0 passes) */
int wavelet_validate_order(WaveletType type, int order) {
  if (order == 0) {
    return type == WAVELET_TYPE_DAUBECHIES || type == WAVELET_TYPE_COIFLET ||
           type == WAVELET_TYPE_SYMLET;
  }
  return table_row(type, order) != NULL;
}

/* reference: src/wavelet.c:185-206 */
int simd_b200_wavelet_filters(WaveletType type, int order, float *highpass, float *lowpass) {
  const float *row = table_row(type, order);
  if (!row) return 0;
  for (int i = 0; i < order; i++) {
    lowpass[i] = row[i];
    highpass[order - 1 - i] = (i & 1) ? row[i] : -row[i];
  }
  return 1;
}

static void check_signal(const float *src, size_t length, const float *hi, const float *lo,
                         int need_even) {
  if (!src || !hi || !lo) die("null array");
  if (length == 0) die("empty signal");
  if (need_even && (length & 1)) die("the signal length must be even");
  if (length > ((size_t)1 << 40)) die("signal too long");
}

/* reference: src/wavelet.c:100-119 (AVX layout) and :64-81 */
float *wavelet_prepare_array(int order, const float *src, size_t length) {
  if (!src) die("null array");
  if (length == 0 || (length & 1)) die("the signal length must be even and positive");
  const size_t alen = pow2_ceil(length);
  const size_t total = alen * (order > 4 ? 4 : 2);
  float *res = mallocf(total);
  if (!res) die("out of memory");
  memset(res, 0, total * sizeof(float));
  memcpy(res, src, length * sizeof(float));
  if (length > 8) {
    /* shifted copies; the reference copies alen - 8 floats even when that runs past the end
     * of src -- here the copy stops at the last sample and the rest stays zero */
    const int shifts = order > 4 ? 3 : 1;
    for (int s = 1; s <= shifts; s++) {
      const size_t from = 2 * (size_t)s;
      size_t count = alen - 8;
      if (from + count > length) count = length - from;
      memcpy(res + s * alen - 8 * (size_t)(s - 1), src + from, count * sizeof(float));
    }
  }
  return res;
}

/* reference: src/wavelet.c:121-136 */
float *wavelet_allocate_destination(int order, size_t sourceLength) {
  if (sourceLength == 0 || sourceLength % 4 != 0) die("the source length must be a positive multiple of 4");
  float *res = mallocf(pow2_ceil(sourceLength) * (order > 4 ? 2 : 1));
  if (!res) die("out of memory");
  return res;
}

/* reference: src/wavelet.c:138-164 */
void wavelet_recycle_source(int order, float *src, size_t length, float **desthihi,
                            float **desthilo, float **destlohi, float **destlolo) {
  if (length == 0 || length % 4 != 0) {
    *desthihi = *desthilo = *destlohi = *destlolo = NULL;
    return;
  }
  size_t lq = pow2_ceil(length) / (order > 4 ? 1 : 2);
  if (length < 16) lq = length / 4;
  *desthihi = src;
  *desthilo = src + lq;
  *destlohi = src + lq * 2;
  *destlolo = src + lq * 3;
}

/* reference: src/wavelet.c:271-323 */
void wavelet_apply_na(WaveletType type, int order, ExtensionType ext, const float *__restrict src,
                      size_t length, float *__restrict desthi, float *__restrict destlo) {
  float hp[B200_WAVELET_MAX_ORDER], lp[B200_WAVELET_MAX_ORDER];
  check_signal(src, length, desthi, destlo, 1);
  if ((int)ext < 0 || (int)ext > (int)EXTENSION_TYPE_ZERO) die("unknown extension type");
  if (!simd_b200_wavelet_filters(type, order, hp, lp)) {
    die("unsupported order (Daubechies / Symlets: 2..76 even; Coiflets: 6, 12, 18, 24, 30)");
  }
  b200_wavelet_run(src, (long long)length, desthi, destlo, hp, lp, order, 1, 1, (int)ext);
}

/* reference: src/wavelet.c:1878-1905 (a switch over hand-unrolled orders, same result) */
void wavelet_apply(WaveletType type, int order, ExtensionType ext, const float *__restrict src,
                   size_t length, float *__restrict desthi, float *__restrict destlo) {
  wavelet_apply_na(type, order, ext, src, length, desthi, destlo);
}

/* reference: src/wavelet.c:325-382, filter layout :208-246 */
void stationary_wavelet_apply_na(WaveletType type, int order, int level, ExtensionType ext,
                                 const float *__restrict src, size_t length,
                                 float *__restrict desthi, float *__restrict destlo) {
  float hp[B200_WAVELET_MAX_ORDER], lp[B200_WAVELET_MAX_ORDER];
  check_signal(src, length, desthi, destlo, 0);
  if ((int)ext < 0 || (int)ext > (int)EXTENSION_TYPE_ZERO) die("unknown extension type");
  if (level < 1 || level > 24) die("the decomposition level must be in 1..24");
  if (!simd_b200_wavelet_filters(type, order, hp, lp)) {
    die("unsupported order (Daubechies / Symlets: 2..76 even; Coiflets: 6, 12, 18, 24, 30)");
  }
  b200_wavelet_run(src, (long long)length, desthi, destlo, hp, lp, order, 1 << (level - 1), 0,
                   (int)ext);
}

/* reference: src/wavelet.c:1907-1940 */
void stationary_wavelet_apply(WaveletType type, int order, int level, ExtensionType ext,
                              const float *__restrict src, size_t length,
                              float *__restrict desthi, float *__restrict destlo) {
  stationary_wavelet_apply_na(type, order, level, ext, src, length, desthi, destlo);
}
