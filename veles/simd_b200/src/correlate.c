/* correlate.c -- cross-correlation entry points of the B200 build.
 *
 * Cross-correlation is convolution with the impulse response read backwards, and on the
 * GPU "backwards" is only an index direction inside the kernels' tap loads.  So every
 * function here is a convolution handle whose `reverse` flag is raised; nothing is copied
 * or flipped on the host.  The ten symbols and their contracts are those of
 * /root/reference/inc/simd/correlate.h (implemented there in src/correlate.c:37-160).
 */
#define LIBSIMD_IMPLEMENTATION
#include <simd/correlate.h>
#include <simd/convolve.h>
#include <simd/b200.h>
#include <stdio.h>
#include <stdlib.h>
#include "../csrc/b200_internal.h"

/* ---- explicit transform variants: initialise as convolution, then flag ------------------ */

/* src/correlate.c:37-42 */
CrossCorrelationFFTHandle cross_correlate_fft_initialize(size_t xLength, size_t hLength) {
  CrossCorrelationFFTHandle whole_signal = convolve_fft_initialize(xLength, hLength);
  whole_signal.reverse = 1;
  return whole_signal;
}

/* src/correlate.c:54-60 */
CrossCorrelationOverlapSaveHandle cross_correlate_overlap_save_initialize(size_t xLength,
                                                                          size_t hLength) {
  CrossCorrelationOverlapSaveHandle segmented = convolve_overlap_save_initialize(xLength, hLength);
  segmented.reverse = 1;
  return segmented;
}

/* src/correlate.c:44-48 and :62-67 -- the flag travels inside the by-value struct */
void cross_correlate_fft(CrossCorrelationFFTHandle handle, const float *x, const float *h,
                         float *result) {
  convolve_fft(handle, x, h, result);
}

void cross_correlate_overlap_save(CrossCorrelationOverlapSaveHandle handle,
                                  const float *__restrict x, const float *__restrict h,
                                  float *result) {
  convolve_overlap_save(handle, x, h, result);
}

/* src/correlate.c:50-52 and :69-72 */
void cross_correlate_fft_finalize(CrossCorrelationFFTHandle handle) {
  convolve_fft_finalize(handle);
}

void cross_correlate_overlap_save_finalize(CrossCorrelationOverlapSaveHandle handle) {
  convolve_overlap_save_finalize(handle);
}

/* ---- direct kernel in correlation form (src/correlate.c:74-126) ------------------------- */

void cross_correlate_simd(int simd, const float *__restrict x, size_t xLength,
                          const float *__restrict h, size_t hLength,
                          float *__restrict result) {
  (void)simd;                      /* both values run the GPU kernel */
  if (x == NULL || h == NULL || result == NULL || xLength == 0 || hLength == 0) {
    fprintf(stderr, "libSimd(b200): precondition failed: cross_correlate_simd needs non-null "
                    "pointers and non-zero lengths\n");
    abort();
  }
  const long long outputs = (long long)(xLength + hLength - 1);
  B200Job job = {
    .x = x, .x_len = (long long)xLength, .x_stride = (long long)xLength,
    .h = h, .h_len = (long long)hLength,
    .out = result, .out_len = outputs, .out_stride = outputs,
    .pad = (long long)hLength - 1, .batch = 1,
    .reverse = 1,                  /* taps are read in forward order: correlation */
  };
  b200_run(NULL, B200_ALG_DIRECT, &job);
}

/* ---- best-method handle (src/correlate.c:128-160) --------------------------------------- */

CrossCorrelationHandle cross_correlate_initialize(size_t xLength, size_t hLength) {
  CrossCorrelationHandle best = convolve_initialize(xLength, hLength);
  /* the inner handle of the chosen transform carries the flag; the direct kernel has no
   * inner handle and is told at call time */
  if (best.algorithm == kConvolutionAlgorithmFFT) best.handle.fft.reverse = 1;
  if (best.algorithm == kConvolutionAlgorithmOverlapSave) best.handle.os.reverse = 1;
  return best;
}

void cross_correlate(CrossCorrelationHandle handle, const float *__restrict x,
                     const float *__restrict h, float *__restrict result) {
  if (handle.algorithm == kConvolutionAlgorithmBruteForce) {
    cross_correlate_simd(1, x, (size_t)handle.x_length, h, (size_t)handle.h_length, result);
  } else {
    convolve(handle, x, h, result);      /* reverse flag already set on the inner handle */
  }
}

void cross_correlate_finalize(CrossCorrelationHandle handle) { convolve_finalize(handle); }

/* ---- additive: correlate and keep only the extrema (simd/b200.h) ------------------------ */

void b200_detect_peaks(const float *data, long long size, int type, void **results_host,
                       size_t *count);

void simd_b200_cross_correlate_peaks(CrossCorrelationHandle handle, const float *x,
                                     const float *h, ExtremumType type,
                                     ExtremumPoint **results, size_t *resultsLength) {
  if (!x || !h || !results || !resultsLength) {
    fprintf(stderr, "libSimd(b200): simd_b200_cross_correlate_peaks: null argument\n");
    abort();
  }
  const size_t n = (size_t)handle.x_length + (size_t)handle.h_length - 1;
  /* the correlation result never leaves the GPU: x and h go up, {position, value} pairs
   * come down (reference chain: src/correlate.c:144-156 then src/detect_peaks.c:58-127) */
  float *d_result = b200_device_alloc(n);
  cross_correlate(handle, x, h, d_result);
  void *host = NULL;
  *resultsLength = 0;
  if (n > 2) b200_detect_peaks(d_result, (long long)n, (int)type, &host, resultsLength);
  *results = (ExtremumPoint *)host;
  b200_device_free(d_result);
}
