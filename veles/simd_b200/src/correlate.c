/* correlate.c -- plain-C host layer of the cross-correlation API (B200 build).
 *
 * Same structure as the reference (/root/reference/src/correlate.c): the FFT and
 * overlap-save variants are the convolution handles with reverse = 1 (the device kernels
 * then read h in forward order instead of reversed), the brute-force variant is the direct
 * kernel in correlation form.
 */
#define LIBSIMD_IMPLEMENTATION
#include <simd/correlate.h>
#include <simd/convolve.h>
#include <simd/b200.h>
#include <stdio.h>
#include <stdlib.h>
#include "../csrc/b200_internal.h"

/* reference: src/correlate.c:37-42 */
CrossCorrelationFFTHandle cross_correlate_fft_initialize(size_t xLength, size_t hLength) {
  CrossCorrelationFFTHandle handle = convolve_fft_initialize(xLength, hLength);
  handle.reverse = 1;
  return handle;
}

/* reference: src/correlate.c:44-48 */
void cross_correlate_fft(CrossCorrelationFFTHandle handle, const float *x, const float *h,
                         float *result) {
  convolve_fft(handle, x, h, result);
}

/* reference: src/correlate.c:50-52 */
void cross_correlate_fft_finalize(CrossCorrelationFFTHandle handle) {
  convolve_fft_finalize(handle);
}

/* reference: src/correlate.c:54-60 */
CrossCorrelationOverlapSaveHandle cross_correlate_overlap_save_initialize(size_t xLength,
                                                                          size_t hLength) {
  CrossCorrelationOverlapSaveHandle handle =
      convolve_overlap_save_initialize(xLength, hLength);
  handle.reverse = 1;
  return handle;
}

/* reference: src/correlate.c:62-67 */
void cross_correlate_overlap_save(CrossCorrelationOverlapSaveHandle handle,
                                  const float *__restrict x, const float *__restrict h,
                                  float *result) {
  convolve_overlap_save(handle, x, h, result);
}

/* reference: src/correlate.c:69-72 */
void cross_correlate_overlap_save_finalize(CrossCorrelationOverlapSaveHandle handle) {
  convolve_overlap_save_finalize(handle);
}

/* reference: src/correlate.c:74-126 */
void cross_correlate_simd(int simd, const float *__restrict x, size_t xLength,
                          const float *__restrict h, size_t hLength,
                          float *__restrict result) {
  (void)simd;
  if (!x || !h || !result || xLength == 0 || hLength == 0) {
    fprintf(stderr, "libSimd(b200): cross_correlate_simd: null pointer or zero length\n");
    abort();
  }
  B200Job job;
  job.x = x; job.x_len = (long long)xLength; job.x_stride = (long long)xLength;
  job.h = h; job.h_len = (long long)hLength;
  job.out = result; job.out_len = (long long)(xLength + hLength - 1);
  job.out_stride = job.out_len;
  job.pad = (long long)hLength - 1;
  job.batch = 1;
  job.reverse = 1;
  b200_run(NULL, B200_ALG_DIRECT, &job);
}

/* reference: src/correlate.c:128-142 */
CrossCorrelationHandle cross_correlate_initialize(size_t xLength, size_t hLength) {
  CrossCorrelationHandle handle = convolve_initialize(xLength, hLength);
  switch (handle.algorithm) {
    case kConvolutionAlgorithmFFT:
      handle.handle.fft.reverse = 1;
      break;
    case kConvolutionAlgorithmOverlapSave:
      handle.handle.os.reverse = 1;
      break;
    case kConvolutionAlgorithmBruteForce:
      break;
  }
  return handle;
}

/* reference: src/correlate.c:144-156 */
void cross_correlate(CrossCorrelationHandle handle, const float *__restrict x,
                     const float *__restrict h, float *__restrict result) {
  switch (handle.algorithm) {
    case kConvolutionAlgorithmFFT:
    case kConvolutionAlgorithmOverlapSave:
      convolve(handle, x, h, result);
      break;
    case kConvolutionAlgorithmBruteForce:
      cross_correlate_simd(1, x, (size_t)handle.x_length, h, (size_t)handle.h_length,
                           result);
      break;
  }
}

/* reference: src/correlate.c:158-160 */
void cross_correlate_finalize(CrossCorrelationHandle handle) {
  convolve_finalize(handle);
}
