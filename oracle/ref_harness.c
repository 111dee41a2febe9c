/* oracle/ref_harness.c -- TEST INFRASTRUCTURE, not product code.
 * Thin helpers linked into oracle/_ref/libSimd_ref.so next to the reference's own,
 * unmodified src/convolve.c, src/correlate.c and src/memory.c.  They only loop the
 * reference's public API (inc/simd/convolve.h, inc/simd/correlate.h) so that the timed
 * CPU baseline can use every host core: the reference has no batched entry point and is
 * single-threaded on this path (src/convolve.c:179-180), so "all cores" means one handle
 * per thread over independent signals. */
#include <stddef.h>
#include <simd/convolve.h>
#include <simd/correlate.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define REF_API __attribute__((visibility("default")))

/* handle path: convolve_initialize -> convolve -> convolve_finalize per signal.
 * algorithm: -1 = reference selector, 0 = brute force (AVX, simd=1). */
REF_API int ref_convolve_batch(const float *x, size_t x_length, size_t x_stride,
                               const float *h, size_t h_length, float *result,
                               size_t result_stride, size_t batch, int reverse,
                               int algorithm, int threads) {
  int used = 1;
#ifdef _OPENMP
  if (threads <= 0) threads = omp_get_max_threads();
  used = threads;
#pragma omp parallel num_threads(threads)
#endif
  {
    ConvolutionHandle handle;
    int have = 0;
    if (algorithm < 0) {
      handle = reverse ? cross_correlate_initialize(x_length, h_length)
                       : convolve_initialize(x_length, h_length);
      have = 1;
    }
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
    for (long b = 0; b < (long)batch; b++) {
      const float *xb = x + b * x_stride;
      float *rb = result + b * result_stride;
      if (algorithm < 0) {
        if (reverse) cross_correlate(handle, xb, h, rb);
        else convolve(handle, xb, h, rb);
      } else {
        if (reverse) cross_correlate_simd(1, xb, x_length, h, h_length, rb);
        else convolve_simd(1, xb, x_length, h, h_length, rb);
      }
    }
    if (have) {
      if (reverse) cross_correlate_finalize(handle); else convolve_finalize(handle);
    }
  }
  return used;
}

REF_API int ref_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* struct layout probe for the ABI test (tests/test_abi.py) */
REF_API void ref_struct_layout(size_t out[16]) {
  out[0] = sizeof(ConvolutionOverlapSaveHandle);
  out[1] = sizeof(ConvolutionFFTHandle);
  out[2] = sizeof(ConvolutionHandle);
  out[3] = offsetof(ConvolutionOverlapSaveHandle, fft_boiler_plate);
  out[4] = offsetof(ConvolutionOverlapSaveHandle, H);
  out[5] = offsetof(ConvolutionOverlapSaveHandle, x_length);
  out[6] = offsetof(ConvolutionOverlapSaveHandle, h_length);
  out[7] = offsetof(ConvolutionOverlapSaveHandle, L);
  out[8] = offsetof(ConvolutionOverlapSaveHandle, reverse);
  out[9] = offsetof(ConvolutionFFTHandle, M);
  out[10] = offsetof(ConvolutionFFTHandle, x_length);
  out[11] = offsetof(ConvolutionFFTHandle, h_length);
  out[12] = offsetof(ConvolutionFFTHandle, inputs);
  out[13] = offsetof(ConvolutionFFTHandle, reverse);
  out[14] = offsetof(ConvolutionHandle, x_length);
  out[15] = offsetof(ConvolutionHandle, handle);
}

/* the reference's converters are header inlines (inc/simd/arithmetic.h): export callable
 * copies of its scalar versions for the oracle tests */
#include <simd/arithmetic.h>
REF_API void ref_int16_to_float_na(const int16_t *data, size_t length, float *res) {
  int16_to_float_na(data, length, res);
}
REF_API void ref_float_to_int16_na(const float *data, size_t length, int16_t *res) {
  float_to_int16_na(data, length, res);
}
