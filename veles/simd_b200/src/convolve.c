/* convolve.c -- plain-C host layer of the convolution API (B200 build).
 *
 * Mirrors the reference's public functions one for one (/root/reference/src/convolve.c);
 * each builds a B200Job and hands it to the CUDA side through the C-ABI shim declared in
 * csrc/b200_internal.h.  Handle bookkeeping (heap int for L / M, by-value structs, who
 * frees what) follows the reference so that callers cannot tell the difference.
 */
#define LIBSIMD_IMPLEMENTATION
#include <simd/convolve.h>
#include <simd/b200.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "../csrc/b200_internal.h"

/* The reference checks preconditions with assert(), which is live in its release builds
 * (common.ac:51-54 sets no -DNDEBUG): a violation aborts the process.  Same here, with a
 * message that names the condition. */
#define SIMD_REQUIRE(cond)                                                              \
  do {                                                                                  \
    if (!(cond)) {                                                                      \
      fprintf(stderr, "libSimd(b200): precondition failed: %s (%s:%d)\n", #cond,        \
              __FILE__, __LINE__);                                                      \
      abort();                                                                          \
    }                                                                                   \
  } while (0)

static void run_full_typed(B200Plan *plan, int algorithm, const float *x, size_t x_length,
                           size_t x_stride, const float *h, size_t h_length, float *result,
                           size_t result_stride, size_t batch, int reverse, int x_int16) {
  B200Job job;
  memset(&job, 0, sizeof(job));
  job.x_int16 = x_int16;
  job.x = x; job.x_len = (long long)x_length; job.x_stride = (long long)x_stride;
  job.h = h; job.h_len = (long long)h_length;
  job.out = result;
  job.out_len = (long long)(x_length + h_length - 1);
  job.out_stride = (long long)result_stride;
  job.pad = (long long)h_length - 1;
  job.batch = (long long)batch;
  job.reverse = reverse;
  if (batch > 1) b200_run_multi(plan, algorithm, &job);   /* one GPU unless simd_b200_use_devices() said more */
  else b200_run(plan, algorithm, &job);
}

static void run_full(B200Plan *plan, int algorithm, const float *x, size_t x_length,
                     size_t x_stride, const float *h, size_t h_length, float *result,
                     size_t result_stride, size_t batch, int reverse) {
  run_full_typed(plan, algorithm, x, x_length, x_stride, h, h_length, result, result_stride,
                 batch, reverse, 0);
}

/* reference: src/convolve.c:40-101 */
void convolve_simd(int simd, const float *__restrict x, size_t xLength,
                   const float *__restrict h, size_t hLength, float *__restrict result) {
  (void)simd;
  SIMD_REQUIRE(x); SIMD_REQUIRE(h); SIMD_REQUIRE(result);
  SIMD_REQUIRE(xLength > 0); SIMD_REQUIRE(hLength > 0);
  run_full(NULL, B200_ALG_DIRECT, x, xLength, xLength, h, hLength, result,
           xLength + hLength - 1, 1, 0);
}

/* reference: src/convolve.c:103-146 */
ConvolutionOverlapSaveHandle convolve_overlap_save_initialize(size_t xLength,
                                                              size_t hLength) {
  SIMD_REQUIRE(hLength < xLength / 2);
  SIMD_REQUIRE(xLength > 0);
  SIMD_REQUIRE(hLength > 0);
  ConvolutionOverlapSaveHandle handle;
  memset(&handle, 0, sizeof(handle));
  int n = 0, real_packed = 0;
  b200_os_pick_mode((long long)hLength, (long long)(xLength + hLength - 1), &n, &real_packed);
  B200Plan *plan = b200_plan_create(B200_ALG_OS, (long long)xLength, (long long)hLength, n,
                                    real_packed);
  handle.fft_plan = plan;
  handle.fft_inverse_plan = plan;
  handle.fft_boiler_plate = b200_plan_scratch(plan);
  handle.H = b200_plan_spectrum(plan);
  handle.x_length = xLength;
  handle.h_length = hLength;
  handle.L = malloc(sizeof(int));
  SIMD_REQUIRE(handle.L);
  *handle.L = n;
  handle.reverse = 0;
  return handle;
}

/* reference: src/convolve.c:148-154 */
void convolve_overlap_save_finalize(ConvolutionOverlapSaveHandle handle) {
  b200_plan_destroy((B200Plan *)handle.fft_plan);
  free(handle.L);
}

/* reference: src/convolve.c:156-229 */
void convolve_overlap_save(ConvolutionOverlapSaveHandle handle, const float *x,
                           const float *h, float *result) {
  SIMD_REQUIRE(x != NULL); SIMD_REQUIRE(h != NULL); SIMD_REQUIRE(result != NULL);
  run_full((B200Plan *)handle.fft_plan, B200_ALG_OS, x, handle.x_length, handle.x_length,
           h, handle.h_length, result, handle.x_length + handle.h_length - 1, 1,
           handle.reverse);
}

/* reference: src/convolve.c:231-278 */
ConvolutionFFTHandle convolve_fft_initialize(size_t xLength, size_t hLength) {
  SIMD_REQUIRE(hLength > 0);
  SIMD_REQUIRE(xLength > 0);
  SIMD_REQUIRE(xLength + hLength - 1 < 2147483647u);   /* int fields, SURVEY Q7 */
  ConvolutionFFTHandle handle;
  memset(&handle, 0, sizeof(handle));
  /* M as the reference reports it: next power of two >= x+h-1 (src/convolve.c:237-244) */
  long long out_len = (long long)xLength + (long long)hLength - 1;
  long long M = 1;
  while (M < out_len) M <<= 1;
  handle.M = malloc(sizeof(int));
  SIMD_REQUIRE(handle.M);
  *handle.M = (int)(M > 2147483647LL ? 2147483647LL : M);
  handle.x_length = (int)xLength;
  handle.h_length = (int)hLength;
  handle.reverse = 0;
  handle.inputs = NULL;
  int n, real_packed = 0;
  if (out_len <= B200_FFT_MAX) {
    n = (int)(M < B200_FFT_MIN ? B200_FFT_MIN : M);   /* whole signal in one transform */
  } else {
    b200_os_pick_mode((long long)hLength, out_len, &n, &real_packed);  /* segmented / partitioned */
  }
  B200Plan *plan = b200_plan_create(B200_ALG_FFT, (long long)xLength, (long long)hLength, n,
                                    real_packed);
  handle.fft_plan = plan;
  handle.fft_inverse_plan = plan;
  return handle;
}

/* reference: src/convolve.c:280-287 */
void convolve_fft_finalize(ConvolutionFFTHandle handle) {
  b200_plan_destroy((B200Plan *)handle.fft_plan);
  free(handle.M);
}

/* reference: src/convolve.c:289-326 */
void convolve_fft(ConvolutionFFTHandle handle, const float *x, const float *h,
                  float *result) {
  SIMD_REQUIRE(x != NULL); SIMD_REQUIRE(h != NULL); SIMD_REQUIRE(result != NULL);
  size_t xl = (size_t)handle.x_length, hl = (size_t)handle.h_length;
  run_full((B200Plan *)handle.fft_plan, B200_ALG_FFT, x, xl, xl, h, hl, result,
           xl + hl - 1, 1, handle.reverse);
}

/* reference: src/convolve.c:328-366 */
ConvolutionHandle convolve_initialize(size_t xLength, size_t hLength) {
  SIMD_REQUIRE(xLength > 0);
  SIMD_REQUIRE(hLength > 0);
  SIMD_REQUIRE(xLength + hLength - 1 < 2147483647u);
  ConvolutionHandle handle;
  memset(&handle, 0, sizeof(handle));
  handle.x_length = (int)xLength;
  handle.h_length = (int)hLength;
  handle.algorithm = simd_b200_select_algorithm(xLength, hLength);
  switch (handle.algorithm) {
    case kConvolutionAlgorithmFFT:
      handle.handle.fft = convolve_fft_initialize(xLength, hLength);
      break;
    case kConvolutionAlgorithmOverlapSave:
      handle.handle.os = convolve_overlap_save_initialize(xLength, hLength);
      break;
    case kConvolutionAlgorithmBruteForce:
      break;
  }
  return handle;
}

/* reference: src/convolve.c:368-379 */
void convolve_finalize(ConvolutionHandle handle) {
  switch (handle.algorithm) {
    case kConvolutionAlgorithmFFT:
      convolve_fft_finalize(handle.handle.fft);
      break;
    case kConvolutionAlgorithmOverlapSave:
      convolve_overlap_save_finalize(handle.handle.os);
      break;
    case kConvolutionAlgorithmBruteForce:
      break;
  }
}

/* reference: src/convolve.c:381-395 */
void convolve(ConvolutionHandle handle, const float *__restrict x,
              const float *__restrict h, float *__restrict result) {
  switch (handle.algorithm) {
    case kConvolutionAlgorithmFFT:
      convolve_fft(handle.handle.fft, x, h, result);
      break;
    case kConvolutionAlgorithmOverlapSave:
      convolve_overlap_save(handle.handle.os, x, h, result);
      break;
    case kConvolutionAlgorithmBruteForce:
      convolve_simd(1, x, (size_t)handle.x_length, h, (size_t)handle.h_length, result);
      break;
  }
}

/* ---- additive surface (simd/b200.h) ---------------------------------------------------- */

static void batch_dispatch(ConvolutionHandle handle, const float *x, size_t x_stride,
                           const float *h, float *result, size_t result_stride,
                           size_t batch, int brute_reverse, int x_int16) {
  SIMD_REQUIRE(x != NULL); SIMD_REQUIRE(h != NULL); SIMD_REQUIRE(result != NULL);
  size_t xl = (size_t)handle.x_length, hl = (size_t)handle.h_length;
  SIMD_REQUIRE(x_stride >= xl);
  SIMD_REQUIRE(result_stride >= xl + hl - 1);
  switch (handle.algorithm) {
    case kConvolutionAlgorithmFFT:
      run_full_typed((B200Plan *)handle.handle.fft.fft_plan, B200_ALG_FFT, x, xl, x_stride, h, hl,
                     result, result_stride, batch, handle.handle.fft.reverse, x_int16);
      break;
    case kConvolutionAlgorithmOverlapSave:
      run_full_typed((B200Plan *)handle.handle.os.fft_plan, B200_ALG_OS, x, xl, x_stride, h, hl,
                     result, result_stride, batch, handle.handle.os.reverse, x_int16);
      break;
    case kConvolutionAlgorithmBruteForce:
      /* The handle was chosen for one signal.  A large batch can cross over to the fused
       * transform kernels (profiles/selector_sweep_r01.md); those then run without a handle,
       * on the per-device context's scratch. */
      /* profiling knob: VELES_B200_NO_BATCH_CROSSOVER=1 keeps the batch on the direct kernel */
      if (getenv("VELES_B200_NO_BATCH_CROSSOVER") ||
          simd_b200_select_algorithm_batch(xl, hl, batch) == kConvolutionAlgorithmBruteForce) {
        run_full_typed(NULL, B200_ALG_DIRECT, x, xl, x_stride, h, hl, result, result_stride, batch,
                       brute_reverse, x_int16);
      } else {
        run_full_typed(NULL, B200_ALG_OS, x, xl, x_stride, h, hl, result, result_stride, batch,
                       brute_reverse, x_int16);
      }
      break;
  }
}

void convolve_batch(ConvolutionHandle handle, const float *x, size_t x_stride,
                    const float *h, float *result, size_t result_stride, size_t batch) {
  batch_dispatch(handle, x, x_stride, h, result, result_stride, batch, 0, 0);
}

void convolve_batch_int16(ConvolutionHandle handle, const int16_t *x, size_t x_stride,
                          const float *h, float *result, size_t result_stride, size_t batch) {
  batch_dispatch(handle, (const float *)(const void *)x, x_stride, h, result, result_stride, batch, 0, 1);
}

void cross_correlate_batch_int16(ConvolutionHandle handle, const int16_t *x, size_t x_stride,
                                 const float *h, float *result, size_t result_stride,
                                 size_t batch) {
  batch_dispatch(handle, (const float *)(const void *)x, x_stride, h, result, result_stride, batch, 1, 1);
}

void cross_correlate_batch(ConvolutionHandle handle, const float *x, size_t x_stride,
                           const float *h, float *result, size_t result_stride,
                           size_t batch) {
  batch_dispatch(handle, x, x_stride, h, result, result_stride, batch, 1, 0);
}

void simd_b200_filter_shard(const float *x, size_t x_length, const float *h,
                            size_t h_length, int reverse, size_t left_pad, float *result,
                            size_t result_length, int algorithm) {
  SIMD_REQUIRE(x != NULL); SIMD_REQUIRE(h != NULL); SIMD_REQUIRE(result != NULL);
  SIMD_REQUIRE(x_length > 0); SIMD_REQUIRE(h_length > 0);
  if (algorithm < 0) {
    algorithm = simd_b200_select_algorithm(result_length, h_length) ==
                        kConvolutionAlgorithmBruteForce
                    ? B200_ALG_DIRECT
                    : B200_ALG_OS;
  }
  B200Job job;
  memset(&job, 0, sizeof(job));
  job.x = x; job.x_len = (long long)x_length; job.x_stride = (long long)x_length;
  job.h = h; job.h_len = (long long)h_length;
  job.out = result; job.out_len = (long long)result_length;
  job.out_stride = (long long)result_length;
  job.pad = (long long)left_pad;
  job.batch = 1;
  job.reverse = reverse;
  /* no handle: the FFT-based path uses the per-device context's scratch */
  b200_run(NULL, algorithm == B200_ALG_DIRECT ? B200_ALG_DIRECT : B200_ALG_OS, &job);
}

static int shard_algorithm(size_t result_length, size_t h_length) {
  return simd_b200_select_algorithm(result_length, h_length) == kConvolutionAlgorithmBruteForce
             ? B200_ALG_DIRECT
             : B200_ALG_OS;
}

void simd_b200_filter_long(const float *x, size_t x_length, const float *h, size_t h_length,
                           int reverse, float *result) {
  SIMD_REQUIRE(x != NULL); SIMD_REQUIRE(h != NULL); SIMD_REQUIRE(result != NULL);
  SIMD_REQUIRE(x_length > 0); SIMD_REQUIRE(h_length > 0);
  B200Job job;
  memset(&job, 0, sizeof(job));
  job.x = x; job.x_len = (long long)x_length; job.x_stride = (long long)x_length;
  job.h = h; job.h_len = (long long)h_length;
  job.out = result; job.out_len = (long long)(x_length + h_length - 1);
  job.out_stride = job.out_len;
  job.pad = (long long)h_length - 1;
  job.batch = 1;
  job.reverse = reverse;
  b200_run_multi(NULL, shard_algorithm(x_length + h_length - 1, h_length), &job);
}

void simd_b200_filter_long_sharded(int n_shards, const int *devices, float *const *x_shards,
                                   const size_t *shard_lengths, const float *const *h_on_device,
                                   size_t h_length, int reverse, float *const *result_shards) {
  SIMD_REQUIRE(n_shards >= 1); SIMD_REQUIRE(h_length > 0);
  size_t longest = 0;
  for (int i = 0; i < n_shards; i++) if (shard_lengths[i] > longest) longest = shard_lengths[i];
  b200_filter_long_sharded(n_shards, devices, x_shards, shard_lengths, h_on_device, h_length,
                           reverse, result_shards, shard_algorithm(longest, h_length));
}

float *simd_b200_mallocf_pinned(size_t length) { return b200_host_alloc_pinned(length); }
void simd_b200_free_pinned(float *ptr) { b200_host_free_pinned(ptr); }
void simd_b200_use_devices(int n) { b200_use_devices(n); }
int simd_b200_devices_in_use(void) { return b200_devices_in_use(); }
void simd_b200_set_stream(void *cuda_stream) { b200_set_stream(cuda_stream); }
void *simd_b200_get_stream(void) { return b200_get_stream(); }
void simd_b200_synchronize(void) { b200_synchronize(); }
int simd_b200_device_count(void) { return b200_device_count(); }
long long simd_b200_kernel_launches(void) { return b200_kernel_launches(); }
void simd_b200_profile(int enable) { b200_profile_enable(enable); }
void simd_b200_profile_collect(double *total_ms, long long *count) {
  b200_profile_collect(total_ms, count);
}
