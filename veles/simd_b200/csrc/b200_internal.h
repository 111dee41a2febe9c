/* b200_internal.h -- private interface between the plain-C host layer (src/*.c) and the
 * CUDA translation units (csrc/*.cu).  C linkage, no CUDA types in the signatures, so the
 * host layer compiles with gcc alone.  Not installed; the public ABI is include/simd/*.h.
 */
#ifndef VELES_SIMD_B200_INTERNAL_H_
#define VELES_SIMD_B200_INTERNAL_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* One unit of device work.  All of it is "valid-region" correlation-form filtering:
 *   out[b][n] = sum_{j < h_len} g[j] * X_b(n + j - pad),   n in [0, out_len)
 *   g[j]   = reverse ? h[j] : h[h_len-1-j]
 *   X_b(i) = x[b*x_stride + i] for 0 <= i < x_len, else 0
 * Full linear convolution (src/convolve.c:49-99) is pad = h_len-1, out_len = x_len+h_len-1,
 * reverse = 0; cross-correlation (src/correlate.c:78-124) is the same with reverse = 1.
 * A halo-extended shard of a long signal is pad = 0.  Pointers may be host or device. */
typedef struct B200Job {
  const float *x;
  long long x_len;
  long long x_stride;
  const float *h;
  long long h_len;
  float *out;
  long long out_len;
  long long out_stride;
  long long pad;
  long long batch;
  int reverse;
  int x_int16;   /* x points at int16_t samples in HOST memory (strides and lengths still count
                  * samples): they cross PCIe as 2 bytes each and are widened on the device */
} B200Job;

/* Algorithms; numeric values follow ConvolutionAlgorithm (convolve_structs.h). */
enum { B200_ALG_DIRECT = 0, B200_ALG_FFT = 1, B200_ALG_OS = 2 };

typedef struct B200Plan B200Plan;   /* opaque; lives behind handle.fft_plan */

/* Plan for the fused FFT kernels.  n_fft = complex transform length (power of two in
 * [B200_FFT_MIN, B200_FFT_MAX]); x_len/h_len/batch_hint size the staging buffers used when
 * callers pass host pointers.  Aborts on CUDA errors (reference convention: assert). */
B200Plan *b200_plan_create(int algorithm, long long x_len, long long h_len, int n_fft,
                           int real_packed);
void b200_plan_destroy(B200Plan *plan);
float *b200_host_alloc_pinned(size_t floats);      /* cudaHostAlloc(portable) / cudaFreeHost */
void b200_host_free_pinned(float *p);
float *b200_device_alloc(size_t floats);           /* cudaMalloc / cudaFree on the current device */
void b200_device_free(float *p);
float *b200_plan_spectrum(const B200Plan *plan);   /* device pointer */
float *b200_plan_scratch(const B200Plan *plan);    /* device pointer */

/* Run a job.  plan may be NULL for B200_ALG_DIRECT.  Host pointers are staged through
 * device buffers and the call returns after the result is back in host memory; device
 * pointers are used in place and the call is asynchronous on the library stream. */
void b200_run(B200Plan *plan, int algorithm, const B200Job *job);

/* Several GPUs from one process (csrc/context.cu).  b200_use_devices(n): host-pointer jobs
 * given to b200_run_multi are spread over n GPUs (0 = all visible; default 1). */
void b200_use_devices(int n);
int b200_devices_in_use(void);
void b200_run_multi(B200Plan *plan, int algorithm, const B200Job *job);
void b200_filter_long_sharded(int n_shards, const int *devices, float *const *x_shards,
                              const size_t *shard_lengths, const float *const *h_on_device,
                              size_t h_len, int reverse, float *const *result_shards,
                              int algorithm);

/* transform-length policy (src/selector.c, pure host logic) */
int b200_os_pick_n(long long h_len, long long out_len);
int b200_os_part_taps(int n_fft, long long h_len);
/* full choice: transform length and whether one real segment per transform (real-packed)
 * beats two paired segments; real-packed needs h_len <= n */
void b200_os_pick_mode(long long h_len, long long out_len, int *n_fft, int *real_packed);

#define B200_FFT_MIN 256
#define B200_FFT_MAX 16384

/* stream / device helpers exported through include/simd/b200.h */
void b200_set_stream(void *cuda_stream);
void *b200_get_stream(void);
void b200_synchronize(void);
int b200_device_count(void);
long long b200_kernel_launches(void);
void b200_profile_enable(int on);
void b200_profile_collect(double *total_ms, long long *count);

#ifdef __cplusplus
}
#endif
#endif  /* VELES_SIMD_B200_INTERNAL_H_ */
