/* simd/b200.h -- ADDITIVE surface of the B200 build.  Nothing here exists in the
 * reference; the 20 reference symbols (convolve.h, correlate.h) stay source- and
 * ABI-compatible without it.  It exists because (hard parts 1-3):
 *   - the reference has no batched entry point (a batch is the caller's loop over
 *     convolve(), one handle per signal) and a GPU wants the whole batch in one launch;
 *   - throughput is only meaningful with data resident in HBM, so callers need to say
 *     which stream the work goes to and when to wait for it;
 *   - a long signal sharded over several GPUs needs "my shard plus a halo" as a primitive.
 */
#ifndef SIMD_B200_EXT_H_
#define SIMD_B200_EXT_H_

#include <stddef.h>
#include <simd/common.h>
#include <simd/attributes.h>
#include <simd/convolve_structs.h>

SIMD_API_BEGIN

/* ---- batched calls --------------------------------------------------------------------
 * `batch` independent signals of handle.x_length samples, signal b at x + b*x_stride, all
 * filtered with the same h; result b (x_length + h_length - 1 floats) goes to
 * result + b*result_stride.  Equivalent to looping convolve() / cross_correlate()
 * (reference src/convolve.c:381-395, src/correlate.c:144-156) over the signals with one
 * handle.  Host or device pointers, as for the single-signal calls. */
void convolve_batch(struct ConvolutionHandle handle, const float *x, size_t x_stride,
                    const float *h, float *result, size_t result_stride, size_t batch)
    NOTNULL(2, 4, 5);
void cross_correlate_batch(struct ConvolutionHandle handle, const float *x, size_t x_stride,
                           const float *h, float *result, size_t result_stride,
                           size_t batch) NOTNULL(2, 4, 5);

/* ---- shard primitive ------------------------------------------------------------------
 * Filters a piece of a longer signal:
 *   result[n] = sum_j g[j] * X(n + j - left_pad),  n in [0, result_length)
 *   g = h reversed (convolution, reverse = 0) or h (cross-correlation, reverse = 1)
 *   X(i) = x[i] for 0 <= i < x_length, else 0
 * left_pad = h_length-1 with result_length = x_length+h_length-1 is the full operation; a
 * GPU that owns samples [s, e) of a long signal and has the h_length-1 samples before s
 * in front of them (the halo) passes left_pad = 0 and result_length = e - s.
 * algorithm: a ConvolutionAlgorithm value, or -1 to let the library choose. */
void simd_b200_filter_shard(const float *x, size_t x_length, const float *h,
                            size_t h_length, int reverse, size_t left_pad, float *result,
                            size_t result_length, int algorithm) NOTNULL(1, 3, 7);

/* ---- selector, exposed for inspection and tests (pure host logic, no CUDA call) -------
 * The decision convolve_initialize makes on B200 (replaces the 2h / 200 / 350 rules of
 * reference src/convolve.c:349-363) and the transform length the FFT-based paths use. */
ConvolutionAlgorithm simd_b200_select_algorithm(size_t x_length, size_t h_length);
int simd_b200_transform_length(size_t x_length, size_t h_length);

/* ---- stream and device control --------------------------------------------------------
 * The library enqueues on one CUDA stream per process: its own non-blocking stream until
 * simd_b200_set_stream() installs the caller's (a cudaStream_t passed as void*; NULL is the
 * legacy default stream).  Calls whose x, h and result are all device pointers return
 * without waiting; simd_b200_synchronize() waits for the library stream.  The device is
 * whatever cudaSetDevice() made current in the calling thread; a handle must be used on
 * the device it was created on. */
void simd_b200_set_stream(void *cuda_stream);
void *simd_b200_get_stream(void);
void simd_b200_synchronize(void);
int simd_b200_device_count(void);

/* Number of kernels this library has launched in this process (for benchmarks that must
 * show the GPU path ran). */
long long simd_b200_kernel_launches(void);

/* Per-kernel device timing for benchmarks: while enabled, every kernel launch is bracketed
 * by CUDA events on the launching stream.  simd_b200_profile_collect() waits for them and
 * fills total_ms[k] / count[k] for k = 0 direct kernel, 1 fused overlap-save kernel,
 * 2 filter-spectrum kernel (arrays of 3), then clears the record. */
void simd_b200_profile(int enable);
void simd_b200_profile_collect(double *total_ms, long long *count) NOTNULL(1, 2);

SIMD_API_END

#endif  /* SIMD_B200_EXT_H_ */
