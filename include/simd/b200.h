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
#include <simd/detect_peaks.h>
#include <stdint.h>

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

/* ---- pinned host buffers ---------------------------------------------------------------
 * mallocf() (simd/memory.h, reference src/memory.c:89-91) returns pageable memory, which the
 * GPU's copy engines cannot read in place: the library then moves it through pinned bounce
 * buffers with a few copy threads (about half the rate of pinned memory end to end).  A
 * caller that can choose its allocator gets the full host-link rate with these two; the
 * memory is page-locked and usable from every GPU.  Release with simd_b200_free_pinned(). */
float *simd_b200_mallocf_pinned(size_t length);
void simd_b200_free_pinned(float *ptr);

/* ---- fewer bytes over PCIe for host-pointer callers -----------------------------------
 * A drop-in caller hands over host memory, and end to end such a call is bound by the host
 * link, not by the kernels (BENCH: ~45 GB/s each way against ~2.3 TB/s of kernel traffic).
 * Two chains of the reference API that a caller would otherwise run as separate calls are
 * offered as one, so that the bulky intermediate never crosses the link:
 *
 * int16 samples (PCM, ADC data): the reference widens them on the host with int16_to_float
 * (inc/simd/arithmetic.h:231-278) and convolves the floats.  Here the int16 samples are
 * what is copied to the GPU (2 bytes per sample instead of 4) and they are widened there,
 * inside the staging pipeline.  x: HOST memory, x_stride in samples; everything else as
 * convolve_batch / cross_correlate_batch; the result is what the float call gives on
 * (float)x[i]. */
void convolve_batch_int16(struct ConvolutionHandle handle, const int16_t *x, size_t x_stride,
                          const float *h, float *result, size_t result_stride, size_t batch)
    NOTNULL(2, 4, 5);
void cross_correlate_batch_int16(struct ConvolutionHandle handle, const int16_t *x,
                                 size_t x_stride, const float *h, float *result,
                                 size_t result_stride, size_t batch) NOTNULL(2, 4, 5);

/* cross_correlate() followed by detect_peaks() (reference src/correlate.c:144-156, then
 * src/detect_peaks.c:58-127) with the x_length+h_length-1 correlation samples kept in GPU
 * memory: only the extrema travel back.  *results as detect_peaks() returns it (malloc()ed,
 * NULL when there is none). */
void simd_b200_cross_correlate_peaks(struct ConvolutionHandle handle, const float *x,
                                     const float *h, ExtremumType type,
                                     ExtremumPoint **results, size_t *resultsLength)
    NOTNULL(2, 3, 5, 6);

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

/* ---- several GPUs from one process ----------------------------------------------------
 * The reference is single-threaded and knows one address space.  A host-pointer caller of
 * the batched calls, or of simd_b200_filter_long, can let the library spread the work over
 * the GPUs of the box: one host thread, one set of staging buffers and one PCIe link per
 * device; batches split by rows, a long signal by contiguous output ranges whose (h-1)
 * samples of halo are read from the caller's buffer.  No collective, no peer traffic.
 *   simd_b200_use_devices(n)   n GPUs from now on, starting at the calling thread's current
 *                              device (0 = all visible).  Default 1: nothing changes for a
 *                              program that manages devices itself.
 * Device-pointer calls always run on the current device. */
void simd_b200_use_devices(int n);
int simd_b200_devices_in_use(void);

/* Full convolution (reverse = 0) or cross-correlation (reverse != 0) of one long signal:
 * result[0 .. x_length+h_length-2], as convolve() / cross_correlate() on a handle made for
 * these lengths (reference src/convolve.c:381-395), without a handle and over
 * simd_b200_devices_in_use() GPUs when the pointers are host memory. */
void simd_b200_filter_long(const float *x, size_t x_length, const float *h, size_t h_length,
                           int reverse, float *result) NOTNULL(1, 3, 6);

/* The same operation on a signal that already lives sharded in the memories of several GPUs
 * (one process, all devices visible).  Shard i = samples [s_i, e_i) of the signal, at
 * x_shards[i] on device devices[i], with h_length-1 writable floats in FRONT of it: the call
 * copies the tail of shard i-1 there (cudaMemcpyPeerAsync over NVLink; the halo lands in
 * local memory before the kernels of that GPU start -- no kernel reads a peer's memory), then
 * every GPU filters its shard.  result_shards[i] receives shard_lengths[i] samples of the
 * result, the last one h_length-1 more; concatenated they are the full result.
 * h_on_device[i] is a copy of h in device i's memory.  Every shard but the last must hold at
 * least h_length-1 samples.  Returns after all GPUs have finished. */
void simd_b200_filter_long_sharded(int n_shards, const int *devices, float *const *x_shards,
                                   const size_t *shard_lengths, const float *const *h_on_device,
                                   size_t h_length, int reverse, float *const *result_shards)
    NOTNULL(2, 3, 4, 5, 8);

/* ---- selector, exposed for inspection and tests (pure host logic, no CUDA call) -------
 * The decision convolve_initialize makes on B200 (replaces the 2h / 200 / 350 rules of
 * reference src/convolve.c:349-363) and the transform length the FFT-based paths use. */
ConvolutionAlgorithm simd_b200_select_algorithm(size_t x_length, size_t h_length);
/* Same decision for `batch` signals sharing one h (what convolve_batch re-evaluates when its
 * handle says brute force: many short-filter signals are faster through the transform path). */
ConvolutionAlgorithm simd_b200_select_algorithm_batch(size_t x_length, size_t h_length,
                                                      size_t batch);
int simd_b200_transform_length(size_t x_length, size_t h_length);

/* ---- stream and device control --------------------------------------------------------
 * The library enqueues on one CUDA stream per device: its own non-blocking stream until
 * simd_b200_set_stream() installs the caller's for the current device (a cudaStream_t passed
 * as void*; NULL is the legacy default stream).  Calls whose x, h and result are all device pointers return
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

/* ---- allocation-free filtering ---------------------------------------------------------
 * No counterpart in the reference (which allocates with malloc in every *_initialize).
 *
 * For callers that must not, or cannot, allocate device memory while they run: code that
 * lives inside a fixed memory budget handed out by its own arena, plug-ins whose host
 * forbids cudaMalloc after start-up, and real-time loops where an allocation (an implicit
 * device synchronisation) would break their deadline.  The handle API cannot serve them:
 * convolve_overlap_save_initialize() allocates the spectrum scratch and staging buffers.
 *
 * result[0 .. x_length+h_length-2] = x (*) h   (reverse = 0, as convolve())
 *                                  = x (x) h   (reverse != 0, as cross_correlate())
 * computed by the same fused overlap-save kernels as convolve_overlap_save().  The only
 * device memory used besides the caller's x, h and result is a pair of arrays built into
 * the library image (128 KiB of spectrum scratch + 127 KiB of twiddle tables per GPU that
 * loads the library, held until the process exits).  The call performs no cudaMalloc,
 * cudaFree or other allocating call, creates no stream or event, and returns after
 * enqueueing its kernels on `cuda_stream` (a cudaStream_t passed as void*, NULL = the
 * legacy default stream); it does not wait for them.
 *
 * Limits and rules
 *   - h_length <= 8192 taps (what the 16384-bin scratch array holds); larger values are
 *     rejected with return code 1 and a message on stderr.  x_length is not limited by it.
 *   - x, h and result must be device (or managed) memory; host pointers are rejected with
 *     return code 2, because staging them would take an allocation.  result must not
 *     overlap x or h.
 *   - The first call on each device does a one-time set-up with that device current: it
 *     looks up the device addresses of the built-in arrays, fills the twiddle tables with a
 *     small kernel on `cuda_stream` and waits for it.  Make that first call before any
 *     stream capture begins.
 *   - All calls on one device share the one scratch array, and it keeps whatever the last
 *     call left in it.  Nothing needs clearing between calls (each call rewrites every bin
 *     it later reads), but two calls whose kernels can run at the same time on the GPU
 *     (different streams, or different host threads without ordering) race on it: order
 *     them on one stream or with events.
 * Returns 0 on success, 1 size beyond the built-in maximum, 2 bad argument, 3 CUDA error
 * during the one-time set-up (which is then retried by the next call).  A kernel launch
 * that fails aborts the process with a message, as everywhere else in this library (the
 * reference's assert convention).  Per-kernel event timing (simd_b200_profile) skips this
 * entry point, so that it never creates an event. */
int simd_b200_filter_noalloc(const float *x, size_t x_length, const float *h,
                             size_t h_length, int reverse, float *result, void *cuda_stream);

SIMD_API_END

#endif  /* SIMD_B200_EXT_H_ */
