/* simd/convolve.h -- 1D linear convolution, B200 (sm_100a) implementation behind the
 * veles.simd C API.  Drop-in for /root/reference/inc/simd/convolve.h: same ten symbols,
 * same signatures, same by-value handles.
 *
 * Behaviour common to every entry point
 *   - result always receives x_length + h_length - 1 floats (the reference's header says
 *     "xLength" for three of them, its code writes x+h-1: src/convolve.c:49,221-226,325).
 *   - x, h, result may each be a host pointer or a device pointer (classified with
 *     cudaPointerGetAttributes).  With host pointers the call returns when result is
 *     complete in host memory.  With device pointers everywhere the call only enqueues work
 *     on the library stream (simd/b200.h) and returns.
 *   - preconditions are enforced like the reference does: a violated one prints a message
 *     and abort()s (the reference uses assert(), live in its release builds).  CUDA errors
 *     do the same.  There is no CPU fallback.
 *   - one thread per handle at a time; distinct handles may be used concurrently.
 */
#ifndef SIMD_B200_CONVOLVE_H_
#define SIMD_B200_CONVOLVE_H_

#include <stddef.h>
#include <simd/common.h>
#include <simd/attributes.h>
#include <simd/convolve_structs.h>

SIMD_API_BEGIN

typedef struct ConvolutionFFTHandle ConvolutionFFTHandle;
typedef struct ConvolutionOverlapSaveHandle ConvolutionOverlapSaveHandle;
typedef struct ConvolutionHandle ConvolutionHandle;

/* Whole-signal FFT method.  Replaces reference inc/simd/convolve.h:48 /
 * src/convolve.c:231-278.  *handle.M = next power of two >= x+h-1, as in the reference.
 * Unlike the reference (whose handle is single-use) the handle may be reused. */
ConvolutionFFTHandle convolve_fft_initialize(size_t xLength, size_t hLength);

/* Replaces reference inc/simd/convolve.h:57-59 / src/convolve.c:289-326. */
void convolve_fft(ConvolutionFFTHandle handle, const float *x, const float *h,
                  float *result) NOTNULL(2, 3, 4);

/* Replaces reference inc/simd/convolve.h:63 / src/convolve.c:280-287. */
void convolve_fft_finalize(ConvolutionFFTHandle handle);

/* Overlap-save method.  Replaces reference inc/simd/convolve.h:72-73 /
 * src/convolve.c:103-146.  Aborts unless hLength < xLength / 2 (src/convolve.c:105). */
ConvolutionOverlapSaveHandle convolve_overlap_save_initialize(size_t xLength,
                                                              size_t hLength);

/* Replaces reference inc/simd/convolve.h:82-85 / src/convolve.c:156-229.  Samples past
 * xLength are zeros (the reference over-reads them in its first block). */
void convolve_overlap_save(ConvolutionOverlapSaveHandle handle, const float *x,
                           const float *h, float *result) NOTNULL(2, 3, 4);

/* Replaces reference inc/simd/convolve.h:89 / src/convolve.c:148-154. */
void convolve_overlap_save_finalize(ConvolutionOverlapSaveHandle handle);

/* Direct ("brute force") method.  Replaces reference inc/simd/convolve.h:99-102 /
 * src/convolve.c:40-101.  simd != 0 and simd == 0 both run the direct GPU kernel (the
 * reference's simd=0 scalar loop is its own test oracle; it lives in oracle/ here). */
void convolve_simd(int simd, const float *__restrict x, size_t xLength,
                   const float *__restrict h, size_t hLength,
                   float *__restrict result) NOTNULL(2, 4, 6);

/* Best-method handle.  Replaces reference inc/simd/convolve.h:111 /
 * src/convolve.c:328-366; the thresholds come from B200 timings (see simd/b200.h,
 * simd_b200_select_algorithm) instead of the reference's 2h / 200 / 350 rules. */
ConvolutionHandle convolve_initialize(size_t xLength, size_t hLength);

/* Replaces reference inc/simd/convolve.h:119-121 / src/convolve.c:381-395. */
void convolve(ConvolutionHandle handle, const float *__restrict x,
              const float *__restrict h, float *__restrict result) NOTNULL(2, 3, 4);

/* Replaces reference inc/simd/convolve.h:125 / src/convolve.c:368-379. */
void convolve_finalize(ConvolutionHandle handle);

SIMD_API_END

#endif  /* SIMD_B200_CONVOLVE_H_ */
