/* simd/correlate.h -- 1D cross-correlation, B200 (sm_100a) implementation behind the
 * veles.simd C API.  Drop-in for /root/reference/inc/simd/correlate.h: same ten symbols.
 *
 * result[k] = sum_m x[m] * h[m + hLength-1-k], k in [0, xLength+hLength-1)
 * (src/correlate.c:78-84,124), i.e. convolution with the time-reversed h; the FFT and
 * overlap-save variants are the convolution handles with reverse = 1
 * (src/correlate.c:37-72).  Pointer, error and threading rules: see simd/convolve.h.
 */
#ifndef SIMD_B200_CORRELATE_H_
#define SIMD_B200_CORRELATE_H_

#include <stddef.h>
#include <simd/common.h>
#include <simd/attributes.h>
#include <simd/convolve_structs.h>

SIMD_API_BEGIN

typedef struct ConvolutionFFTHandle CrossCorrelationFFTHandle;                 /* correlate.h:41 */
typedef struct ConvolutionOverlapSaveHandle CrossCorrelationOverlapSaveHandle; /* correlate.h:66 */
typedef struct ConvolutionHandle CrossCorrelationHandle;                       /* correlate.h:110 */

/* Replaces reference inc/simd/correlate.h:48-49 / src/correlate.c:37-42. */
CrossCorrelationFFTHandle cross_correlate_fft_initialize(size_t xLength, size_t hLength);

/* Replaces reference inc/simd/correlate.h:58-60 / src/correlate.c:44-48. */
void cross_correlate_fft(CrossCorrelationFFTHandle handle, const float *x, const float *h,
                         float *result) NOTNULL(2, 3, 4);

/* Replaces reference inc/simd/correlate.h:64 / src/correlate.c:50-52. */
void cross_correlate_fft_finalize(CrossCorrelationFFTHandle handle);

/* Replaces reference inc/simd/correlate.h:73-74 / src/correlate.c:54-60. */
CrossCorrelationOverlapSaveHandle cross_correlate_overlap_save_initialize(size_t xLength,
                                                                          size_t hLength);

/* Replaces reference inc/simd/correlate.h:84-87 / src/correlate.c:62-67. */
void cross_correlate_overlap_save(CrossCorrelationOverlapSaveHandle handle,
                                  const float *__restrict x, const float *__restrict h,
                                  float *result) NOTNULL(2, 3, 4);

/* Replaces reference inc/simd/correlate.h:91-92 / src/correlate.c:69-72. */
void cross_correlate_overlap_save_finalize(CrossCorrelationOverlapSaveHandle handle);

/* Replaces reference inc/simd/correlate.h:105-108 / src/correlate.c:74-126. */
void cross_correlate_simd(int simd, const float *__restrict x, size_t xLength,
                          const float *__restrict h, size_t hLength,
                          float *__restrict result) NOTNULL(2, 4, 6);

/* Replaces reference inc/simd/correlate.h:118-119 / src/correlate.c:128-142. */
CrossCorrelationHandle cross_correlate_initialize(size_t xLength, size_t hLength);

/* Replaces reference inc/simd/correlate.h:127-129 / src/correlate.c:144-156. */
void cross_correlate(CrossCorrelationHandle handle, const float *__restrict x,
                     const float *__restrict h, float *__restrict result) NOTNULL(2, 3, 4);

/* Replaces reference inc/simd/correlate.h:133-134 / src/correlate.c:158-160. */
void cross_correlate_finalize(CrossCorrelationHandle handle);

SIMD_API_END

#endif  /* SIMD_B200_CORRELATE_H_ */
