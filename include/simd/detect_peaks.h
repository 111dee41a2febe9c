/* simd/detect_peaks.h -- local extrema of a float series, B200 implementation behind the
 * veles.simd API.  Drop-in for /root/reference/inc/simd/detect_peaks.h:40-63 (same enum,
 * struct and function).  This is the first "next" row of the hot-path scope table
 * (SURVEY.md section 8(f)): the classic consumer of a cross-correlation result.
 *
 * A point i in [1, size-2] is an extremum when (d[i]-d[i-1]) * (d[i]-d[i+1]) > 0 in float
 * arithmetic: a maximum if d[i] > d[i-1], a minimum otherwise
 * (/root/reference/src/detect_peaks.c:41-56).  Results are ordered by position.
 */
#ifndef SIMD_B200_DETECT_PEAKS_H_
#define SIMD_B200_DETECT_PEAKS_H_

#include <stddef.h>
#include <simd/common.h>
#include <simd/attributes.h>

SIMD_API_BEGIN

/* reference: detect_peaks.h:40-44 -- values are part of the ABI (bit mask) */
typedef enum {
  kExtremumTypeMaximum = 1,
  kExtremumTypeMinimum,
  kExtremumTypeBoth
} ExtremumType;

/* reference: detect_peaks.h:46-49 */
typedef struct {
  int position;
  float value;
} ExtremumPoint;

/* Replaces reference inc/simd/detect_peaks.h:60-63 / src/detect_peaks.c:58-127.
 * data: host or device pointer, size > 2.  *results is allocated with malloc() (host memory,
 * release with free()) and set to NULL when nothing is found; *resultsLength receives the
 * count.  Both values of `simd` run the GPU kernels; the result is what the reference's
 * scalar loop produces (its AVX loop re-checks block boundaries and reads data[-1] for
 * size < 9: those artefacts are not reproduced). */
void detect_peaks(int simd, const float *data, size_t size, ExtremumType type,
                  ExtremumPoint **results, size_t *resultsLength) NOTNULL(2, 5, 6);

/* Additive: everything stays on the device.  d_results must hold `capacity` points; at most
 * `capacity` are written (in position order) and the total number found is returned, so a
 * return value above capacity means the list was truncated.  Work is enqueued on the library
 * stream; the call waits only for the 8-byte count. */
size_t simd_b200_detect_peaks_device(const float *d_data, size_t size, ExtremumType type,
                                     ExtremumPoint *d_results, size_t capacity) NOTNULL(1);

SIMD_API_END

#endif  /* SIMD_B200_DETECT_PEAKS_H_ */
