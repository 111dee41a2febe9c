/* simd/normalize.h -- the one function of the reference's normalize module that sits next
 * to the convolve / correlate path: minmax1D (the usual step
 * between a correlation result and peak picking).  The 2-D image normalisation functions of
 * /root/reference/inc/simd/normalize.h are out of scope.
 */
#ifndef SIMD_B200_NORMALIZE_H_
#define SIMD_B200_NORMALIZE_H_

#include <stddef.h>
#include <simd/common.h>
#include <simd/attributes.h>

SIMD_API_BEGIN

/* Replaces reference inc/simd/normalize.h:81-90 / src/normalize.c:323-380,425-443,503-521.
 * Minimum and maximum of src[0..length); either result pointer may be NULL (then that value
 * is not produced).  src: host or device pointer, length > 0; min / max: host pointers.
 * NaN samples are ignored (fminf / fmaxf), as the reference's scalar loop ignores them through
 * its `<` / `>` updates -- except that the reference returns NaN when src[0] itself is NaN
 * (it starts from src[0]); here a leading NaN is ignored like any other. */
void minmax1D(int simd, const float *src, int length, float *min, float *max) NOTNULL(2);

/* Additive: device-resident and asynchronous on the library stream.  d_minmax[0] = minimum,
 * d_minmax[1] = maximum (device memory, 2 floats). */
void simd_b200_minmax1D_device(const float *d_src, size_t length, float *d_minmax) NOTNULL(1, 3);

SIMD_API_END

#endif  /* SIMD_B200_NORMALIZE_H_ */
