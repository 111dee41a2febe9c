/* simd/arithmetic.h -- the two sample-format converters of the reference's arithmetic module
 * that feed the convolve / correlate path (PCM sources are
 * int16; converting on the device halves the host-to-device bytes).  In the reference these
 * are `static inline` functions of inc/simd/arithmetic.h (scalar: :43-57, AVX2: :225-278);
 * here they are exported library functions with the same names and arguments.  The rest of
 * that header (int32 / float16 converters, real / complex multiplies, sums) is out of scope.
 */
#ifndef SIMD_B200_ARITHMETIC_H_
#define SIMD_B200_ARITHMETIC_H_

#include <stddef.h>
#include <stdint.h>
#include <simd/common.h>
#include <simd/attributes.h>

SIMD_API_BEGIN

/* res[i] = (float)data[i].  Replaces reference inc/simd/arithmetic.h:43-48 (scalar
 * int16_to_float_na, the semantics followed) and :225-248 (AVX2).  data, res: host or device
 * pointers, both of the same kind; with device pointers the call is asynchronous on the
 * library stream. */
void int16_to_float(const int16_t *data, size_t length, float *res) NOTNULL(1, 3);

/* res[i] = (int16_t)data[i], truncation toward zero (reference :50-57, "simple truncation");
 * values outside [-32768, 32767] saturate like the reference's AVX pack (:250-278) does
 * (the scalar C cast is undefined there); NaN gives 0. */
void float_to_int16(const float *data, size_t length, int16_t *res) NOTNULL(1, 3);

SIMD_API_END

#endif  /* SIMD_B200_ARITHMETIC_H_ */
