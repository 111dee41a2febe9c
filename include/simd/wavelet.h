/* simd/wavelet.h -- one level of the discrete wavelet transform (decimated and stationary),
 * B200 implementation behind the veles.simd API.  Drop-in for
 * /root/reference/inc/simd/wavelet.h:40-162 (same ten functions, same argument meaning).
 * This is the fourth "next" row of the hot-path scope table: the
 * same short-FIR-with-halo stencil as the direct convolution kernel, two filters at once,
 * with decimation and four ways of extending the signal.
 *
 *   decimated  (src/wavelet.c:271-323):   desthi[i] = sum_j hp[j] * S(2i + j),  i < length/2
 *                                         destlo[i] = sum_j lp[j] * S(2i + j)
 *   stationary (src/wavelet.c:325-382):   desthi[i] = sum_j hp[j] * S(i + j*2^(level-1)),
 *                                         i < length (and the same with lp)
 *   S(n) = src[n] for n < length, else the extension (wavelet_types.h) of n - length;
 *   lp = the table row of (type, order), hp[j] = (-1)^(order-j) * lp[order-1-j]
 *   (src/wavelet.c:185-206).
 *
 * src / desthi / destlo may be host or device pointers (all three of the same kind).  With
 * host pointers a call returns when the results are in host memory; with device pointers it
 * is asynchronous on the library stream (simd/b200.h).  The *_na functions and the plain ones
 * run the same GPU kernel: there is no CPU path.  Precondition failures abort with a message
 * (the reference asserts).
 *
 * Filter tables: computed from first principles (tools/gen_wavelet_tables.py).  Daubechies
 * (all orders), Symlets up to 46 taps and Coiflets up to 18 taps are bit-identical to the
 * reference's float tables; Symlets 48..74 and Coiflet 24 differ in the last bits (relative
 * L2 <= 1e-5); Symlet 76 and Coiflet 30 differ by 4e-5 / 2e-5 because the reference carries
 * literature values that are themselves that far from the exact filters.
 */
#ifndef SIMD_B200_WAVELET_H_
#define SIMD_B200_WAVELET_H_

#include <stddef.h>
#include <simd/common.h>
#include <simd/attributes.h>
#include <simd/wavelet_types.h>

SIMD_API_BEGIN

/* Replaces reference wavelet.h:40-44 / src/wavelet.c:83-98.  1 if (type, order) has a table
 * row: Daubechies and Symlets 2, 4, ..., 76; Coiflets 6, 12, ..., 30 (order 0 also answers 1,
 * as in the reference). */
int wavelet_validate_order(WaveletType type, int order);

/* Replaces reference wavelet.h:46-60 / src/wavelet.c:100-119.  Returns a malloc'ed host copy
 * of src (release with free()) in the layout the reference's AVX build produces: the signal,
 * then copies shifted by 2 (and by 4 and 6 when order > 4) at power-of-two offsets, so that
 * callers which size or inspect the block the reference's way keep working.  The GPU kernels
 * only read the first `length` floats. */
float *wavelet_prepare_array(int order, const float *src, size_t length)
    NOTNULL(2) MALLOC WARN_UNUSED_RESULT;

/* Replaces reference wavelet.h:62-72 / src/wavelet.c:121-136.  A malloc'ed host block large
 * enough for one of desthi / destlo (and for a further wavelet_prepare_array-style level). */
float *wavelet_allocate_destination(int order, size_t sourceLength) MALLOC WARN_UNUSED_RESULT;

/* Replaces reference wavelet.h:74-79 / src/wavelet.c:138-164.  Splits a prepared source into
 * the four regions a second decomposition level writes to. */
void wavelet_recycle_source(int order, float *src, size_t length, float **desthihi,
                            float **desthilo, float **destlohi, float **destlolo)
    NOTNULL(2, 4, 5, 6, 7);

/* Replaces reference wavelet.h:81-101 / src/wavelet.c:1878-1905.  length > 0 and even;
 * desthi / destlo hold length/2 floats. */
void wavelet_apply(WaveletType type, int order, ExtensionType ext, const float *__restrict src,
                   size_t length, float *__restrict desthi, float *__restrict destlo)
    NOTNULL(4, 6, 7);

/* Replaces reference wavelet.h:103-121 / src/wavelet.c:271-323. */
void wavelet_apply_na(WaveletType type, int order, ExtensionType ext, const float *__restrict src,
                      size_t length, float *__restrict desthi, float *__restrict destlo)
    NOTNULL(4, 6, 7);

/* Replaces reference wavelet.h:123-145 / src/wavelet.c:1907-1940.  level >= 1; desthi / destlo
 * hold length floats. */
void stationary_wavelet_apply(WaveletType type, int order, int level, ExtensionType ext,
                              const float *__restrict src, size_t length,
                              float *__restrict desthi, float *__restrict destlo)
    NOTNULL(5, 7, 8);

/* Replaces reference wavelet.h:147-162 / src/wavelet.c:325-382. */
void stationary_wavelet_apply_na(WaveletType type, int order, int level, ExtensionType ext,
                                 const float *__restrict src, size_t length,
                                 float *__restrict desthi, float *__restrict destlo)
    NOTNULL(5, 7, 8);

/* Additive (no reference counterpart): the filter pair the kernels use for (type, order),
 * `order` floats each.  Returns 0 (and writes nothing) when the order is not supported. */
int simd_b200_wavelet_filters(WaveletType type, int order, float *highpass, float *lowpass)
    NOTNULL(3, 4);

SIMD_API_END

#endif  /* SIMD_B200_WAVELET_H_ */
