/* oracle/fftf_shim/fftf/api.h -- TEST INFRASTRUCTURE, not product code.
 *
 * Stand-in for <fftf/api.h> of Samsung FFTF (third-party, not vendored in the reference,
 * not installable here; required by /root/reference/configure.ac:47-50 and included at
 * /root/reference/src/convolve.c:37).  Only the names the reference's convolve path uses
 * are declared (call sites: src/convolve.c:131-143,150-151,175,198,216,264-276,285-286,
 * 309,323); semantics are documented in ../../fft_pow2.h.
 */
#ifndef ORACLE_FFTF_SHIM_API_H_
#define ORACLE_FFTF_SHIM_API_H_

#ifdef __cplusplus
extern "C" {
#endif

typedef enum { FFTF_TYPE_COMPLEX = 0, FFTF_TYPE_REAL = 1 } FFTFType;
typedef enum { FFTF_DIRECTION_FORWARD = -1, FFTF_DIRECTION_BACKWARD = 1 } FFTFDirection;
typedef enum { FFTF_DIMENSION_1D = 1 } FFTFDimension;
typedef int FFTFOptions;
#define FFTF_NO_OPTIONS 0

typedef struct FFTFInstance FFTFInstance;

FFTFInstance *fftf_init(FFTFType type, FFTFDirection direction,
                        FFTFDimension dimension, const int *lengths,
                        FFTFOptions options, const float *input, float *output);

FFTFInstance *fftf_init_batch(FFTFType type, FFTFDirection direction,
                              FFTFDimension dimension, const int *lengths,
                              FFTFOptions options, int batchSize,
                              const float *const *inputs, float *const *outputs);

void fftf_calc(const FFTFInstance *instance);

void fftf_destroy(FFTFInstance *instance);

#ifdef __cplusplus
}
#endif
#endif  /* ORACLE_FFTF_SHIM_API_H_ */
