/* simd/wavelet_types.h -- enumerations of the wavelet module; values are part of the ABI.
 * Drop-in for /root/reference/inc/simd/wavelet_types.h:38-53. */
#ifndef SIMD_B200_WAVELET_TYPES_H_
#define SIMD_B200_WAVELET_TYPES_H_

#include <simd/common.h>

SIMD_API_BEGIN

/* reference: wavelet_types.h:38-42 */
typedef enum {
  WAVELET_TYPE_DAUBECHIES,
  WAVELET_TYPE_COIFLET,
  WAVELET_TYPE_SYMLET
} WaveletType;

/* reference: wavelet_types.h:44-53.  How the signal continues past its last sample
 * (shown for the signal 1 2 3). */
typedef enum {
  EXTENSION_TYPE_PERIODIC,   /* 1 2 3 | 1 2 3 */
  EXTENSION_TYPE_MIRROR,     /* 1 2 3 | 3 2 1 */
  EXTENSION_TYPE_CONSTANT,   /* 1 2 3 | 3 3 3 */
  EXTENSION_TYPE_ZERO        /* 1 2 3 | 0 0 0 */
} ExtensionType;

SIMD_API_END

#endif  /* SIMD_B200_WAVELET_TYPES_H_ */
