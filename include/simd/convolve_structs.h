/* simd/convolve_structs.h -- public handle structs of the convolution API.
 *
 * ABI CONTRACT.  These structs are passed BY VALUE across the library boundary, so their
 * size and field offsets are frozen to those of the reference
 * (/root/reference/inc/simd/convolve_structs.h:39-74; x86-64: 64 / 48 / 80 bytes, offsets
 * checked by tests/test_abi.py).  Field names and types are unchanged; what the pointer
 * fields point AT is different on the B200 build:
 *
 *   fft_plan          -> opaque device plan (B200Plan): stream-ordered scratch, staging
 *                        buffers, twiddle tables.  Owned by the handle, freed by *_finalize.
 *   fft_inverse_plan  -> the same plan (kept non-NULL because reference callers may assert
 *                        on it); never freed separately.
 *   fft_boiler_plate  -> DEVICE pointer to the handle's staging area (was: host scratch).
 *   H                 -> DEVICE pointer to the filter spectrum in kernel register order
 *                        (was: host half-spectrum).  Do not dereference on the host.
 *   L / M             -> host-readable int, as in the reference: the transform length in
 *                        use.  L is the B200 segment length (complex points per segment
 *                        pair), which is tuned for the GPU and is NOT the reference's
 *                        4*2^floor(log2 h) (src/convolve.c:116-121).  M is the reference's
 *                        next-power-of-two of x+h-1 (src/convolve.c:237-244).
 *   inputs            -> NULL (the reference's two host FFT buffers have no counterpart).
 *   reverse           -> as in the reference: non-zero = cross-correlation.  It stays in
 *                        the by-value struct because cross_correlate_initialize sets it
 *                        after the inner initialise (src/correlate.c:131-140).
 */
#ifndef SIMD_B200_CONVOLVE_STRUCTS_H_
#define SIMD_B200_CONVOLVE_STRUCTS_H_

#include <stddef.h>
#include <simd/common.h>

SIMD_API_BEGIN

/* reference: convolve_structs.h:39-48 */
struct ConvolutionOverlapSaveHandle {
  void *fft_plan;
  void *fft_inverse_plan;
  float *fft_boiler_plate;
  float *H;
  size_t x_length;
  size_t h_length;
  int *L;
  int reverse;
};

/* reference: convolve_structs.h:50-58 */
struct ConvolutionFFTHandle {
  void *fft_plan;
  void *fft_inverse_plan;
  int *M;
  int x_length;
  int h_length;
  float **inputs;
  int reverse;
};

/* reference: convolve_structs.h:60-64 -- numeric values are part of the ABI */
typedef enum {
  kConvolutionAlgorithmBruteForce,   /* 0: direct kernel            */
  kConvolutionAlgorithmFFT,          /* 1: whole-signal transform   */
  kConvolutionAlgorithmOverlapSave   /* 2: segmented transform      */
} ConvolutionAlgorithm;

/* reference: convolve_structs.h:66-74 */
struct ConvolutionHandle {
  ConvolutionAlgorithm algorithm;
  int x_length;
  int h_length;
  union {
    struct ConvolutionFFTHandle fft;
    struct ConvolutionOverlapSaveHandle os;
  } handle;
};

SIMD_API_END

#endif  /* SIMD_B200_CONVOLVE_STRUCTS_H_ */
