// b200_device.cuh -- shared declarations of the CUDA translation units.
#pragma once

#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include "b200_internal.h"

// The reference reports precondition failures with assert()/abort() and has no error
// codes (src/convolve.c:44-48,105-107); CUDA failures follow the same convention.
#define B200_CUDA_CHECK(expr)                                                         \
  do {                                                                                \
    cudaError_t err__ = (expr);                                                       \
    if (err__ != cudaSuccess) {                                                       \
      fprintf(stderr, "libSimd(b200): %s failed at %s:%d: %s\n", #expr, __FILE__,     \
              __LINE__, cudaGetErrorString(err__));                                   \
      abort();                                                                        \
    }                                                                                 \
  } while (0)

#define B200_REQUIRE(cond, msg)                                                       \
  do {                                                                                \
    if (!(cond)) {                                                                    \
      fprintf(stderr, "libSimd(b200): %s (%s) at %s:%d\n", msg, #cond, __FILE__,      \
              __LINE__);                                                              \
      abort();                                                                        \
    }                                                                                 \
  } while (0)

typedef long long i64;

// Device-pointer job (everything already resident in HBM).
struct DevJob {
  const float *x;
  i64 x_len, x_stride;
  const float *h;
  i64 h_len;
  float *out;
  i64 out_len, out_stride;
  i64 pad;
  i64 batch;
  int reverse;
};

// direct.cu
void b200_direct_launch(const DevJob &job, cudaStream_t stream);

// os_fft.cu
struct OsDevPlan {
  int n_fft;            // complex transform length
  int part_taps;        // impulse-response taps handled per partition
  int max_parts;        // capacity of `spectrum` in partitions
  bool real_packed;     // one real segment of 2*n_fft samples per transform (single partition)
  float2 *spectrum;     // [max_parts][n_fft] filter spectra in register order (scratch)
  const float2 *twiddles;  // W_n^j, j < n_fft/2
};
const float2 *b200_twiddles(int n_fft);          // cached per device, never freed
void b200_os_launch(const OsDevPlan &plan, const DevJob &job, bool single_segment,
                    cudaStream_t stream);
// Same, with caller-supplied spectrum scratch and twiddles (no allocation on this path).
void b200_os_launch_with(const float2 *twiddles, float2 *spectrum, int n_fft, int part_taps,
                         const DevJob &job, bool single_segment, cudaStream_t stream);

void b200_os_launch_real(const float2 *twiddles, float4 *coeffs, int n_fft, const DevJob &job,
                         cudaStream_t stream);

// convert.cu: device-to-device int16 -> float on a given stream (staging of int16 input)
void b200_int16_to_float_on(const short *d_in, i64 n, float *d_out, cudaStream_t stream);

void b200_count_launch(int n);

// Optional per-kernel CUDA-event timing (simd_b200_profile): kind 0 = direct kernel,
// 1 = fused overlap-save kernel, 2 = filter-spectrum kernel.
enum { B200_PROF_DIRECT = 0, B200_PROF_FUSED = 1, B200_PROF_SPECTRUM = 2, B200_PROF_KINDS = 3 };
void b200_prof_begin(int kind, cudaStream_t stream);
void b200_prof_end(int kind, cudaStream_t stream);
void b200_prof_suppress(int on);   // per calling thread
