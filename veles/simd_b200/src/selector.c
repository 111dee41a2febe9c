/* selector.c -- handle-time algorithm selection and transform-length policy (pure host
 * logic: no CUDA call, usable on a machine without a GPU).
 *
 * Replaces the reference's hard-coded rules (src/convolve.c:349-363: overlap-save if
 * x > 2h and x > 200; FFT if x <= 2h and x > 350; else brute force), which were tuned on
 * 2013 CPUs against FFTF.  The enum values and their meaning are kept:
 *   BruteForce  = direct kernel                 (cost ~ x*h, one launch)
 *   FFT         = whole-signal transform        (x <= 2h: h is not "short" relative to x)
 *   OverlapSave = segmented transform           (x > 2h)
 * Thresholds below come from B200 timings (profiles/selector_sweep_r01.md).
 */
#define LIBSIMD_IMPLEMENTATION
#include <simd/b200.h>
#include <stdlib.h>
#include "../csrc/b200_internal.h"

/* Direct kernel wins while the impulse response is short (FP32-bound at 2*h FLOP/output,
 * the fused FFT path costs ~50-100 FLOP/output plus a second launch for the spectrum) or
 * while the whole problem is too small to amortise two launches. */
#define B200_DIRECT_MAX_TAPS 64
#define B200_DIRECT_MAX_WORK (1LL << 21)     /* x*h multiply-adds */

ConvolutionAlgorithm simd_b200_select_algorithm(size_t x_length, size_t h_length) {
  long long work = (long long)x_length * (long long)h_length;
  if (h_length <= B200_DIRECT_MAX_TAPS || work <= B200_DIRECT_MAX_WORK) {
    return kConvolutionAlgorithmBruteForce;
  }
  if (x_length > h_length * 2) return kConvolutionAlgorithmOverlapSave;
  return kConvolutionAlgorithmFFT;
}

/* Transform length for an impulse response of h_len taps producing out_len outputs per
 * signal.  Cost model: transforms executed (segment pairs x partitions) x N*(log2 N + 2);
 * impulse responses that would leave less than half a transform of useful output are
 * split into partitions of N/2 taps. */
int b200_os_pick_n(long long h_len, long long out_len) {
  /* tuning knob for selector sweeps (profiles/): VELES_B200_FORCE_N=<power of two> */
  const char *force = getenv("VELES_B200_FORCE_N");
  if (force) {
    int n = atoi(force);
    if (n >= B200_FFT_MIN && n <= B200_FFT_MAX && (n & (n - 1)) == 0) return n;
  }
  int best = B200_FFT_MAX;
  double best_cost = 1e300;
  for (int n = B200_FFT_MIN; n <= B200_FFT_MAX; n *= 2) {
    long long taps = h_len < n / 2 ? h_len : n / 2;
    long long parts = (h_len + taps - 1) / taps;
    long long step = n - taps + 1;
    long long nseg = (out_len + step - 1) / step;
    long long pairs = (nseg + 1) / 2;
    int lg = 0;
    while ((1 << lg) < n) lg++;
    double cost = (double)parts * (double)pairs * (double)n * (lg + 2.0);
    if (cost < best_cost * 0.999) { best_cost = cost; best = n; }
  }
  return best;
}

int b200_os_part_taps(int n_fft, long long h_len) {
  long long cap = n_fft / 2;
  return (int)(h_len < cap ? h_len : cap);
}

static int next_pow2_ge(long long v) {
  int n = 1;
  while (n < v) n <<= 1;
  return n;
}

int simd_b200_transform_length(size_t x_length, size_t h_length) {
  long long out_len = (long long)x_length + (long long)h_length - 1;
  if (x_length <= h_length * 2 && out_len <= B200_FFT_MAX) {
    int n = next_pow2_ge(out_len);
    return n < B200_FFT_MIN ? B200_FFT_MIN : n;     /* whole signal in one transform */
  }
  return b200_os_pick_n((long long)h_length, out_len);
}
