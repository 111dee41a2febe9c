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

/* Cost model fitted to B200 timings (profiles/selector_sweep_r02.md, device-resident data).
 *
 *   direct kernel        t = 9 us + max(1.64e-6 us per output                      (HBM floor)
 *                                       0.9e-6 us per output + 3.5e-8 us per multiply-add)
 *                        (one launch; 3.8e-8 above 512 taps, where fewer CTAs fit an SM)
 *   fused overlap-save   t = 15 us + max(1.85e-6 us per output                     (HBM floor)
 *                                        c(N) * outputs/67.1e6 * N/(N-taps+1) per partition)
 *                        (two launches; c(N) = us per 67.1e6 outputs at 100 % segment use)
 *
 * c(N) is not monotone in N: transforms up to 2048 points run many CTAs per SM and overlap
 * shared-memory traffic with arithmetic, the 8192/16384-point plans run 2 / 1 CTA per SM. */
static const struct { int n; double c; double real_ratio; } kFusedCost[] = {
  /* N, paired cost c(N), cost of a real-packed transform relative to a paired one
   * (tools/real_ratio.py; 256 / 512 extrapolated from the C3 shape) */
  {256, 104.0, 1.6}, {512, 147.0, 1.5}, {1024, 150.0, 1.34}, {2048, 149.0, 1.20},
  {4096, 192.0, 1.12}, {8192, 203.0, 1.39}, {16384, 237.0, 1.35},
};
#define B200_DIRECT_US_PER_MAC 3.5e-8
#define B200_DIRECT_US_PER_MAC_LONG 3.8e-8     /* more than 512 taps */
#define B200_DIRECT_US_PER_OUTPUT 0.9e-6
#define B200_DIRECT_FLOOR_US_PER_OUTPUT 1.64e-6
#define B200_FUSED_FLOOR_US_PER_OUTPUT 1.85e-6
#define B200_DIRECT_LAUNCH_US 9.0
#define B200_FUSED_LAUNCH_US 15.0

static double fused_cost_us(int n, double c, long long h_len, long long out_len,
                            long long batch) {
  long long taps = h_len < n / 2 ? h_len : n / 2;
  long long parts = (h_len + taps - 1) / taps;
  double step = (double)(n - taps + 1);
  /* whole segment pairs are executed, so short outputs pay for a full pair */
  double nseg = (double)((out_len + (long long)step - 1) / (long long)step);
  double pairs = (double)(((long long)nseg + 1) / 2);
  double points = pairs * 2.0 * n * (double)batch;       /* real samples transformed */
  double work = (double)parts * c * points / 67.1e6;
  double floor_us = B200_FUSED_FLOOR_US_PER_OUTPUT * (double)out_len * (double)batch;
  return B200_FUSED_LAUNCH_US + (work > floor_us ? work : floor_us);
}

/* Transform length for an impulse response of h_len taps producing out_len outputs per
 * signal; impulse responses longer than N/2 taps are split into partitions of N/2 taps. */
int b200_os_pick_n(long long h_len, long long out_len) {
  /* tuning knob for selector sweeps (profiles/): VELES_B200_FORCE_N=<power of two> */
  const char *force = getenv("VELES_B200_FORCE_N");
  if (force) {
    int n = atoi(force);
    if (n >= B200_FFT_MIN && n <= B200_FFT_MAX && (n & (n - 1)) == 0) return n;
  }
  int best = B200_FFT_MAX;
  double best_cost = 1e300;
  for (unsigned i = 0; i < sizeof(kFusedCost) / sizeof(kFusedCost[0]); i++) {
    double cost = fused_cost_us(kFusedCost[i].n, kFusedCost[i].c, h_len, out_len, 1);
    if (cost < best_cost * 0.999) { best_cost = cost; best = kFusedCost[i].n; }
  }
  return best;
}

/* Real-packed mode: one real segment of 2N samples per N-point transform.  Its transform
 * costs real_ratio times the paired one (one more shared-memory exchange, two more barriers,
 * twice the coefficient bytes) but yields 2N - taps + 1 outputs instead of 2(N - taps + 1). */
static double real_cost_us(int n, double c, double ratio, long long h_len, long long out_len,
                           long long batch) {
  double step = (double)(2 * n - (h_len - 1) - 1);
  double nseg = (double)((out_len + (long long)step - 1) / (long long)step);
  return B200_FUSED_LAUNCH_US + ratio * c * nseg * 2.0 * n * (double)batch / 67.1e6;
}

void b200_os_pick_mode(long long h_len, long long out_len, int *n_fft, int *real_packed) {
  int n_a = b200_os_pick_n(h_len, out_len);
  *n_fft = n_a;
  *real_packed = 0;
  const char *force = getenv("VELES_B200_REAL");        /* tuning knob: 0 / 1 */
  if (force && atoi(force) == 0) return;
  double best = 1e300;
  for (unsigned i = 0; i < sizeof(kFusedCost) / sizeof(kFusedCost[0]); i++)
    if (kFusedCost[i].n == n_a) best = fused_cost_us(n_a, kFusedCost[i].c, h_len, out_len, 1);
  if (force && atoi(force) == 1) best = 1e300;
  const char *force_n = getenv("VELES_B200_FORCE_N");
  for (unsigned i = 0; i < sizeof(kFusedCost) / sizeof(kFusedCost[0]); i++) {
    int n = kFusedCost[i].n;
    if (h_len > n) continue;                            /* all taps in one transform */
    if (force_n && atoi(force_n) != n) continue;
    double cost = real_cost_us(n, kFusedCost[i].c, kFusedCost[i].real_ratio, h_len, out_len, 1);
    if (cost < best * 0.98) { best = cost; *n_fft = n; *real_packed = 1; }
  }
}

/* Handle-time decision for `batch` signals (the reference API has no batch: batch = 1). */
ConvolutionAlgorithm simd_b200_select_algorithm_batch(size_t x_length, size_t h_length,
                                                      size_t batch) {
  long long out_len = (long long)x_length + (long long)h_length - 1;
  /* direct kernel: throughput term when the grid fills the GPU, otherwise the time of its
   * waves of 2048-output tiles (one 128-thread CTA does 128 multiply-adds per clock, ~4 CTAs
   * per SM on 148 SMs) -- a few tiles with a long h are slow however small the total is */
  double tiles = (double)((out_len + 2047) / 2048) * (double)batch;
  double waves = (double)((long long)((tiles + 591.0) / 592.0));
  double per_wave_us = (double)h_length * 8.2e-3;
  /* fewer tiles than SMs: the small-problem kernel (512-output tiles, 4 outputs per thread)
   * runs instead; its serial chain per thread is a quarter as long */
  if (tiles < 148.0 && h_length <= 1024) per_wave_us = (double)h_length * 2.1e-3;
  double outputs = (double)out_len * (double)batch;
  double direct_work = B200_DIRECT_US_PER_OUTPUT * outputs +
                       (h_length > 512 ? B200_DIRECT_US_PER_MAC_LONG : B200_DIRECT_US_PER_MAC) *
                           (double)batch * (double)x_length * (double)h_length;
  if (B200_DIRECT_FLOOR_US_PER_OUTPUT * outputs > direct_work)
    direct_work = B200_DIRECT_FLOOR_US_PER_OUTPUT * outputs;
  if (waves * per_wave_us > direct_work) direct_work = waves * per_wave_us;
  double direct = B200_DIRECT_LAUNCH_US + direct_work;
  int n = b200_os_pick_n((long long)h_length, out_len);
  double c = 255.0;
  for (unsigned i = 0; i < sizeof(kFusedCost) / sizeof(kFusedCost[0]); i++)
    if (kFusedCost[i].n == n) c = kFusedCost[i].c;
  double fused = fused_cost_us(n, c, (long long)h_length, out_len, (long long)batch);
  if (direct <= fused) return kConvolutionAlgorithmBruteForce;
  /* the same integer test convolve_overlap_save_initialize asserts (h < x/2, reference
   * src/convolve.c:105), so that x = 2h+1 goes to the FFT method instead of aborting there */
  if (x_length / 2 > h_length) return kConvolutionAlgorithmOverlapSave;
  return kConvolutionAlgorithmFFT;
}

ConvolutionAlgorithm simd_b200_select_algorithm(size_t x_length, size_t h_length) {
  return simd_b200_select_algorithm_batch(x_length, h_length, 1);
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
