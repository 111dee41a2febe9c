/* oracle/peaks_oracle.c -- TEST INFRASTRUCTURE, not product code.
 * CPU restatement of detect_peaks (scalar path), /root/reference/src/detect_peaks.c:41-56
 * (check_peak) and :118-125 (the simd == 0 loop).  Parity status: PINNED against the
 * reference's own tests (tests/detect_peaks.cc:43-100: sin and nasty_peaks) and against the
 * reference's source compiled into oracle/_ref (bit-exact, simd = 0).
 * Output: positions[i], values[i] for the first `capacity` extrema; returns the total found.
 */
#include <stddef.h>

__attribute__((visibility("default")))
size_t oracle_detect_peaks(const float *data, size_t size, int type, int *positions,
                           float *values, size_t capacity) {
  size_t n = 0;
  for (long i = 1; i < (long)size - 1; i++) {            /* :118-125 */
    float prev = data[i - 1], curr = data[i], next = data[i + 1];
    float delta1 = curr - prev, delta2 = curr - next;    /* :47-48 */
    if (delta1 * delta2 > 0) {                           /* :49 */
      if ((delta1 > 0 && (type & 1) != 0) || (delta1 < 0 && (type & 2) != 0)) {   /* :50-51 */
        if (n < capacity) { positions[n] = (int)i; values[n] = curr; }
        n++;
      }
    }
  }
  return n;
}

/* minmax1D, scalar path: /root/reference/src/normalize.c:425-443 (minmax1D_novec).  Pinned
 * against the reference's test (tests/normalize.cc:65-81) and oracle/_ref (bit-exact). */
__attribute__((visibility("default")))
void oracle_minmax1D(const float *src, int length, float *min_out, float *max_out) {
  float mn = src[0], mx = src[0];
  for (int i = 0; i < length; i++) {
    float val = src[i];
    if (val < mn) mn = val;
    if (val > mx) mx = val;
  }
  if (min_out) *min_out = mn;
  if (max_out) *max_out = mx;
}

/* int16 <-> float converters, scalar semantics of /root/reference/inc/simd/arithmetic.h:43-57
 * (int16_to_float_na, float_to_int16_na: plain casts, truncation toward zero).  Out-of-range
 * floats saturate as the reference's AVX pack does (:250-278); the scalar cast is undefined
 * there, so parity tests stay in range.  Pinned bit-exact against the reference's own inline
 * functions called through oracle/_ref (ref_harness.c). */
#include <stdint.h>
__attribute__((visibility("default")))
void oracle_int16_to_float(const int16_t *data, size_t length, float *res) {
  for (size_t i = 0; i < length; i++) res[i] = (float)data[i];
}
__attribute__((visibility("default")))
void oracle_float_to_int16(const float *data, size_t length, int16_t *res) {
  for (size_t i = 0; i < length; i++) {
    float v = data[i];
    if (v != v) { res[i] = 0; continue; }
    if (v >= 32767.f) { res[i] = 32767; continue; }
    if (v <= -32768.f) { res[i] = -32768; continue; }
    res[i] = (int16_t)v;
  }
}
