/* detect_peaks.c -- plain-C host layer of detect_peaks (B200 build).
 * Mirrors /root/reference/src/detect_peaks.c:58-127: same signature, same ownership of the
 * result (malloc()ed, NULL when empty); the work happens in csrc/peaks.cu.
 */
#define LIBSIMD_IMPLEMENTATION
#include <simd/detect_peaks.h>
#include <stdio.h>
#include <stdlib.h>

unsigned long long b200_detect_peaks_device(const float *d_data, long long size, int type,
                                            void *d_results, unsigned long long capacity);
void b200_detect_peaks(const float *data, long long size, int type, void **results_host,
                       size_t *count);

/* reference: src/detect_peaks.c:58-127 (asserts :60-63) */
void detect_peaks(int simd, const float *data, size_t size, ExtremumType type,
                  ExtremumPoint **results, size_t *resultsLength) {
  (void)simd;
  if (!data || !results || !resultsLength || size <= 2) {
    fprintf(stderr, "libSimd(b200): detect_peaks: null argument or size <= 2\n");
    abort();
  }
  void *host = NULL;
  b200_detect_peaks(data, (long long)size, (int)type, &host, resultsLength);
  *results = (ExtremumPoint *)host;
}

size_t simd_b200_detect_peaks_device(const float *d_data, size_t size, ExtremumType type,
                                     ExtremumPoint *d_results, size_t capacity) {
  return (size_t)b200_detect_peaks_device(d_data, (long long)size, (int)type, d_results,
                                          (unsigned long long)capacity);
}
