/* normalize.c -- plain-C host layer of minmax1D (B200 build); mirrors the dispatch of
 * /root/reference/src/normalize.c:503-521.  The work happens in csrc/reduce.cu. */
#define LIBSIMD_IMPLEMENTATION
#include <simd/normalize.h>
#include <stdio.h>
#include <stdlib.h>

void b200_minmax1d(const float *src, long long n, float *min, float *max);
void b200_minmax1d_device(const float *d_src, long long n, float *d_minmax);

/* reference: src/normalize.c:503-521 */
void minmax1D(int simd, const float *src, int length, float *min, float *max) {
  (void)simd;
  if (!src || length <= 0) {
    fprintf(stderr, "libSimd(b200): minmax1D: null source or length <= 0\n");
    abort();
  }
  if (!min && !max) return;        /* nothing requested (the reference computes nothing either) */
  b200_minmax1d(src, (long long)length, min, max);
}

void simd_b200_minmax1D_device(const float *d_src, size_t length, float *d_minmax) {
  b200_minmax1d_device(d_src, (long long)length, d_minmax);
}
