/* arithmetic.c -- plain-C host layer of the int16 <-> float converters (B200 build).
 * Same names and arguments as the inline helpers of
 * /root/reference/inc/simd/arithmetic.h:43-57,225-278; the work happens in csrc/convert.cu. */
#define LIBSIMD_IMPLEMENTATION
#include <simd/arithmetic.h>

void b200_int16_to_float(const short *data, long long n, float *res);
void b200_float_to_int16(const float *data, long long n, short *res);

void int16_to_float(const int16_t *data, size_t length, float *res) {
  b200_int16_to_float((const short *)data, (long long)length, res);
}

void float_to_int16(const float *data, size_t length, int16_t *res) {
  b200_float_to_int16(data, (long long)length, (short *)res);
}
