/* memory.c -- host memory helpers kept from the reference API (simd/memory.h).
 * Plain portable C: these are not on the GPU path (zero padding and impulse-response
 * reversal happen while the kernels stage their operands); they exist so programs that
 * allocate and fill their host buffers with them keep working.
 */
#define LIBSIMD_IMPLEMENTATION
#define _POSIX_C_SOURCE 200112L
#include <simd/memory.h>
#include <stdlib.h>

/* reference: src/memory.c:77-87 (64-byte alignment) */
void *malloc_aligned(size_t size) {
  void *ptr = NULL;
  if (posix_memalign(&ptr, 64, size ? size : 64) != 0) return NULL;
  return ptr;
}

/* reference: src/memory.c:89-91 */
float *mallocf(size_t length) { return (float *)malloc_aligned(length * sizeof(float)); }

/* reference: src/memory.c:93-123 */
void memsetf(float *ptr, float value, size_t length) {
  for (size_t i = 0; i < length; i++) ptr[i] = value;
}

/* reference: src/memory.c:144-174 */
float *rmemcpyf(float *__restrict dest, const float *__restrict src, size_t length) {
  for (size_t i = 0; i < length; i++) dest[i] = src[length - 1 - i];
  return dest;
}
