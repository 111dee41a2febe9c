/* simd/memory.h -- the four host memory helpers the convolution path and its callers use.
 * Plain C on the host, as in the reference; kept so that programs (and the reference's own
 * tests / benchmarks, tests/convolve.cc, tests/benchmark.inc) that allocate their buffers
 * with them keep linking.  The rest of the reference's memory.h (zeropadding*, crmemcpyf,
 * malloc_aligned_offset, align_complement_*) is not on the path and is out of scope.
 */
#ifndef SIMD_B200_MEMORY_H_
#define SIMD_B200_MEMORY_H_

#include <stddef.h>
#include <simd/common.h>
#include <simd/attributes.h>

SIMD_API_BEGIN

/* 64-byte aligned allocation; free() releases it.  reference: src/memory.c:77-87 */
void *malloc_aligned(size_t size) MALLOC;

/* malloc_aligned(length * sizeof(float)).  reference: src/memory.c:89-91 */
float *mallocf(size_t length) MALLOC;

/* ptr[i] = value for i < length.  reference: src/memory.c:93-123 */
void memsetf(float *ptr, float value, size_t length) NOTNULL(1);

/* dest[i] = src[length-1-i]; returns dest.  reference: src/memory.c:144-174 */
float *rmemcpyf(float *__restrict dest, const float *__restrict src, size_t length)
    NOTNULL(1, 2);

SIMD_API_END

#endif  /* SIMD_B200_MEMORY_H_ */
