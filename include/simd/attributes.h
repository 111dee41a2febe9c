/* simd/attributes.h -- function attribute shorthands used by the public headers.
 * Names follow /root/reference/inc/simd/attributes.h:34-61 (NOTNULL, MALLOC, INLINE,
 * UNUSED, WARN_UNUSED_RESULT); NOTNULL is dropped while building the library itself
 * (LIBSIMD_IMPLEMENTATION) so that the implementation can still check for NULL and abort
 * with a message instead of having the check optimised away.
 */
#ifndef SIMD_B200_ATTRIBUTES_H_
#define SIMD_B200_ATTRIBUTES_H_

#ifndef INLINE
#define INLINE static inline __attribute__((always_inline))
#endif
#ifndef MALLOC
#define MALLOC __attribute__((__malloc__))
#endif
#ifndef NOTNULL
#ifdef LIBSIMD_IMPLEMENTATION
#define NOTNULL(...)
#else
#define NOTNULL(...) __attribute__((__nonnull__(__VA_ARGS__)))
#endif
#endif
#ifndef UNUSED
#define UNUSED __attribute__((unused))
#endif
#ifndef WARN_UNUSED_RESULT
#define WARN_UNUSED_RESULT __attribute__((__warn_unused_result__))
#endif

#endif  /* SIMD_B200_ATTRIBUTES_H_ */
