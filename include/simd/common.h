/* simd/common.h -- linkage / visibility macros of the libSimd (B200) public headers.
 *
 * Same macro names and meaning as /root/reference/inc/simd/common.h:34-51, so code written
 * against the reference headers keeps compiling: SIMD_API_BEGIN / SIMD_API_END wrap the
 * declarations in extern "C" (for C++ callers) with default symbol visibility.
 */
#ifndef SIMD_B200_COMMON_H_
#define SIMD_B200_COMMON_H_

#ifdef __cplusplus
#define SIMD_EXTERN_C_OPEN extern "C" {
#define SIMD_EXTERN_C_CLOSE }
#else
#define SIMD_EXTERN_C_OPEN
#define SIMD_EXTERN_C_CLOSE
#endif

#if defined(__GNUC__) && __GNUC__ >= 4
#define SIMD_VISIBILITY_PUSH _Pragma("GCC visibility push(default)")
#define SIMD_VISIBILITY_POP _Pragma("GCC visibility pop")
#else
#define SIMD_VISIBILITY_PUSH
#define SIMD_VISIBILITY_POP
#endif

#define SIMD_API_BEGIN SIMD_EXTERN_C_OPEN SIMD_VISIBILITY_PUSH
#define SIMD_API_END SIMD_EXTERN_C_CLOSE SIMD_VISIBILITY_POP

#endif  /* SIMD_B200_COMMON_H_ */
