// reduce.cu -- minmax1D on the GPU (replaces /root/reference/src/normalize.c:323-380 AVX,
// :425-443 scalar, dispatch :503-521).  Two launches: a grid-stride pass with 16-byte loads
// that leaves one {min, max} per CTA, and a one-CTA pass over those partials.  4 B of HBM
// traffic per sample; min / max are exact, so the result equals the reference bit for bit
// (for inputs without NaN).
#include "b200_device.cuh"
#include <float.h>
#include <mutex>

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ void warp_minmax(float &lo, float &hi) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
}

__device__ __forceinline__ void block_minmax(float &lo, float &hi) {
  __shared__ float s_lo[kThreads / 32], s_hi[kThreads / 32];
  warp_minmax(lo, hi);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) { s_lo[wid] = lo; s_hi[wid] = hi; }
  __syncthreads();
  if (wid == 0) {
    lo = lane < kThreads / 32 ? s_lo[lane] : s_lo[0];
    hi = lane < kThreads / 32 ? s_hi[lane] : s_hi[0];
    warp_minmax(lo, hi);
  }
}

__global__ void __launch_bounds__(kThreads) minmax_partial_kernel(const float *src, i64 n, bool vec,
                                                                  float2 *partials) {
  float lo = src[0], hi = src[0];
  const i64 tid = (i64)blockIdx.x * kThreads + threadIdx.x;
  const i64 stride = (i64)gridDim.x * kThreads;
  if (vec) {
    const float4 *s4 = reinterpret_cast<const float4 *>(src);
    const i64 n4 = n >> 2;
    for (i64 i = tid; i < n4; i += stride) {
      const float4 v = s4[i];
      lo = fminf(fminf(lo, v.x), fminf(v.y, fminf(v.z, v.w)));
      hi = fmaxf(fmaxf(hi, v.x), fmaxf(v.y, fmaxf(v.z, v.w)));
    }
    for (i64 i = (n4 << 2) + tid; i < n; i += stride) { lo = fminf(lo, src[i]); hi = fmaxf(hi, src[i]); }
  } else {
    for (i64 i = tid; i < n; i += stride) { lo = fminf(lo, src[i]); hi = fmaxf(hi, src[i]); }
  }
  block_minmax(lo, hi);
  if (threadIdx.x == 0) partials[blockIdx.x] = make_float2(lo, hi);
}

__global__ void __launch_bounds__(kThreads) minmax_final_kernel(const float2 *partials, int n, float *out) {
  float lo = partials[0].x, hi = partials[0].y;
  for (int i = threadIdx.x; i < n; i += kThreads) { lo = fminf(lo, partials[i].x); hi = fmaxf(hi, partials[i].y); }
  block_minmax(lo, hi);
  if (threadIdx.x == 0) { out[0] = lo; out[1] = hi; }
}

struct ReduceScratch { float2 *partials = nullptr; float *result = nullptr; };
ReduceScratch g_reduce[16];
std::mutex g_reduce_mu;
constexpr int kMaxBlocks = 148 * 8;

void minmax_run(const float *d_src, i64 n, float *d_out, cudaStream_t stream) {
  int dev = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  B200_REQUIRE(dev < 16, "device index out of range");
  ReduceScratch &sc = g_reduce[dev];
  {
    std::lock_guard<std::mutex> lock(g_reduce_mu);
    if (!sc.partials) {
      B200_CUDA_CHECK(cudaMalloc(&sc.partials, sizeof(float2) * kMaxBlocks));
      B200_CUDA_CHECK(cudaMalloc(&sc.result, sizeof(float) * 2));
    }
  }
  i64 blocks = (n / 4 + kThreads - 1) / kThreads;
  if (blocks < 1) blocks = 1;
  if (blocks > kMaxBlocks) blocks = kMaxBlocks;
  const bool vec = (((unsigned long long)d_src) & 15ull) == 0;
  minmax_partial_kernel<<<(unsigned)blocks, kThreads, 0, stream>>>(d_src, n, vec, sc.partials);
  minmax_final_kernel<<<1, kThreads, 0, stream>>>(sc.partials, (int)blocks, d_out ? d_out : sc.result);
  B200_CUDA_CHECK(cudaGetLastError());
  b200_count_launch(2);
}

bool reduce_is_device_pointer(const void *p) {
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged;
}

}  // namespace

extern "C" void b200_minmax1d_device(const float *d_src, long long n, float *d_minmax) {
  B200_REQUIRE(d_src && d_minmax && n > 0, "minmax1D: null argument or empty array");
  minmax_run(d_src, n, d_minmax, (cudaStream_t)b200_get_stream());
}

extern "C" void b200_minmax1d(const float *src, long long n, float *min, float *max) {
  B200_REQUIRE(src && n > 0, "minmax1D: null source or empty array");
  cudaStream_t stream = (cudaStream_t)b200_get_stream();
  const float *d_src = src;
  float *staged = nullptr;
  if (!reduce_is_device_pointer(src)) {
    B200_CUDA_CHECK(cudaMalloc(&staged, sizeof(float) * (size_t)n));
    B200_CUDA_CHECK(cudaMemcpyAsync(staged, src, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice, stream));
    d_src = staged;
  }
  minmax_run(d_src, n, nullptr, stream);
  int dev = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  float host[2];
  B200_CUDA_CHECK(cudaMemcpyAsync(host, g_reduce[dev].result, sizeof(host), cudaMemcpyDeviceToHost, stream));
  B200_CUDA_CHECK(cudaStreamSynchronize(stream));
  if (staged) B200_CUDA_CHECK(cudaFree(staged));
  if (min) *min = host[0];
  if (max) *max = host[1];
}
