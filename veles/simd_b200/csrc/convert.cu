// convert.cu -- int16 <-> float32 sample converters on the GPU (replace the inline helpers of
// /root/reference/inc/simd/arithmetic.h:43-57, 225-278).  Pure streaming kernels: 6 B of HBM
// traffic per sample, 8 samples per thread through 16-byte accesses when the pointers allow.
#include "b200_device.cuh"
#include <stdint.h>

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ short to_i16(float v) {
  int i = __float2int_rz(v);                 // truncation; saturates at the int range, NaN -> 0
  i = max(-32768, min(32767, i));
  return (short)i;
}

__global__ void __launch_bounds__(kThreads) i16_to_f32_kernel(const short *in, float *out, i64 n, bool vec) {
  const i64 tid = (i64)blockIdx.x * kThreads + threadIdx.x;
  const i64 stride = (i64)gridDim.x * kThreads;
  if (vec) {
    const i64 n8 = n >> 3;
    for (i64 i = tid; i < n8; i += stride) {
      const int4 q = reinterpret_cast<const int4 *>(in)[i];     // 8 samples
      const short *s = reinterpret_cast<const short *>(&q);
      float4 a = make_float4((float)s[0], (float)s[1], (float)s[2], (float)s[3]);
      float4 b = make_float4((float)s[4], (float)s[5], (float)s[6], (float)s[7]);
      reinterpret_cast<float4 *>(out)[2 * i] = a;
      reinterpret_cast<float4 *>(out)[2 * i + 1] = b;
    }
    for (i64 i = (n8 << 3) + tid; i < n; i += stride) out[i] = (float)in[i];
  } else {
    for (i64 i = tid; i < n; i += stride) out[i] = (float)in[i];
  }
}

__global__ void __launch_bounds__(kThreads) f32_to_i16_kernel(const float *in, short *out, i64 n, bool vec) {
  const i64 tid = (i64)blockIdx.x * kThreads + threadIdx.x;
  const i64 stride = (i64)gridDim.x * kThreads;
  if (vec) {
    const i64 n8 = n >> 3;
    for (i64 i = tid; i < n8; i += stride) {
      const float4 a = reinterpret_cast<const float4 *>(in)[2 * i];
      const float4 b = reinterpret_cast<const float4 *>(in)[2 * i + 1];
      int4 q;
      short *s = reinterpret_cast<short *>(&q);
      s[0] = to_i16(a.x); s[1] = to_i16(a.y); s[2] = to_i16(a.z); s[3] = to_i16(a.w);
      s[4] = to_i16(b.x); s[5] = to_i16(b.y); s[6] = to_i16(b.z); s[7] = to_i16(b.w);
      reinterpret_cast<int4 *>(out)[i] = q;
    }
    for (i64 i = (n8 << 3) + tid; i < n; i += stride) out[i] = to_i16(in[i]);
  } else {
    for (i64 i = tid; i < n; i += stride) out[i] = to_i16(in[i]);
  }
}

bool conv_is_device_pointer(const void *p) {
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged;
}

unsigned grid_for(i64 n) {
  i64 blocks = (n / 8 + kThreads - 1) / kThreads;
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 16) blocks = 148 * 16;
  return (unsigned)blocks;
}

// direction 0: int16 -> float, 1: float -> int16
void convert(const void *in, void *out, i64 n, int direction) {
  B200_REQUIRE(in && out, "converter: null pointer");
  if (n <= 0) return;
  cudaStream_t stream = (cudaStream_t)b200_get_stream();
  const size_t in_bytes = (size_t)n * (direction == 0 ? 2 : 4), out_bytes = (size_t)n * (direction == 0 ? 4 : 2);
  const bool ind = conv_is_device_pointer(in), outd = conv_is_device_pointer(out);
  const void *d_in = in;
  void *d_out = out, *tmp_in = nullptr, *tmp_out = nullptr;
  if (!ind) {
    B200_CUDA_CHECK(cudaMalloc(&tmp_in, in_bytes));
    B200_CUDA_CHECK(cudaMemcpyAsync(tmp_in, in, in_bytes, cudaMemcpyHostToDevice, stream));
    d_in = tmp_in;
  }
  if (!outd) {
    B200_CUDA_CHECK(cudaMalloc(&tmp_out, out_bytes));
    d_out = tmp_out;
  }
  const bool vec = ((((unsigned long long)d_in) | ((unsigned long long)d_out)) & 15ull) == 0;
  if (direction == 0)
    i16_to_f32_kernel<<<grid_for(n), kThreads, 0, stream>>>((const short *)d_in, (float *)d_out, n, vec);
  else
    f32_to_i16_kernel<<<grid_for(n), kThreads, 0, stream>>>((const float *)d_in, (short *)d_out, n, vec);
  B200_CUDA_CHECK(cudaGetLastError());
  b200_count_launch(1);
  if (!outd) B200_CUDA_CHECK(cudaMemcpyAsync(out, d_out, out_bytes, cudaMemcpyDeviceToHost, stream));
  if (!ind || !outd) {
    B200_CUDA_CHECK(cudaStreamSynchronize(stream));
    if (tmp_in) B200_CUDA_CHECK(cudaFree(tmp_in));
    if (tmp_out) B200_CUDA_CHECK(cudaFree(tmp_out));
  }
}

}  // namespace

void b200_int16_to_float_on(const short *d_in, i64 n, float *d_out, cudaStream_t stream) {
  if (n <= 0) return;
  const bool vec = ((((unsigned long long)d_in) | ((unsigned long long)d_out)) & 15ull) == 0;
  i16_to_f32_kernel<<<grid_for(n), kThreads, 0, stream>>>(d_in, d_out, n, vec);
  B200_CUDA_CHECK(cudaGetLastError());
  b200_count_launch(1);
}

extern "C" void b200_int16_to_float(const short *data, long long n, float *res) { convert(data, res, n, 0); }
extern "C" void b200_float_to_int16(const float *data, long long n, short *res) { convert(data, res, n, 1); }
