// dwt.cu -- one level of the discrete wavelet transform on the GPU: the highpass and the
// lowpass filter applied together, decimated by two or stationary (dilated taps).
//
// Replaces /root/reference/src/wavelet.c: the scalar loops wavelet_apply_na (:271-323) and
// stationary_wavelet_apply_na (:325-382) and their hand-unrolled AVX/NEON variants
// (:395-1876), including the extension of the signal past its end (:248-269):
//
//   hi[i] = sum_{k < order} hp[k] * S(step*i + k*dil),   lo[i] likewise with lp,
//   S(n)  = src[n] for n < length, else ext(n - length)
//   decimated: step 2, dil 1, i < length/2;  stationary: step 1, dil 2^(level-1), i < length.
//
// Layout.  A CTA produces kTile consecutive outputs.  Its input window (step*kTile samples
// plus the (order-1)*dil samples the last output reaches past them) is staged in shared
// memory once, extension applied while staging, so the tap loop has no bounds logic.  For
// the decimated transform the window is stored de-interleaved -- even samples in one array,
// odd samples in another -- so that outputs i, i+1, ... read consecutive words for every tap:
// 16-byte conflict-free shared-memory reads into a sliding register window, 16-byte global
// loads and stores.  The two filters ride in the kernel parameters (constant bank).
//
// Roofline: 4 B read + 4 B written per input sample (decimated: two half-length outputs;
// stationary: 4 B read + 8 B written), `order` (stationary: 2*order) FMAs per sample ->
// HBM-bound for the short filters this transform is normally used with.
//
// Windows that do not fit shared memory (stationary transform at a high level) take the
// unstaged kernel, which applies the extension rule per tap.
#include "b200_device.cuh"
#include <mutex>

namespace {

constexpr int kThreads = 256;
constexpr int kR = 8;                    // outputs per thread
constexpr int kTile = kThreads * kR;     // outputs per CTA
constexpr int kMaxOrder = 76;
constexpr int kMaxHalo = 24576;          // floats of window past the tile that may be staged

struct WaveletParams {
  const float *src;
  i64 length;
  float *hi, *lo;
  i64 n_out;
  int order, dil, ext;
  int halo;                              // (order - 1) * dil
  float hp[kMaxOrder], lp[kMaxOrder];
};

// S(n) for n >= length: the reference's initialize_extension (src/wavelet.c:248-269)
__device__ __forceinline__ float extended(const WaveletParams &p, i64 n) {
  const i64 e = n - p.length;
  switch (p.ext) {
    case 0: return p.src[e % p.length];                       // periodic
    case 1: return p.src[p.length - 1 - (e % p.length)];      // mirror
    case 2: return p.src[p.length - 1];                       // constant
    default: return 0.f;                                      // zero
  }
}
__device__ __forceinline__ float sample(const WaveletParams &p, i64 n) {
  return n < p.length ? p.src[n] : extended(p, n);
}

// ---- decimated transform --------------------------------------------------------------------
// Every thread owns two groups of four consecutive outputs.  Four consecutive outputs and four
// consecutive tap pairs touch seven consecutive words of each parity array, so one 16-byte
// shared-memory read per array and group feeds 32 FMAs: the tap loop is bound by the FP32 pipe,
// not by shared-memory bandwidth, for long filters too.
constexpr int kGroup = 4;                              // consecutive outputs per group
constexpr int kGroups = kR / kGroup;                   // groups per thread
constexpr int kGroupStride = kThreads * kGroup;        // outputs between a thread's groups

__device__ __forceinline__ void load4(float (&w)[8], int at, const float *src) {
  const float4 v = *reinterpret_cast<const float4 *>(src);
  w[at] = v.x; w[at + 1] = v.y; w[at + 2] = v.z; w[at + 3] = v.w;
}

__global__ void __launch_bounds__(kThreads) wavelet_decimated_kernel(const __grid_constant__ WaveletParams p) {
  extern __shared__ float4 smem4[];
  const int npairs = p.order / 2;                // tap pairs (order is even)
  const int half = kTile + ((npairs + 3) & ~3) + 4;   // words per parity array incl. window slack
  const int pitch = (half + 3) & ~3;
  float *ev = reinterpret_cast<float *>(smem4);  // ev[j] = S(base + 2j)
  float *od = ev + pitch;                        // od[j] = S(base + 2j + 1)
  const int tid = threadIdx.x;
  const i64 out0 = (i64)blockIdx.x * kTile;
  const i64 base = 2 * out0;
  const int pairs = half;                        // sample pairs to stage

  const bool aligned = ((reinterpret_cast<unsigned long long>(p.src) & 15ull) == 0);
  if (aligned && base + 2 * (i64)pairs <= p.length) {
    // two pairs per 16-byte load (pairs is a multiple of 4)
    const float4 *s4 = reinterpret_cast<const float4 *>(p.src + base);
    for (int q = tid; q < pairs / 2; q += kThreads) {
      const float4 v = s4[q];
      *reinterpret_cast<float2 *>(ev + 2 * q) = make_float2(v.x, v.z);
      *reinterpret_cast<float2 *>(od + 2 * q) = make_float2(v.y, v.w);
    }
  } else {
    for (int q = tid; q < pairs; q += kThreads) {
      const i64 n = base + 2 * (i64)q;
      ev[q] = sample(p, n);
      od[q] = sample(p, n + 1);
    }
  }
  __syncthreads();

  const bool vec = ((reinterpret_cast<unsigned long long>(p.hi) | reinterpret_cast<unsigned long long>(p.lo)) & 15ull) == 0;
  // one group at a time: 24 live floats per thread keep six CTAs on an SM
#pragma unroll 1
  for (int g = 0; g < kGroups; g++) {
    const float *e0 = ev + g * kGroupStride + kGroup * tid, *o0 = od + g * kGroupStride + kGroup * tid;
    float ah[kGroup], al[kGroup], we[8], wo[8];
#pragma unroll
    for (int j = 0; j < kGroup; j++) ah[j] = al[j] = 0.f;
    load4(we, 0, e0);
    load4(wo, 0, o0);
    for (int q0 = 0; q0 < npairs; q0 += 4) {
      load4(we, 4, e0 + q0 + 4);
      load4(wo, 4, o0 + q0 + 4);
#pragma unroll
      for (int qq = 0; qq < 4; qq++) {
        if (q0 + qq < npairs) {                  // warp-uniform; taps ascend as in the reference
          const int k = 2 * (q0 + qq);
          const float h0 = p.hp[k], l0 = p.lp[k], h1 = p.hp[k + 1], l1 = p.lp[k + 1];
#pragma unroll
          for (int j = 0; j < kGroup; j++) {
            ah[j] = fmaf(h0, we[j + qq], ah[j]);
            al[j] = fmaf(l0, we[j + qq], al[j]);
            ah[j] = fmaf(h1, wo[j + qq], ah[j]);
            al[j] = fmaf(l1, wo[j + qq], al[j]);
          }
        }
      }
#pragma unroll
      for (int j = 0; j < 4; j++) { we[j] = we[j + 4]; wo[j] = wo[j + 4]; }
    }
    const i64 i = out0 + g * kGroupStride + kGroup * tid;
    if (vec && i + kGroup <= p.n_out) {
      *reinterpret_cast<float4 *>(p.hi + i) = make_float4(ah[0], ah[1], ah[2], ah[3]);
      *reinterpret_cast<float4 *>(p.lo + i) = make_float4(al[0], al[1], al[2], al[3]);
    } else {
#pragma unroll
      for (int j = 0; j < kGroup; j++) {
        if (i + j < p.n_out) { p.hi[i + j] = ah[j]; p.lo[i + j] = al[j]; }
      }
    }
  }
}

// ---- stationary transform, window staged -------------------------------------------------------
__global__ void __launch_bounds__(kThreads) wavelet_stationary_kernel(const __grid_constant__ WaveletParams p) {
  extern __shared__ float4 smem4[];
  float *win = reinterpret_cast<float *>(smem4);   // win[j] = S(out0 + j)
  const int tid = threadIdx.x;
  const i64 out0 = (i64)blockIdx.x * kTile;
  const int wlen = kTile + p.halo;

  const bool aligned = ((reinterpret_cast<unsigned long long>(p.src) & 15ull) == 0);
  if (aligned && out0 + wlen <= p.length) {
    const float4 *s4 = reinterpret_cast<const float4 *>(p.src + out0);
    float4 *w4 = reinterpret_cast<float4 *>(win);
    for (int q = tid; q < wlen / 4; q += kThreads) w4[q] = s4[q];
    for (int q = (wlen & ~3) + tid; q < wlen; q += kThreads) win[q] = p.src[out0 + q];
  } else {
    for (int q = tid; q < wlen; q += kThreads) win[q] = sample(p, out0 + q);
  }
  __syncthreads();

  float ah[kR], al[kR];
#pragma unroll
  for (int r = 0; r < kR; r++) ah[r] = al[r] = 0.f;
  const float *a = win + tid;
  for (int k = 0; k < p.order; k++, a += p.dil) {
    const float h = p.hp[k], l = p.lp[k];
#pragma unroll
    for (int r = 0; r < kR; r++) {
      const float v = a[r * kThreads];
      ah[r] = fmaf(h, v, ah[r]);
      al[r] = fmaf(l, v, al[r]);
    }
  }
#pragma unroll
  for (int r = 0; r < kR; r++) {
    const i64 i = out0 + tid + r * kThreads;
    if (i < p.n_out) { p.hi[i] = ah[r]; p.lo[i] = al[r]; }
  }
}

// ---- any step / dilation, nothing staged (windows too long for shared memory) -----------------
__global__ void __launch_bounds__(kThreads) wavelet_unstaged_kernel(const __grid_constant__ WaveletParams p, int step) {
  const i64 i = (i64)blockIdx.x * kThreads + threadIdx.x;
  if (i >= p.n_out) return;
  float ah = 0.f, al = 0.f;
  i64 n = step * i;
  for (int k = 0; k < p.order; k++, n += p.dil) {
    const float v = sample(p, n);
    ah = fmaf(p.hp[k], v, ah);
    al = fmaf(p.lp[k], v, al);
  }
  p.hi[i] = ah;
  p.lo[i] = al;
}

void wavelet_launch(const WaveletParams &p, bool decimate, cudaStream_t stream) {
  static bool configured[16] = {false};
  static std::mutex mu;
  int dev = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  B200_REQUIRE(dev < 16, "device index out of range");
  {
    std::lock_guard<std::mutex> lock(mu);
    if (!configured[dev]) {
      B200_CUDA_CHECK(cudaFuncSetAttribute(wavelet_stationary_kernel,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)(sizeof(float) * (kTile + kMaxHalo + 4))));
      configured[dev] = true;
    }
  }
  const i64 tiles = (p.n_out + kTile - 1) / kTile;
  B200_REQUIRE(tiles < 2147483647LL, "wavelet grid too large");
  if (decimate) {
    const int half = kTile + ((p.order / 2 + 3) & ~3) + 4;
    const size_t smem = sizeof(float) * (size_t)(2 * ((half + 3) & ~3));
    wavelet_decimated_kernel<<<(unsigned)tiles, kThreads, smem, stream>>>(p);
  } else if (p.halo <= kMaxHalo) {
    const size_t smem = sizeof(float) * (size_t)(kTile + p.halo + 4);
    wavelet_stationary_kernel<<<(unsigned)tiles, kThreads, smem, stream>>>(p);
  } else {
    const i64 blocks = (p.n_out + kThreads - 1) / kThreads;
    B200_REQUIRE(blocks < 2147483647LL, "wavelet grid too large");
    wavelet_unstaged_kernel<<<(unsigned)blocks, kThreads, 0, stream>>>(p, 1);
  }
  B200_CUDA_CHECK(cudaGetLastError());
  b200_count_launch(1);
}

bool on_device(const void *ptr) {
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, ptr) != cudaSuccess) { cudaGetLastError(); return false; }
  return attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged;
}

// grow-only staging for host-pointer callers, one per device
struct Staging { float *buf = nullptr; size_t floats = 0; };
Staging g_staging[16];
std::mutex g_staging_mu;

}  // namespace

extern "C" void b200_wavelet_run(const float *src, long long length, float *desthi, float *destlo,
                                 const float *highpass, const float *lowpass, int order,
                                 int dilation, int decimate, int ext) {
  B200_REQUIRE(src && desthi && destlo && length > 0, "wavelet: null array or empty signal");
  B200_REQUIRE(order >= 2 && order <= kMaxOrder && dilation >= 1, "wavelet: bad filter");
  cudaStream_t stream = (cudaStream_t)b200_get_stream();
  WaveletParams p;
  p.length = length;
  p.n_out = decimate ? length / 2 : length;
  p.order = order; p.dil = dilation; p.ext = ext;
  const i64 halo = (i64)(order - 1) * dilation;
  p.halo = halo > (i64)kMaxHalo ? kMaxHalo + 1 : (int)halo;
  for (int k = 0; k < kMaxOrder; k++) {
    p.hp[k] = k < order ? highpass[k] : 0.f;
    p.lp[k] = k < order ? lowpass[k] : 0.f;
  }
  const bool dev_src = on_device(src), dev_hi = on_device(desthi), dev_lo = on_device(destlo);
  B200_REQUIRE(dev_src == dev_hi && dev_src == dev_lo,
               "wavelet: src, desthi and destlo must all be host or all be device pointers");
  if (dev_src) {
    p.src = src; p.hi = desthi; p.lo = destlo;
    wavelet_launch(p, decimate != 0, stream);
    return;
  }
  // host pointers: stage in, run, copy both halves back, return when they have landed
  int dev = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  B200_REQUIRE(dev < 16, "device index out of range");
  std::lock_guard<std::mutex> lock(g_staging_mu);
  Staging &st = g_staging[dev];
  const size_t n_in = (size_t)length, n_out = (size_t)p.n_out;
  const size_t in_pad = (n_in + 3) & ~(size_t)3, out_pad = (n_out + 3) & ~(size_t)3;
  const size_t need = in_pad + 2 * out_pad;
  if (st.floats < need) {
    if (st.buf) { B200_CUDA_CHECK(cudaStreamSynchronize(stream)); B200_CUDA_CHECK(cudaFree(st.buf)); }
    B200_CUDA_CHECK(cudaMalloc(&st.buf, sizeof(float) * need));
    st.floats = need;
  }
  float *d_src = st.buf, *d_hi = st.buf + in_pad, *d_lo = d_hi + out_pad;
  B200_CUDA_CHECK(cudaMemcpyAsync(d_src, src, sizeof(float) * n_in, cudaMemcpyHostToDevice, stream));
  p.src = d_src; p.hi = d_hi; p.lo = d_lo;
  wavelet_launch(p, decimate != 0, stream);
  B200_CUDA_CHECK(cudaMemcpyAsync(desthi, d_hi, sizeof(float) * n_out, cudaMemcpyDeviceToHost, stream));
  B200_CUDA_CHECK(cudaMemcpyAsync(destlo, d_lo, sizeof(float) * n_out, cudaMemcpyDeviceToHost, stream));
  B200_CUDA_CHECK(cudaStreamSynchronize(stream));
}
