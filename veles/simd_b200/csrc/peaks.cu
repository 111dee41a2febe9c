// peaks.cu -- detect_peaks on the GPU: ordered stream compaction of local extrema.
//
// Replaces /root/reference/src/detect_peaks.c:41-127.  Three launches:
//   peaks_count_kernel   every warp classifies its 256 samples and adds its count to its
//                        tile's (2048 samples) -- one atomic per warp, no barrier
//   peaks_sums_kernel    sums of the tile counts per group (32 tiles) and per block of groups
//                        (1024 tiles); a few microseconds
//   peaks_write_kernel   every CTA re-classifies its tile and writes {position, value} at
//                        offset + in-tile rank, which keeps the list ordered by position; the
//                        offset is the sum of the block sums before its block, the group sums
//                        before its group and the tile counts before it in the group (a
//                        hundred or so small L2 loads by one warp while the samples are in
//                        flight: no scan over all tiles); tiles without extrema are not read again
// The series is read twice (8 B per sample) and 8 B are written per extremum: HBM-bound.
#include "b200_device.cuh"
#include <mutex>
#include <vector>

namespace {

constexpr int kThreads = 256;
constexpr int kPerThread = 8;
constexpr int kTile = kThreads * kPerThread;

// Thread t of a tile owns the 8 consecutive samples starting at c0 = tile*kTile + t*8 (a
// multiple of 8, so two aligned 16-byte loads fetch them when the series itself is 16-byte
// aligned); its left / right neighbours come from the adjacent lanes by shuffle, the warp's
// two edge lanes load theirs.  Bit r of the result <=> sample c0 + r is an interior point and
// an extremum of the requested kind (KIND = the reference's bit mask: 1 maxima, 2 minima,
// 3 both).  vals[] receives the 8 samples.
//
// The reference tests (v - prev) * (v - next) > 0 in float arithmetic
// (/root/reference/src/detect_peaks.c:41-56).  IEEE subtraction and multiplication are exactly
// odd in each operand, so with the forward differences dd[r] = sample r - sample r-1 that is
// dd[r] * dd[r+1] < 0 bit for bit: one subtraction per sample instead of two.  A maximum is
// dd[r] > 0.  Interior points (1 <= i <= size-2) are two mask operations per thread.
template <int KIND>
__device__ __forceinline__ unsigned classify(const float *d, i64 c0, i64 size, bool vec,
                                             float (&vals)[kPerThread]) {
  const int lane = threadIdx.x & 31;
  if (vec && c0 + kPerThread <= size) {
    const float4 a = *reinterpret_cast<const float4 *>(d + c0);
    const float4 b = *reinterpret_cast<const float4 *>(d + c0 + 4);
    vals[0] = a.x; vals[1] = a.y; vals[2] = a.z; vals[3] = a.w;
    vals[4] = b.x; vals[5] = b.y; vals[6] = b.z; vals[7] = b.w;
  } else {
#pragma unroll
    for (int r = 0; r < kPerThread; r++) vals[r] = c0 + r < size ? d[c0 + r] : 0.f;
  }
  float left = __shfl_up_sync(0xffffffffu, vals[kPerThread - 1], 1);
  float right = __shfl_down_sync(0xffffffffu, vals[0], 1);
  if (lane == 0) left = c0 > 0 && c0 - 1 < size ? d[c0 - 1] : 0.f;
  if (lane == 31) right = c0 + kPerThread < size ? d[c0 + kPerThread] : 0.f;
  float dd[kPerThread + 1];
  dd[0] = __fsub_rn(vals[0], left);
#pragma unroll
  for (int r = 1; r < kPerThread; r++) dd[r] = __fsub_rn(vals[r], vals[r - 1]);
  dd[kPerThread] = __fsub_rn(right, vals[kPerThread - 1]);
  unsigned m = 0;
#pragma unroll
  for (int r = 0; r < kPerThread; r++) {
    bool e = __fmul_rn(dd[r], dd[r + 1]) < 0.f;
    if (KIND == 1) e = e && dd[r] > 0.f;
    if (KIND == 2) e = e && dd[r] < 0.f;
    m |= (unsigned)e << r;
  }
  if (c0 == 0) m &= ~1u;                                   // sample 0 is not interior
  const i64 last = size - 2 - c0;                          // highest interior r of this thread
  if (last < kPerThread - 1) m &= last < 0 ? 0u : (2u << (int)last) - 1u;
  return m;
}

__device__ __forceinline__ int block_exclusive_scan(int v, int *total, int *smem) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  if (lane == 31) smem[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    int w = lane < kThreads / 32 ? smem[lane] : 0;
    int winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int n = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += n;
    }
    smem[32 + lane] = winc - w;           // exclusive warp offsets
    if (lane == 31) smem[64] = winc;      // block total
  }
  __syncthreads();
  *total = smem[64];
  return smem[32 + wid] + inc - v;
}

// counts[] zeroed before the launch; every warp adds the extrema of its 256 samples to its
// tile's count (no CTA barrier: the warps of a tile do not wait for one another)
template <int KIND>
__global__ void __launch_bounds__(kThreads) peaks_count_kernel(const float *d, i64 size, bool vec,
                                                               unsigned *counts) {
  const i64 c0 = (i64)blockIdx.x * kTile + (i64)threadIdx.x * kPerThread;
  float vals[kPerThread];
  const unsigned c = __reduce_add_sync(0xffffffffu, (unsigned)__popc(classify<KIND>(d, c0, size, vec, vals)));
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(counts + blockIdx.x, c);
}

typedef unsigned long long u64;
constexpr int kGroupTiles = 32;          // tiles per group sum
constexpr int kBlockTiles = 1024;        // tiles per block sum (32 groups)

// group_sums[g] = sum of the 32 tile counts of group g, block_sums[b] = sum of the 1024 tile
// counts of block b; one CTA of 1024 threads per block, one warp per group
__global__ void __launch_bounds__(kBlockTiles) peaks_sums_kernel(const unsigned *counts, i64 tiles,
                                                                unsigned *group_sums, u64 *block_sums) {
  __shared__ unsigned warp_sum[kBlockTiles / 32];
  const i64 tile = (i64)blockIdx.x * kBlockTiles + threadIdx.x;
  const unsigned c = __reduce_add_sync(0xffffffffu, tile < tiles ? counts[tile] : 0u);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) {
    warp_sum[wid] = c;
    if (tile < tiles) group_sums[(i64)blockIdx.x * (kBlockTiles / kGroupTiles) + wid] = c;   // group exists
  }
  __syncthreads();
  if (wid == 0) {
    u64 v = warp_sum[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) block_sums[blockIdx.x] = v;
  }
}

template <int KIND>
__global__ void __launch_bounds__(kThreads) peaks_write_kernel(const float *d, i64 size,
                                                               bool vec, const unsigned *counts,
                                                               const unsigned *group_sums,
                                                               const u64 *block_sums, int2 *out,
                                                               u64 capacity) {
  __shared__ int smem[65];
  __shared__ int2 staged[kTile];          // the tile's points in rank order
  __shared__ u64 s_base;
  // nothing to write for this tile (smooth series: most tiles): do not read it a second time
  if (counts[blockIdx.x] == 0) return;
  const i64 c0 = (i64)blockIdx.x * kTile + (i64)threadIdx.x * kPerThread;
  float vals[kPerThread];
  unsigned m = classify<KIND>(d, c0, size, vec, vals);
  // last warp: extrema before this tile = blocks before its block + groups before its group
  // (inside the block) + tiles before it (inside the group)
  if (threadIdx.x >= kThreads - 32) {
    const int lane = threadIdx.x & 31;
    const unsigned tile = blockIdx.x, blk = tile / kBlockTiles, group = tile / kGroupTiles;
    u64 v = 0;
    for (unsigned k = lane; k < blk; k += 32) v += block_sums[k];
    const unsigned g = blk * (kBlockTiles / kGroupTiles) + lane;      // 32 groups per block
    if (g < group) v += group_sums[g];
    const unsigned t = group * kGroupTiles + lane;
    if (t < tile) v += counts[t];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_base = v;
  }
  int total;
  int rank = block_exclusive_scan(__popc(m), &total, smem);
  const int pos0 = (int)c0;               // positions fit an int (checked by the host)
  int2 *slot = staged + rank;
#pragma unroll
  for (int r = 0; r < kPerThread; r++) {
    if (m & (1u << r)) *slot++ = make_int2(pos0 + r, __float_as_int(vals[r]));
  }
  __syncthreads();
  // coalesced copy of the compacted tile to its place in the global list (clipped once to the
  // caller's capacity)
  const u64 base = s_base;
  int limit = total;
  if (base >= capacity) limit = 0;
  else if (capacity - base < (u64)total) limit = (int)(capacity - base);
  int2 *dst = out + base;
#pragma unroll 4
  for (int k = threadIdx.x; k < limit; k += kThreads) dst[k] = staged[k];
}

struct PeakScratch {       // grow-only per-device scratch, one allocation:
  u64 *block_sums = nullptr;      //   [blocks] u64 | [groups] u32 | [tiles] u32
  size_t tiles_cap = 0;
};
PeakScratch g_peaks[16];
std::recursive_mutex g_peaks_mu;   // recursive: the host-facing call holds it across its two phases

// returns the number of extrema found; writes at most `capacity` of them (device memory)
// reuse_counts: the tile counts and sums of the previous call (same data, same type, lock held
// since) are still in the scratch -- go straight to the write pass
unsigned long long peaks_run(const float *d_data, i64 size, int type, int2 *d_out,
                             unsigned long long capacity, bool count_only, cudaStream_t stream,
                             bool reuse_counts = false) {
  int dev = 0;
  B200_CUDA_CHECK(cudaGetDevice(&dev));
  B200_REQUIRE(dev < 16, "device index out of range");
  const i64 tiles = (size + kTile - 1) / kTile;
  const bool vec = (((unsigned long long)d_data) & 15ull) == 0;
  B200_REQUIRE(tiles < 2147483647LL && size < 2147483647LL, "series too long for int positions");
  const int kind = type & 3;            // the reference tests these two bits only
  if (kind == 0) return 0;
  std::lock_guard<std::recursive_mutex> lock(g_peaks_mu);
  PeakScratch &sc = g_peaks[dev];
  if ((size_t)tiles > sc.tiles_cap) {
    if (sc.block_sums) {
      B200_CUDA_CHECK(cudaDeviceSynchronize());
      cudaFree(sc.block_sums);
    }
    const size_t cap = (size_t)tiles + (size_t)tiles / 4 + 64;
    B200_CUDA_CHECK(cudaMalloc(&sc.block_sums, sizeof(u64) * (cap / kBlockTiles + 2) +
                                                   sizeof(unsigned) * (cap / kGroupTiles + 2) +
                                                   sizeof(unsigned) * cap));
    sc.tiles_cap = cap;
  }
  const size_t blocks = (size_t)((tiles + kBlockTiles - 1) / kBlockTiles);
  u64 *block_sums = sc.block_sums;
  unsigned *group_sums = reinterpret_cast<unsigned *>(block_sums + sc.tiles_cap / kBlockTiles + 2);
  unsigned *counts = group_sums + sc.tiles_cap / kGroupTiles + 2;
  auto count_kernel = kind == 1 ? peaks_count_kernel<1> : kind == 2 ? peaks_count_kernel<2> : peaks_count_kernel<3>;
  auto write_kernel = kind == 1 ? peaks_write_kernel<1> : kind == 2 ? peaks_write_kernel<2> : peaks_write_kernel<3>;
  if (!reuse_counts) {
    // the tile counts are accumulated atomically: zero them (the sums are plainly written)
    B200_CUDA_CHECK(cudaMemsetAsync(counts, 0, sizeof(unsigned) * (size_t)tiles, stream));
    count_kernel<<<(unsigned)tiles, kThreads, 0, stream>>>(d_data, size, vec, counts);
    peaks_sums_kernel<<<(unsigned)blocks, kBlockTiles, 0, stream>>>(counts, tiles, group_sums, block_sums);
    b200_count_launch(2);
  }
  if (!count_only) {
    write_kernel<<<(unsigned)tiles, kThreads, 0, stream>>>(d_data, size, vec, counts, group_sums, block_sums,
                                                           d_out, capacity);
    b200_count_launch(1);
  }
  B200_CUDA_CHECK(cudaGetLastError());
  std::vector<u64> host(blocks);
  B200_CUDA_CHECK(cudaMemcpyAsync(host.data(), block_sums, sizeof(u64) * blocks, cudaMemcpyDeviceToHost, stream));
  B200_CUDA_CHECK(cudaStreamSynchronize(stream));
  unsigned long long total = 0;
  for (u64 v : host) total += v;
  return total;
}

bool peaks_is_device_pointer(const void *p) {
  cudaPointerAttributes attr;
  if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged;
}

}  // namespace

// C-ABI shims used by src/detect_peaks.c -------------------------------------------------------

extern "C" unsigned long long b200_detect_peaks_device(const float *d_data, long long size, int type,
                                                       void *d_results, unsigned long long capacity) {
  B200_REQUIRE(d_data && size > 2, "detect_peaks: null data or size <= 2");
  cudaStream_t stream = (cudaStream_t)b200_get_stream();
  return peaks_run(d_data, size, type, reinterpret_cast<int2 *>(d_results), capacity,
                   d_results == nullptr || capacity == 0, stream);
}

// Host-facing: data may be host or device; *results_host is malloc()ed (NULL when empty).
extern "C" void b200_detect_peaks(const float *data, long long size, int type, void **results_host,
                                  size_t *count) {
  B200_REQUIRE(data && results_host && count, "detect_peaks: null argument");
  B200_REQUIRE(size > 2, "detect_peaks: size must exceed 2");
  cudaStream_t stream = (cudaStream_t)b200_get_stream();
  *results_host = nullptr;
  *count = 0;
  const float *d_data = data;
  float *staged = nullptr;
  if (!peaks_is_device_pointer(data)) {
    B200_CUDA_CHECK(cudaMalloc(&staged, sizeof(float) * (size_t)size));
    B200_CUDA_CHECK(cudaMemcpyAsync(staged, data, sizeof(float) * (size_t)size, cudaMemcpyHostToDevice, stream));
    d_data = staged;
  }
  // pass 1: count; then an exact-size output; pass 2: the write pass alone, on the counts
  // pass 1 left in the scratch (the lock keeps other callers off it in between)
  std::lock_guard<std::recursive_mutex> lock(g_peaks_mu);
  unsigned long long n = peaks_run(d_data, size, type, nullptr, 0, true, stream);
  if (n > 0) {
    int2 *d_out = nullptr;
    B200_CUDA_CHECK(cudaMalloc(&d_out, sizeof(int2) * n));
    peaks_run(d_data, size, type, d_out, n, false, stream, true);
    void *host = malloc(sizeof(int2) * n);
    B200_REQUIRE(host, "detect_peaks: out of host memory");
    B200_CUDA_CHECK(cudaMemcpyAsync(host, d_out, sizeof(int2) * n, cudaMemcpyDeviceToHost, stream));
    B200_CUDA_CHECK(cudaStreamSynchronize(stream));
    B200_CUDA_CHECK(cudaFree(d_out));
    *results_host = host;
    *count = (size_t)n;
  }
  if (staged) B200_CUDA_CHECK(cudaFree(staged));
}
