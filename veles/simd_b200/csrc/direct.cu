// direct.cu -- the "brute force" path on the GPU: shared-memory tiled, register-blocked
// direct convolution / cross-correlation for short impulse responses.
//
// Replaces the AVX loops of the reference: convolve_simd (/root/reference/src/convolve.c:
// 40-101) and cross_correlate_simd (/root/reference/src/correlate.c:74-126).  Both are the
// same correlation-form sum  out[n] = sum_j g[j] * X(n + j - pad)  with g = reversed h for
// convolution and g = h for correlation, so one kernel serves both; the reference's
// per-output beg/end clamps (convolve.c:51-55, correlate.c:80-84) become a zero-filled halo.
//
// Layout: a CTA owns a strip of consecutive TILE = R*TPB output tiles of one signal.  For
// each tile the x window [n0 - pad, n0 - pad + TILE + taps - 1) is staged in shared memory
// with cp.async (LDGSTS, zero-filled outside the signal) into one of two buffers, so the
// next tile's loads are in flight while the current tile is computed.  Every thread
// accumulates R = 16 consecutive outputs in registers, sliding a 32-sample register window
// over the tile so that each 16-byte shared-memory read feeds 64 FMAs.  The x tile is stored
// with 4 pad floats per 16 so that the 64-byte-strided float4 reads of a quarter warp hit 8
// distinct bank groups.  Results go back through shared memory so the global stores are
// fully coalesced for any row alignment.
//
// Impulse responses longer than kMaxHC taps (not what the selector sends here, but legal
// through convolve_simd) take the unpipelined multi-pass kernel.
//
// Roofline: 2*h_len FLOP and 8 B of HBM traffic per output -> FP32-bound above h_len ~ 45.
#include "b200_device.cuh"

namespace {

constexpr int kR = 16;          // outputs per thread
constexpr int kTPB = 128;       // threads per CTA
constexpr int kTile = kR * kTPB;
constexpr int kMaxHC = 1024;    // taps staged per pass over the tile
constexpr int kMaxStrip = 16;   // tiles per CTA in the pipelined kernel

__device__ __forceinline__ int padx(int q) { return q + ((q >> 4) << 2); }

struct DirectParams {
  const float *x;
  i64 x_len, x_stride;
  const float *h;
  i64 h_len;
  float *out;
  i64 out_len, out_stride;
  i64 pad;
  i64 tiles_per_signal;
  i64 strips_per_signal;
  int strip;            // tiles per CTA
  int reverse;
  int hc_max;           // taps per staging pass (multiple of 16)
  int xbuf_floats;      // size of one padded x buffer
  int lead;             // pipelined kernel: tile t covers outputs [t*kTile - lead, ...): chosen so that
                        // every tile's x window starts on a 16-byte boundary (0..3)
  int wide;             // ... and is then staged with 16-byte cp.async
};

__device__ __forceinline__ void cp_async4(float *smem_dst, const float *gsrc, bool valid) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  int sz = valid ? 4 : 0;      // src-size 0 -> the 4 destination bytes are zero-filled
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(d), "l"(gsrc), "r"(sz));
}
// 16 bytes; `bytes` of them (0, 4, 8, 12 or 16) come from global memory, the rest is zero-filled
__device__ __forceinline__ void cp_async16(float *smem_dst, const float *gsrc, int bytes) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(gsrc), "r"(bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }

// 16 taps starting at tap j0 against the 32-sample window [lo | hi] (lo = samples j0 .. j0+15
// of the thread's row, hi = the next 16): 256 FMAs
__device__ __forceinline__ void taps16(float (&acc)[kR], const float *gs, int j0, const float (&lo)[16],
                                       const float (&hi)[16]) {
#pragma unroll
  for (int q4 = 0; q4 < 4; q4++) {
    float4 g = *reinterpret_cast<const float4 *>(gs + j0 + 4 * q4);
    const float gg[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
    for (int jj = 0; jj < 4; jj++) {
#pragma unroll
      for (int r = 0; r < kR; r++) {
        const int i = r + 4 * q4 + jj;                   // compile-time: which half of the window
        acc[r] = fmaf(gg[jj], i < 16 ? lo[i] : hi[i - 16], acc[r]);
      }
    }
  }
}
__device__ __forceinline__ void load16(float (&w)[16], const float *src) {
#pragma unroll
  for (int i = 0; i < 4; i++) {
    float4 v = *reinterpret_cast<const float4 *>(src + 4 * i);
    w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
  }
}

// acc[r] += sum_{j < hc} gs[j] * xs[t0 + r + j], r < 16; xrow = xs + padx(t0)
// The sliding 32-sample window lives in two 16-register halves that swap roles from one block
// of 16 taps to the next (two blocks per loop trip), so it never has to be shifted.
__device__ __forceinline__ void accumulate_taps(float (&acc)[kR], const float *gs,
                                                const float *xrow, int hc) {
  const int hc16 = hc & ~15;
  float wa[16], wb[16];
  if (hc16 > 0) load16(wa, xrow);
  int j0 = 0;
  for (; j0 + 32 <= hc16; j0 += 32) {
    load16(wb, xrow + ((j0 + 16) >> 4) * 20);
    taps16(acc, gs, j0, wa, wb);
    load16(wa, xrow + ((j0 + 32) >> 4) * 20);
    taps16(acc, gs, j0 + 16, wb, wa);
  }
  if (j0 < hc16) {                                       // an odd block of 16 taps
    load16(wb, xrow + ((j0 + 16) >> 4) * 20);
    taps16(acc, gs, j0, wa, wb);
  }
  // remaining (< 16) taps, four at a time with exact tap predicates
  for (int j0 = hc16; j0 < hc; j0 += 4) {
    const int blk = j0 >> 4, off = j0 & 15;    // off in {0,4,8,12}
    const float *xa = xrow + blk * 20 + off;
    float wr[20];
#pragma unroll
    for (int i = 0; i < 5; i++) {
      // the 20-sample window may straddle one row boundary (rows are 16 data + 4 pad)
      const float *src = xa + 4 * i + (off + 4 * i >= 16 ? 4 : 0);
      float4 v = *reinterpret_cast<const float4 *>(src);
      wr[4 * i] = v.x; wr[4 * i + 1] = v.y; wr[4 * i + 2] = v.z; wr[4 * i + 3] = v.w;
    }
    float4 g = *reinterpret_cast<const float4 *>(gs + j0);
    const float gg[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
    for (int jj = 0; jj < 4; jj++) {
      if (j0 + jj < hc) {
#pragma unroll
        for (int r = 0; r < kR; r++) acc[r] = fmaf(gg[jj], wr[r + jj], acc[r]);
      }
    }
  }
}

// ---- pipelined kernel: all taps fit one staging pass (h_len <= kMaxHC) -------------------
__global__ void __launch_bounds__(kTPB, 6) direct_kernel(DirectParams p) {
  extern __shared__ float4 smem4[];
  float *gs = reinterpret_cast<float *>(smem4);          // hc_max + 16 taps
  float *xbuf0 = gs + p.hc_max + 16;                     // two padded x windows
  float *obuf = xbuf0 + 2 * p.xbuf_floats;               // padded output tile

  const int tid = threadIdx.x;
  const i64 b = blockIdx.x / p.strips_per_signal;
  const i64 tile0 = (blockIdx.x % p.strips_per_signal) * p.strip;
  const i64 tile_end = tile0 + p.strip < p.tiles_per_signal ? tile0 + p.strip : p.tiles_per_signal;
  const float *xb = p.x + b * p.x_stride;
  float *ob = p.out + b * p.out_stride;
  const int hc = (int)p.h_len;
  const int wlen = kTile + hc - 1;
  const int t0 = tid * kR;

  auto stage = [&](i64 tile, float *xs) {
    const i64 xi0 = tile * kTile - p.lead - p.pad;
    if (p.wide) {
      // xi0 is a multiple of 4 and the row starts on a 16-byte boundary: whole 16-byte units,
      // four samples each, never straddling a 16-sample padded row; a unit is inside the signal,
      // outside it (zero-filled), or -- at the signal's end only -- partly inside
      // The samples that exist are window positions [lo, hi) (clamped to the window; lo is a
      // multiple of 4 like xi0): one 64-bit computation per tile, 32-bit tests per unit.
      const int span = 4 * ((wlen + 3) >> 2);
      const i64 lo64 = -xi0, hi64 = p.x_len - xi0;
      const int lo = lo64 <= 0 ? 0 : (lo64 < span ? (int)lo64 : span);
      const int hi = hi64 >= span ? span : (hi64 > 0 ? (int)hi64 : 0);
      // thread's unit q = 4*tid, then every 4*kTPB samples = 32 padded rows = 5*kTPB floats on
      const i64 off = xi0 + 4 * tid;
      float *d = xs + padx(4 * tid);
      if (lo == 0 && hi == span) {
        const float *s = xb + off;
        for (int q = 4 * tid; q < span; q += 4 * kTPB, s += 4 * kTPB, d += 5 * kTPB) cp_async16(d, s, 16);
      } else {
        for (int q = 4 * tid; q < span; q += 4 * kTPB, d += 5 * kTPB) {
          int n = hi - q;
          n = q < lo || n <= 0 ? 0 : (n < 4 ? n : 4);
          cp_async16(d, n ? xb + off + (q - 4 * tid) : xb, 4 * n);
        }
      }
      cp_async_commit();
      return;
    }
    const float *s = xb + xi0 + tid;
    // q advances by kTPB = 128 (a multiple of the 16-float row), so the padded address
    // advances by a constant 128 + 32 floats: no per-element index arithmetic
    float *d = xs + padx(tid);
    const int iters = (wlen - tid + kTPB - 1) / kTPB;
    if (xi0 >= 0 && xi0 + wlen <= p.x_len) {
#pragma unroll 4
      for (int k = 0; k < iters; k++, d += kTPB + kTPB / 4, s += kTPB) cp_async4(d, s, true);
    } else {
      i64 i = xi0 + tid;
#pragma unroll 2
      for (int k = 0; k < iters; k++, d += kTPB + kTPB / 4, s += kTPB, i += kTPB) {
        const bool ok = i >= 0 && i < p.x_len;
        cp_async4(d, ok ? s : xb, ok);
      }
    }
    cp_async_commit();
  };

  stage(tile0, xbuf0);
  for (int j = tid; j < hc; j += kTPB) gs[j] = p.reverse ? p.h[j] : p.h[p.h_len - 1 - j];

  int buf = 0;
  for (i64 tile = tile0; tile < tile_end; tile++, buf ^= 1) {
    float *xs = xbuf0 + buf * p.xbuf_floats;
    cp_async_wait_all();
    __syncthreads();                       // tile's window (and the taps) visible to all;
                                           // the other buffer is no longer being read
    if (tile + 1 < tile_end) stage(tile + 1, xbuf0 + (buf ^ 1) * p.xbuf_floats);

    const i64 n0 = tile * kTile - p.lead;              // first output of the tile (tile 0: may be < 0)
    const i64 remaining = p.out_len - n0;
    const int cnt = remaining < kTile ? (int)remaining : kTile;
    const bool active = ((tid & ~31) * kR) < cnt;    // warp-uniform
    float acc[kR];
#pragma unroll
    for (int r = 0; r < kR; r++) acc[r] = 0.f;
    if (active) accumulate_taps(acc, gs, xs + padx(t0), hc);

    // results -> their own shared region (padded rows; last read before this iteration's
    // top barrier) -> coalesced global stores
    if (active) {
      float *orow = obuf + padx(t0);
#pragma unroll
      for (int i = 0; i < 4; i++) {
        *reinterpret_cast<float4 *>(orow + 4 * i) =
            make_float4(acc[4 * i], acc[4 * i + 1], acc[4 * i + 2], acc[4 * i + 3]);
      }
    }
    // Every warp wrote the 512 consecutive outputs of its own 32 threads and copies exactly
    // those out again: a warp-level barrier is enough, the CTA keeps a single barrier per tile.
    __syncwarp();
    const int wbase = (tid & ~31) * kR + (tid & 31);       // this lane's first output of the warp's 512
    float *o = ob + n0 + wbase;
    const float *sr = obuf + padx(wbase);
    if (n0 >= 0 && cnt == kTile) {
      // a whole tile: 16 copies per thread at compile-time offsets (32 outputs = 2 padded rows apart)
#pragma unroll
      for (int k = 0; k < kR; k++) o[k * 32] = sr[k * 40];
    } else {
      // last tile of a signal, or tile 0 of a signal whose tiles were shifted for alignment (its
      // first `lead` outputs do not exist)
      for (int k = 0; k < kR; k++) {
        const i64 n = n0 + wbase + k * 32;
        if (n >= 0 && wbase + k * 32 < cnt) o[k * 32] = sr[k * 40];
      }
    }
  }
}

// ---- small problems: fewer outputs per thread, more CTAs ---------------------------------------
// A single short signal gives the pipelined kernel only a handful of 2048-output tiles: a few
// SMs work through 16 outputs x h_len taps per thread while the rest idle, and the call is
// bound by that serial chain.  Here a CTA owns kSmallTile = 512 outputs, four consecutive ones
// per thread: four outputs and four taps touch seven consecutive window words, so one 16-byte
// shared-memory read feeds 16 FMAs from a sliding 8-word register window; one staging pass,
// no pipeline.  Same summation order per output as direct_kernel (taps ascending).
constexpr int kSmallR = 4;
constexpr int kSmallTile = kSmallR * kTPB;

__global__ void __launch_bounds__(kTPB) direct_small_kernel(DirectParams p) {
  extern __shared__ float4 smem4[];
  const int hc = (int)p.h_len;
  const int hc4 = (hc + 3) & ~3;
  float *gs = reinterpret_cast<float *>(smem4);          // hc4 taps (zero beyond hc: never read)
  float *xs = gs + hc4;                                  // kSmallTile + hc4 + 4 window words

  const int tid = threadIdx.x;
  const i64 b = blockIdx.x / p.tiles_per_signal;
  const i64 n0 = (blockIdx.x % p.tiles_per_signal) * kSmallTile;
  const float *xb = p.x + b * p.x_stride;
  const i64 xi0 = n0 - p.pad;
  const int wlen = kSmallTile + hc4 + 4;
  for (int q = tid; q < wlen; q += kTPB) {
    const i64 i = xi0 + q;
    xs[q] = (i >= 0 && i < p.x_len) ? xb[i] : 0.f;
  }
  for (int j = tid; j < hc4; j += kTPB) gs[j] = j < hc ? (p.reverse ? p.h[j] : p.h[p.h_len - 1 - j]) : 0.f;
  __syncthreads();

  const float *w0 = xs + kSmallR * tid;
  float acc[kSmallR], w[8];
#pragma unroll
  for (int r = 0; r < kSmallR; r++) acc[r] = 0.f;
  {
    const float4 v = *reinterpret_cast<const float4 *>(w0);
    w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
  }
  for (int j0 = 0; j0 < hc; j0 += 4) {
    const float4 v = *reinterpret_cast<const float4 *>(w0 + j0 + 4);
    w[4] = v.x; w[5] = v.y; w[6] = v.z; w[7] = v.w;
    const float4 g = *reinterpret_cast<const float4 *>(gs + j0);
    const float gg[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
    for (int jj = 0; jj < 4; jj++) {
      if (j0 + jj < hc) {                                // warp-uniform: exact tap count
#pragma unroll
        for (int r = 0; r < kSmallR; r++) acc[r] = fmaf(gg[jj], w[r + jj], acc[r]);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; i++) w[i] = w[i + 4];
  }
  float *ob = p.out + b * p.out_stride;
  const i64 n = n0 + kSmallR * tid;
  if (n + kSmallR <= p.out_len && ((reinterpret_cast<unsigned long long>(ob + n) & 15ull) == 0)) {
    *reinterpret_cast<float4 *>(ob + n) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  } else {
#pragma unroll
    for (int r = 0; r < kSmallR; r++) {
      if (n + r < p.out_len) ob[n + r] = acc[r];
    }
  }
}

// ---- unpipelined kernel for very long impulse responses (several staging passes) ---------
__global__ void __launch_bounds__(kTPB, 4) direct_long_kernel(DirectParams p) {
  extern __shared__ float4 smem4[];
  float *gs = reinterpret_cast<float *>(smem4);
  float *xs = gs + p.hc_max + 16;

  const int tid = threadIdx.x;
  const i64 tile = blockIdx.x;
  const i64 b = tile / p.tiles_per_signal;
  const i64 n0 = (tile % p.tiles_per_signal) * kTile;
  const float *xb = p.x + b * p.x_stride;
  const i64 remaining = p.out_len - n0;
  const int cnt = remaining < kTile ? (int)remaining : kTile;
  const int t0 = tid * kR;
  const bool active = ((tid & ~31) * kR) < cnt;

  float acc[kR];
#pragma unroll
  for (int r = 0; r < kR; r++) acc[r] = 0.f;

  for (i64 c0 = 0; c0 < p.h_len; c0 += p.hc_max) {
    const int hc = (int)((p.h_len - c0) < p.hc_max ? (p.h_len - c0) : p.hc_max);
    if (c0 > 0) __syncthreads();
    for (int j = tid; j < hc; j += kTPB) {
      i64 gj = c0 + j;
      gs[j] = p.reverse ? p.h[gj] : p.h[p.h_len - 1 - gj];
    }
    const i64 xi0 = n0 + c0 - p.pad;
    const int wlen = kTile + hc - 1;
#pragma unroll 4
    for (int q = tid; q < wlen; q += kTPB) {
      i64 i = xi0 + q;
      xs[padx(q)] = (i >= 0 && i < p.x_len) ? xb[i] : 0.f;
    }
    __syncthreads();
    if (active) accumulate_taps(acc, gs, xs + padx(t0), hc);
  }

  __syncthreads();
  if (active) {
    float *orow = xs + padx(t0);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      *reinterpret_cast<float4 *>(orow + 4 * i) =
          make_float4(acc[4 * i], acc[4 * i + 1], acc[4 * i + 2], acc[4 * i + 3]);
    }
  }
  __syncthreads();
  float *ob = p.out + b * p.out_stride + n0;
#pragma unroll 4
  for (int i = tid; i < cnt; i += kTPB) ob[i] = xs[padx(i)];
}

}  // namespace

void b200_direct_launch(const DevJob &job, cudaStream_t stream) {
  if (job.batch <= 0 || job.out_len <= 0) return;
  DirectParams p;
  p.x = job.x; p.x_len = job.x_len; p.x_stride = job.x_stride;
  p.h = job.h; p.h_len = job.h_len;
  p.out = job.out; p.out_len = job.out_len; p.out_stride = job.out_stride;
  p.pad = job.pad; p.reverse = job.reverse;
  p.lead = 0; p.wide = 0;
  p.tiles_per_signal = (job.out_len + kTile - 1) / kTile;
  i64 hc = job.h_len < kMaxHC ? job.h_len : kMaxHC;
  p.hc_max = (int)((hc + 15) & ~15LL);
  // padded x window (+ two slack rows for the float4 tail reads)
  p.xbuf_floats = ((kTile + p.hc_max + 32) / 16 + 1) * 20;
  const i64 tiles = p.tiles_per_signal * job.batch;
  b200_prof_begin(B200_PROF_DIRECT, stream);
  if (job.h_len <= kMaxHC && tiles < 148) {
    // less than one big tile per SM: spread the work over more, smaller CTAs
    p.strip = 1;
    p.tiles_per_signal = (job.out_len + kSmallTile - 1) / kSmallTile;
    p.strips_per_signal = p.tiles_per_signal;
    const i64 blocks = p.tiles_per_signal * job.batch;
    const int hc4 = ((int)job.h_len + 3) & ~3;
    size_t smem = sizeof(float) * (size_t)(hc4 + kSmallTile + hc4 + 4);
    direct_small_kernel<<<(unsigned)blocks, kTPB, smem, stream>>>(p);
  } else if (job.h_len <= kMaxHC) {
    // strips long enough to amortise the pipeline prologue, short enough to fill the GPU
    i64 strip = tiles / (148 * 10);
    if (strip < 1) strip = 1;
    if (strip > kMaxStrip) strip = kMaxStrip;
    if (strip > p.tiles_per_signal) strip = p.tiles_per_signal;
    // shift the tiles so that each one's x window starts 16-byte aligned in global memory; then
    // it is staged in 16-byte units (needs rows that start on 16-byte boundaries themselves)
    p.wide = ((((unsigned long long)job.x) & 15ull) == 0 && (job.x_stride & 3) == 0 &&
              !getenv("VELES_B200_DIRECT_NARROW")) ? 1 : 0;
    p.lead = p.wide ? (int)(((-job.pad) % 4 + 4) % 4) : 0;
    p.tiles_per_signal = (job.out_len + p.lead + kTile - 1) / kTile;
    if (strip > p.tiles_per_signal) strip = p.tiles_per_signal;
    p.strip = (int)strip;
    p.strips_per_signal = (p.tiles_per_signal + strip - 1) / strip;
    const i64 blocks = p.strips_per_signal * job.batch;
    B200_REQUIRE(blocks < 2147483647LL, "direct kernel grid too large");
    size_t smem = sizeof(float) * (size_t)(p.hc_max + 16 + 2 * p.xbuf_floats + (kTile / 16 + 1) * 20);
    direct_kernel<<<(unsigned)blocks, kTPB, smem, stream>>>(p);
  } else {
    p.strip = 1;
    p.strips_per_signal = p.tiles_per_signal;
    B200_REQUIRE(tiles < 2147483647LL, "direct kernel grid too large");
    size_t smem = sizeof(float) * (size_t)(p.hc_max + 16 + p.xbuf_floats);
    direct_long_kernel<<<(unsigned)tiles, kTPB, smem, stream>>>(p);
  }
  b200_prof_end(B200_PROF_DIRECT, stream);
  B200_CUDA_CHECK(cudaGetLastError());
  b200_count_launch(1);
}
