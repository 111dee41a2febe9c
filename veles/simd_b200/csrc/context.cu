// context.cu -- device context, plans and the host<->device boundary of libSimd (B200).
//
// The reference API takes host float* (inc/simd/convolve.h:57-59,82-85,99-102,119-121).
// b200_run() classifies every pointer with cudaPointerGetAttributes: host pointers are
// staged through device buffers (and the call returns with the result in host memory),
// device / managed pointers are used in place and the work is left asynchronous on the
// library stream.  There is no CPU fallback: without a usable CUDA device every compute
// entry point aborts with a message.
#include "b200_device.cuh"
#include <atomic>
#include <mutex>
#include <string.h>
#include <vector>

namespace {

constexpr int kMaxDev = 16;

struct Staging {        // grow-only device staging for host-pointer callers (two slots)
  float *x[2] = {nullptr, nullptr}, *out[2] = {nullptr, nullptr}, *h = nullptr;
  size_t x_cap[2] = {0, 0}, out_cap[2] = {0, 0}, h_cap = 0;
};

struct Pipe {           // copy streams / events of the host-pointer pipeline
  bool ready = false;
  cudaStream_t s_in = nullptr, s_out = nullptr;
  cudaEvent_t ev_in[2], ev_comp[2], ev_out[2];
};

struct DeviceCtx {
  cudaStream_t own_stream = nullptr;   // created lazily, non-blocking
  Staging staging;                     // used by handle-less calls (convolve_simd, ...)
  Pipe pipe;
  float2 *spectrum = nullptr;          // scratch for handle-less FFT jobs
  size_t spectrum_cap = 0;
};

std::mutex g_mu;                 // guards contexts, the shared staging and user stream
DeviceCtx g_ctx[kMaxDev];
std::atomic<void *> g_user_stream[kMaxDev];       // per device: a stream belongs to one device
std::atomic<int> g_have_user_stream[kMaxDev];
std::atomic<long long> g_launches{0};

int current_device() {
  int dev = -1;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) {
    fprintf(stderr,
            "libSimd(b200): no usable CUDA device (%s). This library has no CPU fallback.\n",
            cudaGetErrorString(err));
    abort();
  }
  B200_REQUIRE(dev >= 0 && dev < kMaxDev, "device index out of range");
  return dev;
}

cudaStream_t current_stream() {
  int dev = current_device();
  if (g_have_user_stream[dev].load()) return (cudaStream_t)g_user_stream[dev].load();
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_ctx[dev].own_stream) {
    B200_CUDA_CHECK(cudaStreamCreateWithFlags(&g_ctx[dev].own_stream, cudaStreamNonBlocking));
  }
  return g_ctx[dev].own_stream;
}

bool is_device_pointer(const void *p) {
  cudaPointerAttributes attr;
  cudaError_t err = cudaPointerGetAttributes(&attr, p);
  if (err != cudaSuccess) {
    cudaGetLastError();   // clear; treat as plain host memory
    return false;
  }
  return attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged;
}

void grow(float **buf, size_t *cap, size_t need) {
  if (need <= *cap) return;
  if (*buf) B200_CUDA_CHECK(cudaFree(*buf));
  size_t n = need + need / 8 + 256;
  B200_CUDA_CHECK(cudaMalloc(buf, n * sizeof(float)));
  *cap = n;
}

}  // namespace

struct B200Plan {
  int magic;
  int algorithm;
  int device;
  i64 x_len, h_len;
  OsDevPlan os;
  Staging staging;
  std::mutex mu;       // one thread per handle at a time (reference: shared scratch)
};

void b200_count_launch(int n) { g_launches.fetch_add(n); }

// ---- optional per-kernel event timing -----------------------------------------------------
namespace {
std::atomic<int> g_prof_on{0};
std::mutex g_prof_mu;
struct ProfPair { cudaEvent_t e0, e1; int kind; };
std::vector<ProfPair> g_prof_pairs;
cudaEvent_t g_prof_open[B200_PROF_KINDS];
}  // namespace

// simd_b200_filter_noalloc promises to create no event: it switches profiling off for the
// calling thread around its launches.
static thread_local int g_prof_suppressed = 0;
void b200_prof_suppress(int on) { g_prof_suppressed = on; }

void b200_prof_begin(int kind, cudaStream_t stream) {
  if (!g_prof_on.load() || g_prof_suppressed) return;
  std::lock_guard<std::mutex> lock(g_prof_mu);
  B200_CUDA_CHECK(cudaEventCreate(&g_prof_open[kind]));
  B200_CUDA_CHECK(cudaEventRecord(g_prof_open[kind], stream));
}

void b200_prof_end(int kind, cudaStream_t stream) {
  if (!g_prof_on.load() || g_prof_suppressed) return;
  std::lock_guard<std::mutex> lock(g_prof_mu);
  ProfPair p;
  p.e0 = g_prof_open[kind];
  p.kind = kind;
  B200_CUDA_CHECK(cudaEventCreate(&p.e1));
  B200_CUDA_CHECK(cudaEventRecord(p.e1, stream));
  g_prof_pairs.push_back(p);
}

extern "C" void b200_profile_enable(int on) { g_prof_on.store(on); }

// Waits for the recorded events, adds each kind's elapsed time and launch count to
// total_ms[kind] / count[kind] (arrays of B200_PROF_KINDS) and clears the record.
extern "C" void b200_profile_collect(double *total_ms, long long *count) {
  std::lock_guard<std::mutex> lock(g_prof_mu);
  for (int k = 0; k < B200_PROF_KINDS; k++) { total_ms[k] = 0; count[k] = 0; }
  for (ProfPair &p : g_prof_pairs) {
    B200_CUDA_CHECK(cudaEventSynchronize(p.e1));
    float ms = 0;
    B200_CUDA_CHECK(cudaEventElapsedTime(&ms, p.e0, p.e1));
    total_ms[p.kind] += ms;
    count[p.kind] += 1;
    cudaEventDestroy(p.e0);
    cudaEventDestroy(p.e1);
  }
  g_prof_pairs.clear();
}

extern "C" long long b200_kernel_launches(void) { return g_launches.load(); }

extern "C" void b200_set_stream(void *s) {
  int dev = current_device();
  g_user_stream[dev].store(s);
  g_have_user_stream[dev].store(1);
}

extern "C" void *b200_get_stream(void) { return (void *)current_stream(); }

extern "C" void b200_synchronize(void) {
  B200_CUDA_CHECK(cudaStreamSynchronize(current_stream()));
}

extern "C" int b200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

extern "C" B200Plan *b200_plan_create(int algorithm, long long x_len, long long h_len,
                                      int n_fft, int real_packed) {
  B200Plan *plan = new B200Plan();
  plan->magic = 0xB200;
  plan->algorithm = algorithm;
  plan->device = current_device();
  plan->x_len = x_len;
  plan->h_len = h_len;
  B200_REQUIRE(n_fft >= B200_FFT_MIN && n_fft <= B200_FFT_MAX && (n_fft & (n_fft - 1)) == 0,
               "transform length must be a power of two in range");
  plan->os.n_fft = n_fft;
  // whole-signal transform (FFT algorithm, x+h-1 fits): one partition holds all taps
  const bool whole = algorithm == B200_ALG_FFT && x_len + h_len - 1 <= n_fft;
  plan->os.part_taps = whole ? (int)h_len : b200_os_part_taps(n_fft, h_len);
  plan->os.max_parts = (int)((h_len + plan->os.part_taps - 1) / plan->os.part_taps);
  plan->os.real_packed = real_packed != 0 && !whole && h_len <= n_fft;
  if (plan->os.real_packed) { plan->os.part_taps = (int)h_len; plan->os.max_parts = 1; }
  plan->os.twiddles = b200_twiddles(n_fft);
  // paired mode: float2 per bin and partition; real-packed mode: float4 {alpha, beta} per bin
  B200_CUDA_CHECK(cudaMalloc(&plan->os.spectrum,
                             sizeof(float2) * (size_t)n_fft * plan->os.max_parts *
                                 (plan->os.real_packed ? 2 : 1)));
  // staging for host-pointer callers is sized now so the call path does not allocate
  {
    const size_t cap = (size_t)16 << 20;     // one pipeline piece at most (see b200_run)
    size_t xs = (size_t)x_len, os = (size_t)(x_len + h_len - 1);
    grow(&plan->staging.x[0], &plan->staging.x_cap[0], xs < cap ? xs : cap);
    grow(&plan->staging.out[0], &plan->staging.out_cap[0], os < cap ? os : cap);
    grow(&plan->staging.h, &plan->staging.h_cap, (size_t)h_len);
  }
  return plan;
}

extern "C" void b200_plan_destroy(B200Plan *plan) {
  if (!plan) return;
  B200_REQUIRE(plan->magic == 0xB200, "not a libSimd(b200) plan");
  int prev = current_device();
  if (prev != plan->device) B200_CUDA_CHECK(cudaSetDevice(plan->device));
  // outstanding asynchronous work may still read the spectrum / staging buffers
  B200_CUDA_CHECK(cudaDeviceSynchronize());
  cudaFree(plan->os.spectrum);
  for (int sl = 0; sl < 2; sl++) {
    cudaFree(plan->staging.x[sl]);
    cudaFree(plan->staging.out[sl]);
  }
  cudaFree(plan->staging.h);
  if (prev != plan->device) B200_CUDA_CHECK(cudaSetDevice(prev));
  plan->magic = 0;
  delete plan;
}

extern "C" float *b200_plan_spectrum(const B200Plan *plan) { return (float *)plan->os.spectrum; }
extern "C" float *b200_plan_scratch(const B200Plan *plan) { return plan->staging.x[0]; }

namespace {

void launch_device_job(B200Plan *plan, int algorithm, const DevJob &dj, cudaStream_t stream) {
  if (algorithm == B200_ALG_DIRECT) {
    b200_direct_launch(dj, stream);
    return;
  }
  if (plan) {
    bool single = algorithm == B200_ALG_FFT && plan->os.max_parts == 1 &&
                  dj.pad == dj.h_len - 1 && dj.x_len + dj.h_len - 1 <= plan->os.n_fft &&
                  dj.out_len <= plan->os.n_fft;
    b200_os_launch(plan->os, dj, single, stream);
    return;
  }
  // handle-less FFT job (extension API with algorithm forced): context-owned scratch
  int dev = current_device();
  int n = 0, real_packed = 0;
  b200_os_pick_mode(dj.h_len, dj.out_len, &n, &real_packed);
  int taps = b200_os_part_taps(n, dj.h_len);
  int parts = real_packed ? 2 : (int)((dj.h_len + taps - 1) / taps);   // float4 = 2 float2 per bin
  const float2 *tw = b200_twiddles(n);
  std::lock_guard<std::mutex> lock(g_mu);
  DeviceCtx &c = g_ctx[dev];
  size_t need = (size_t)n * parts;
  if (need > c.spectrum_cap) {
    if (c.spectrum) { B200_CUDA_CHECK(cudaDeviceSynchronize()); B200_CUDA_CHECK(cudaFree(c.spectrum)); }
    B200_CUDA_CHECK(cudaMalloc(&c.spectrum, sizeof(float2) * need));
    c.spectrum_cap = need;
  }
  if (real_packed) b200_os_launch_real(tw, reinterpret_cast<float4 *>(c.spectrum), n, dj, stream);
  else b200_os_launch_with(tw, c.spectrum, n, taps, dj, false, stream);
}

}  // namespace

extern "C" void b200_run(B200Plan *plan, int algorithm, const B200Job *job) {
  B200_REQUIRE(job->x && job->h && job->out, "null pointer argument");
  B200_REQUIRE(job->x_len > 0 && job->h_len > 0, "zero-length signal");
  if (plan) B200_REQUIRE(plan->magic == 0xB200, "not a libSimd(b200) plan (stale handle?)");
  if (job->batch <= 0 || job->out_len <= 0) return;
  int dev = current_device();
  if (plan) B200_REQUIRE(plan->device == dev, "handle used on a different device than it was created on");
  cudaStream_t stream = current_stream();

  const bool xd = is_device_pointer(job->x);
  const bool hd = is_device_pointer(job->h);
  const bool od = is_device_pointer(job->out);

  DevJob dj;
  dj.x = job->x; dj.x_len = job->x_len; dj.x_stride = job->x_stride;
  dj.h = job->h; dj.h_len = job->h_len;
  dj.out = job->out; dj.out_len = job->out_len; dj.out_stride = job->out_stride;
  dj.pad = job->pad; dj.batch = job->batch; dj.reverse = job->reverse;

  if (xd && hd && od) {
    if (plan) {
      std::lock_guard<std::mutex> lock(plan->mu);
      launch_device_job(plan, algorithm, dj, stream);
    } else {
      launch_device_job(plan, algorithm, dj, stream);
    }
    return;
  }

  // ---- host-pointer path: stage, run, copy back, synchronise --------------------------
  // The job is cut into pieces (groups of whole signals, or -- for one long signal -- time
  // slices with their (h-1)-sample halo) that flow through two staging slots on three
  // streams, so the H2D copy of piece i+1, the kernels of piece i and the D2H copy of piece
  // i-1 overlap (PCIe is full duplex).
  std::unique_lock<std::mutex> plan_lock, ctx_lock;
  Staging *st;
  if (plan) {
    plan_lock = std::unique_lock<std::mutex>(plan->mu);
    st = &plan->staging;
  } else {
    // handle-less calls share the per-device staging: serialise them
    static std::mutex staged_call_mu;
    ctx_lock = std::unique_lock<std::mutex>(staged_call_mu);
    st = &g_ctx[dev].staging;
  }
  // the copy streams and events are one set per device: host-pointer calls of different
  // handles take turns (they would share the PCIe link anyway)
  static std::mutex pipe_mu[kMaxDev];
  std::lock_guard<std::mutex> pipe_lock(pipe_mu[dev]);
  Pipe &pipe = g_ctx[dev].pipe;
  {
    std::lock_guard<std::mutex> lock(g_mu);
    if (!pipe.ready) {
      B200_CUDA_CHECK(cudaStreamCreateWithFlags(&pipe.s_in, cudaStreamNonBlocking));
      B200_CUDA_CHECK(cudaStreamCreateWithFlags(&pipe.s_out, cudaStreamNonBlocking));
      for (int i = 0; i < 2; i++) {
        B200_CUDA_CHECK(cudaEventCreateWithFlags(&pipe.ev_in[i], cudaEventDisableTiming));
        B200_CUDA_CHECK(cudaEventCreateWithFlags(&pipe.ev_comp[i], cudaEventDisableTiming));
        B200_CUDA_CHECK(cudaEventCreateWithFlags(&pipe.ev_out[i], cudaEventDisableTiming));
      }
      pipe.ready = true;
    }
  }

  const i64 x_row = job->x_len, o_row = job->out_len;
  const i64 budget = (i64)16 << 20;              // floats per staging buffer (64 MiB)
  const bool slice_time = job->batch == 1 && o_row > budget;
  i64 rows_per_piece = 1, outs_per_piece = o_row;
  if (slice_time) {
    outs_per_piece = budget / 4;                            // 16 MiB slices of one long signal
  } else {
    rows_per_piece = budget / (x_row > o_row ? x_row : o_row);
    if (rows_per_piece < 1) rows_per_piece = 1;
    if (rows_per_piece > job->batch) rows_per_piece = job->batch;
    // Aim for ~32 pieces of at least ~8 MiB: the pipeline's fill (first H2D) and drain (last
    // D2H) are not overlapped, so they should be a small fraction of the whole.
    {
      const i64 row_floats = x_row > o_row ? x_row : o_row;
      i64 min_rows = ((i64)2 << 20) / row_floats;           // >= 8 MiB per piece
      if (min_rows < 1) min_rows = 1;
      i64 want = job->batch / 32;
      if (want < min_rows) want = min_rows;
      if (rows_per_piece > want) rows_per_piece = want;
      if (rows_per_piece < 1) rows_per_piece = 1;
    }
  }
  const i64 piece_x = slice_time ? outs_per_piece + job->h_len : rows_per_piece * x_row;
  const i64 piece_o = slice_time ? outs_per_piece : rows_per_piece * o_row;
  const i64 n_pieces = slice_time ? (o_row + outs_per_piece - 1) / outs_per_piece
                                  : (job->batch + rows_per_piece - 1) / rows_per_piece;
  for (int sl = 0; sl < (n_pieces > 1 ? 2 : 1); sl++) {
    if (!xd) grow(&st->x[sl], &st->x_cap[sl], (size_t)piece_x);
    if (!od) grow(&st->out[sl], &st->out_cap[sl], (size_t)piece_o);
  }
  if (!hd) {
    grow(&st->h, &st->h_cap, (size_t)job->h_len);
    B200_CUDA_CHECK(cudaMemcpyAsync(st->h, job->h, sizeof(float) * job->h_len,
                                    cudaMemcpyHostToDevice, stream));
    dj.h = st->h;
  }

  for (i64 pi = 0; pi < n_pieces; pi++) {
    const int sl = (int)(pi & 1);
    // --- describe the piece ---------------------------------------------------------------
    const float *src_x; float *dst_o;
    i64 px_len, po_len, p_rows, p_pad;
    if (slice_time) {
      const i64 o0 = pi * outs_per_piece;
      const i64 o1 = o0 + outs_per_piece < o_row ? o0 + outs_per_piece : o_row;
      i64 lo = o0 - job->pad; if (lo < 0) lo = 0;
      i64 hi = o1 - job->pad + job->h_len - 1; if (hi > x_row) hi = x_row;
      if (hi < lo) hi = lo;
      src_x = job->x + lo; px_len = hi - lo; p_pad = job->pad - o0 + lo;
      dst_o = job->out + o0; po_len = o1 - o0; p_rows = 1;
    } else {
      const i64 b0 = pi * rows_per_piece;
      p_rows = (job->batch - b0) < rows_per_piece ? (job->batch - b0) : rows_per_piece;
      src_x = job->x + b0 * job->x_stride; px_len = x_row; p_pad = job->pad;
      dst_o = job->out + b0 * job->out_stride; po_len = o_row;
    }
    dj.batch = p_rows; dj.x_len = px_len; dj.out_len = po_len; dj.pad = p_pad;
    // --- H2D ------------------------------------------------------------------------------
    if (!xd) {
      if (pi >= 2) B200_CUDA_CHECK(cudaStreamWaitEvent(pipe.s_in, pipe.ev_comp[sl], 0));
      if (px_len > 0) {
        if (p_rows == 1 || job->x_stride == px_len) {   // contiguous: no 2-D pitch limit
          B200_CUDA_CHECK(cudaMemcpyAsync(st->x[sl], src_x, sizeof(float) * px_len * p_rows,
                                          cudaMemcpyHostToDevice, pipe.s_in));
        } else {
          B200_CUDA_CHECK(cudaMemcpy2DAsync(st->x[sl], sizeof(float) * px_len, src_x,
                                            sizeof(float) * job->x_stride, sizeof(float) * px_len,
                                            (size_t)p_rows, cudaMemcpyHostToDevice, pipe.s_in));
        }
      }
      B200_CUDA_CHECK(cudaEventRecord(pipe.ev_in[sl], pipe.s_in));
      B200_CUDA_CHECK(cudaStreamWaitEvent(stream, pipe.ev_in[sl], 0));
      dj.x = st->x[sl]; dj.x_stride = px_len;
    } else {
      dj.x = src_x; dj.x_stride = job->x_stride;
    }
    // --- kernels --------------------------------------------------------------------------
    if (!od) {
      if (pi >= 2) B200_CUDA_CHECK(cudaStreamWaitEvent(stream, pipe.ev_out[sl], 0));
      dj.out = st->out[sl]; dj.out_stride = po_len;
    } else {
      dj.out = dst_o; dj.out_stride = job->out_stride;
    }
    if (px_len > 0) {
      launch_device_job(plan, algorithm, dj, stream);
    } else {
      B200_CUDA_CHECK(cudaMemsetAsync(dj.out, 0, sizeof(float) * po_len, stream));
    }
    B200_CUDA_CHECK(cudaEventRecord(pipe.ev_comp[sl], stream));
    // --- D2H ------------------------------------------------------------------------------
    if (!od) {
      B200_CUDA_CHECK(cudaStreamWaitEvent(pipe.s_out, pipe.ev_comp[sl], 0));
      if (p_rows == 1 || job->out_stride == po_len) {
        B200_CUDA_CHECK(cudaMemcpyAsync(dst_o, st->out[sl], sizeof(float) * po_len * p_rows,
                                        cudaMemcpyDeviceToHost, pipe.s_out));
      } else {
        B200_CUDA_CHECK(cudaMemcpy2DAsync(dst_o, sizeof(float) * job->out_stride, st->out[sl],
                                          sizeof(float) * po_len, sizeof(float) * po_len,
                                          (size_t)p_rows, cudaMemcpyDeviceToHost, pipe.s_out));
      }
      B200_CUDA_CHECK(cudaEventRecord(pipe.ev_out[sl], pipe.s_out));
    }
  }
  if (!od) B200_CUDA_CHECK(cudaStreamSynchronize(pipe.s_out));
  B200_CUDA_CHECK(cudaStreamSynchronize(stream));
}
