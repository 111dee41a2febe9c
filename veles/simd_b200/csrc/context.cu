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
#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
#include <string.h>
#include <thread>
#include <vector>

namespace {

constexpr int kMaxDev = 16;

struct Staging {        // grow-only device staging for host-pointer callers (two slots)
  float *x[2] = {nullptr, nullptr}, *out[2] = {nullptr, nullptr}, *h = nullptr;
  size_t x_cap[2] = {0, 0}, out_cap[2] = {0, 0}, h_cap = 0;
  float *x16[2] = {nullptr, nullptr};      // int16 samples as they arrive (2 B each), before
  size_t x16_cap[2] = {0, 0};              // the on-device conversion; capacity in floats
};

struct Pipe {           // copy streams / events of the host-pointer pipeline
  bool ready = false;
  cudaStream_t s_in = nullptr, s_out = nullptr;
  cudaEvent_t ev_in[2], ev_comp[2], ev_out[2];
  // pinned bounce slots for callers whose buffers are pageable (malloc / mallocf memory, what
  // a caller of the reference API has): created on first use, grow-only
  char *pin_in[2] = {nullptr, nullptr}, *pin_out[2] = {nullptr, nullptr};
  size_t pin_in_cap[2] = {0, 0}, pin_out_cap[2] = {0, 0};
};

// A handful of host threads that copy between pageable and pinned memory.  cudaMemcpyAsync on
// pageable memory is staged by the driver on the calling thread at ~11 GB/s (H2D) and does not
// overlap with anything (tools/h2d_ceiling.cu); one memcpy thread manages ~14 GB/s, several
// together keep a PCIe 5 x16 link (~50 GB/s per direction) busy.
class CopyPool {
 public:
  static CopyPool &get() {
    // never destroyed: its detached workers sleep on the condition variable until the process
    // ends, and a static destructor would pull it from under them (exit() would hang)
    static CopyPool *pool = new CopyPool;
    return *pool;
  }
  int threads() const { return (int)workers_.size(); }
  // copies `rows` rows of `width` bytes (pitches in bytes); returns when all of it is done
  void copy2d(char *dst, size_t dst_pitch, const char *src, size_t src_pitch, size_t width, size_t rows) {
    if (width == 0 || rows == 0) return;
    const size_t chunk = (size_t)2 << 20;
    std::vector<std::function<void()>> tasks;
    if (dst_pitch == width && src_pitch == width) {       // one contiguous block
      const size_t total = width * rows;
      for (size_t off = 0; off < total; off += chunk) {
        const size_t n = total - off < chunk ? total - off : chunk;
        tasks.emplace_back([=]() { memcpy(dst + off, src + off, n); });
      }
    } else {
      size_t rows_per = chunk / width > 0 ? chunk / width : 1;
      for (size_t r0 = 0; r0 < rows; r0 += rows_per) {
        const size_t r1 = r0 + rows_per < rows ? r0 + rows_per : rows;
        tasks.emplace_back([=]() {
          for (size_t r = r0; r < r1; r++) memcpy(dst + r * dst_pitch, src + r * src_pitch, width);
        });
      }
    }
    run(tasks);
  }

 private:
  CopyPool() {
    unsigned hc = std::thread::hardware_concurrency();
    int n = (int)(hc / 2);
    if (n < 2) n = 2;
    if (n > 12) n = 12;
    for (int i = 0; i < n; i++) workers_.emplace_back([this]() { loop(); });
    for (auto &w : workers_) w.detach();                  // live for the life of the process
  }
  struct Batch { std::atomic<size_t> left; std::mutex m; std::condition_variable cv; };
  void run(std::vector<std::function<void()>> &tasks) {
    if (tasks.size() == 1) { tasks[0](); return; }
    Batch b;
    b.left.store(tasks.size());
    {
      std::lock_guard<std::mutex> lock(m_);
      for (auto &t : tasks) q_.emplace_back(&b, std::move(t));
    }
    cv_.notify_all();
    std::unique_lock<std::mutex> lock(b.m);
    b.cv.wait(lock, [&]() { return b.left.load() == 0; });
  }
  void loop() {
    for (;;) {
      std::pair<Batch *, std::function<void()>> job;
      {
        std::unique_lock<std::mutex> lock(m_);
        cv_.wait(lock, [&]() { return !q_.empty(); });
        job = std::move(q_.front());
        q_.pop_front();
      }
      job.second();
      Batch *b = job.first;
      std::lock_guard<std::mutex> lock(b->m);           // the waiter cannot leave before we are out
      if (b->left.fetch_sub(1) == 1) b->cv.notify_all();
    }
  }
  std::vector<std::thread> workers_;
  std::mutex m_;
  std::condition_variable cv_;
  std::deque<std::pair<Batch *, std::function<void()>>> q_;
};

bool is_pageable_host(const void *p) {
  cudaPointerAttributes attr;
  cudaError_t err = cudaPointerGetAttributes(&attr, p);
  if (err != cudaSuccess) { cudaGetLastError(); return true; }
  return attr.type == cudaMemoryTypeUnregistered;
}

void grow_pinned(char **buf, size_t *cap, size_t need) {
  if (need <= *cap) return;
  if (*buf) B200_CUDA_CHECK(cudaFreeHost(*buf));
  B200_CUDA_CHECK(cudaHostAlloc(buf, need, cudaHostAllocDefault));
  *cap = need;
}

// Host-pointer calls too small for the pipeline below to pay (a reference-API caller hands
// over a few KB per call): one pinned, device-mapped bounce buffer per device.
struct SmallPipe {
  float *pin = nullptr;                // cudaHostAlloc(Mapped): [x | h | result]
  size_t cap = 0;                      // floats
  std::mutex mu;
};
constexpr long long kSmallJobFloats = 64 * 1024;   // x + h + result of one call (256 KiB)

struct DeviceCtx {
  cudaStream_t own_stream = nullptr;   // created lazily, non-blocking
  SmallPipe small;
  Staging staging;                     // used by handle-less calls (convolve_simd, ...)
  Pipe pipe;
  float2 *spectrum = nullptr;          // scratch for handle-less FFT jobs
  size_t spectrum_cap = 0;
};

std::mutex g_mu;                 // guards contexts, the shared staging and user stream
std::mutex g_staged_call_mu;     // handle-less host-pointer calls share the per-device staging
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
  int real_packed_arg;
  OsDevPlan os;
  Staging staging;
  std::mutex mu;       // one thread per handle at a time (reference: shared scratch)
  // the same plan on the other GPUs of the box, created on first use by the multi-GPU calls
  // (simd_b200_use_devices) and destroyed with this one
  B200Plan *peer[kMaxDev] = {};
  std::mutex peer_mu;
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
  plan->real_packed_arg = real_packed;
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
  for (int d = 0; d < kMaxDev; d++) {
    if (plan->peer[d]) { b200_plan_destroy(plan->peer[d]); plan->peer[d] = nullptr; }
  }
  int prev = current_device();
  if (prev != plan->device) B200_CUDA_CHECK(cudaSetDevice(plan->device));
  // outstanding asynchronous work may still read the spectrum / staging buffers
  B200_CUDA_CHECK(cudaDeviceSynchronize());
  cudaFree(plan->os.spectrum);
  for (int sl = 0; sl < 2; sl++) {
    cudaFree(plan->staging.x[sl]);
    cudaFree(plan->staging.out[sl]);
    cudaFree(plan->staging.x16[sl]);
  }
  cudaFree(plan->staging.h);
  if (prev != plan->device) B200_CUDA_CHECK(cudaSetDevice(prev));
  plan->magic = 0;
  delete plan;
}

// pinned host memory for callers that do not link the CUDA runtime themselves (portable: every
// GPU's copy engine reads it in place, which is what the multi-GPU host-pointer path wants)
extern "C" float *b200_host_alloc_pinned(size_t floats) {
  float *p = nullptr;
  current_device();     // aborts with a message when there is no usable GPU
  B200_CUDA_CHECK(cudaHostAlloc(&p, sizeof(float) * (floats ? floats : 1), cudaHostAllocPortable));
  return p;
}
extern "C" void b200_host_free_pinned(float *p) {
  if (p) B200_CUDA_CHECK(cudaFreeHost(p));
}

// plain device buffers for the host layer (cross-correlate -> peaks keeps its result here)
extern "C" float *b200_device_alloc(size_t floats) {
  float *p = nullptr;
  B200_CUDA_CHECK(cudaMalloc(&p, sizeof(float) * (floats ? floats : 1)));
  return p;
}
extern "C" void b200_device_free(float *p) {
  if (p) B200_CUDA_CHECK(cudaFree(p));
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

  // ---- small host-pointer jobs --------------------------------------------------------
  // One memcpy into a pinned bounce buffer, one stream, no events, one synchronise:
  //   direct kernel   the kernel reads x / h from and writes the result to the bounce buffer
  //                   itself (mapped host memory: one launch, no copy commands at all)
  //   transform path  one H2D copy of [x | h], the two kernels, one D2H copy of the result
  // This is the path a drop-in caller of convolve() / convolve_simd() with the reference's
  // test shapes takes (tests/convolve.cc:139-166: x = 1024, h = 50).
  B200_REQUIRE(!(job->x_int16 && xd), "int16 samples are converted while they are staged: pass host memory");
  if (!xd && !hd && !od && job->batch == 1 && !job->x_int16 &&
      job->x_len + job->h_len + job->out_len <= kSmallJobFloats) {
    SmallPipe &sp = g_ctx[dev].small;
    std::unique_lock<std::mutex> plan_lock;
    if (plan) plan_lock = std::unique_lock<std::mutex>(plan->mu);
    std::lock_guard<std::mutex> lock(sp.mu);
    // 64-float (256-byte) aligned sections: wide loads and bulk copies stay legal
    const size_t xo = 0, ho = ((size_t)job->x_len + 63) & ~(size_t)63;
    const size_t oo = ho + (((size_t)job->h_len + 63) & ~(size_t)63);
    const size_t total = oo + (size_t)job->out_len;
    if (total > sp.cap) {
      if (sp.pin) { B200_CUDA_CHECK(cudaStreamSynchronize(stream)); B200_CUDA_CHECK(cudaFreeHost(sp.pin)); }
      sp.cap = (size_t)kSmallJobFloats + 256;
      B200_CUDA_CHECK(cudaHostAlloc(&sp.pin, sp.cap * sizeof(float), cudaHostAllocMapped));
    }
    memcpy(sp.pin + xo, job->x, sizeof(float) * (size_t)job->x_len);
    memcpy(sp.pin + ho, job->h, sizeof(float) * (size_t)job->h_len);
    if (algorithm == B200_ALG_DIRECT) {
      dj.x = sp.pin + xo; dj.h = sp.pin + ho; dj.out = sp.pin + oo;
      launch_device_job(plan, algorithm, dj, stream);
    } else {
      std::unique_lock<std::mutex> ctx_lock;
      if (!plan) ctx_lock = std::unique_lock<std::mutex>(g_staged_call_mu);
      Staging *st = plan ? &plan->staging : &g_ctx[dev].staging;
      grow(&st->x[0], &st->x_cap[0], oo + (size_t)job->h_len);   // [x | h] in one piece
      grow(&st->out[0], &st->out_cap[0], (size_t)job->out_len);
      B200_CUDA_CHECK(cudaMemcpyAsync(st->x[0], sp.pin, sizeof(float) * (ho + (size_t)job->h_len),
                                      cudaMemcpyHostToDevice, stream));
      dj.x = st->x[0] + xo; dj.h = st->x[0] + ho; dj.out = st->out[0];
      launch_device_job(plan, algorithm, dj, stream);
      B200_CUDA_CHECK(cudaMemcpyAsync(sp.pin + oo, st->out[0], sizeof(float) * (size_t)job->out_len,
                                      cudaMemcpyDeviceToHost, stream));
    }
    B200_CUDA_CHECK(cudaStreamSynchronize(stream));
    memcpy(job->out, sp.pin + oo, sizeof(float) * (size_t)job->out_len);
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
    ctx_lock = std::unique_lock<std::mutex>(g_staged_call_mu);
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
  // pageable caller memory goes through the pinned bounce slots (worth it from ~1 MiB per piece)
  const size_t es = job->x_int16 ? sizeof(short) : sizeof(float);
  const bool x_page = !xd && (size_t)piece_x * es >= ((size_t)1 << 20) && is_pageable_host(job->x);
  const bool o_page = !od && (size_t)piece_o * 4 >= ((size_t)1 << 20) && is_pageable_host(job->out);
  for (int sl = 0; sl < (n_pieces > 1 ? 2 : 1); sl++) {
    if (x_page) grow_pinned(&pipe.pin_in[sl], &pipe.pin_in_cap[sl], (size_t)piece_x * es);
    if (o_page) grow_pinned(&pipe.pin_out[sl], &pipe.pin_out_cap[sl], (size_t)piece_o * 4);
  }
  struct Unload { float *dst; i64 len, rows, pitch; int slot; bool pending; } prev = {nullptr, 0, 0, 0, 0, false};
  auto unload = [&](Unload &u) {          // pinned bounce slot -> the caller's result rows
    if (!u.pending) return;
    B200_CUDA_CHECK(cudaEventSynchronize(pipe.ev_out[u.slot]));
    CopyPool::get().copy2d(reinterpret_cast<char *>(u.dst), sizeof(float) * (size_t)u.pitch, pipe.pin_out[u.slot],
                           sizeof(float) * (size_t)u.len, sizeof(float) * (size_t)u.len, (size_t)u.rows);
    u.pending = false;
  };
  for (int sl = 0; sl < (n_pieces > 1 ? 2 : 1); sl++) {
    if (!xd) grow(&st->x[sl], &st->x_cap[sl], (size_t)piece_x);
    if (job->x_int16) grow(&st->x16[sl], &st->x16_cap[sl], ((size_t)piece_x + 1) / 2);
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
    i64 x_off;                       // first sample of the piece, in samples from job->x
    i64 px_len, po_len, p_rows, p_pad;
    if (slice_time) {
      const i64 o0 = pi * outs_per_piece;
      const i64 o1 = o0 + outs_per_piece < o_row ? o0 + outs_per_piece : o_row;
      i64 lo = o0 - job->pad; if (lo < 0) lo = 0;
      i64 hi = o1 - job->pad + job->h_len - 1; if (hi > x_row) hi = x_row;
      if (hi < lo) hi = lo;
      x_off = lo; px_len = hi - lo; p_pad = job->pad - o0 + lo;
      dst_o = job->out + o0; po_len = o1 - o0; p_rows = 1;
    } else {
      const i64 b0 = pi * rows_per_piece;
      p_rows = (job->batch - b0) < rows_per_piece ? (job->batch - b0) : rows_per_piece;
      x_off = b0 * job->x_stride; px_len = x_row; p_pad = job->pad;
      dst_o = job->out + b0 * job->out_stride; po_len = o_row;
    }
    src_x = job->x_int16 ? nullptr : job->x + x_off;
    dj.batch = p_rows; dj.x_len = px_len; dj.out_len = po_len; dj.pad = p_pad;
    // --- H2D ------------------------------------------------------------------------------
    // the caller's rows as the copy engine will read them: in place, or compacted into the
    // pinned bounce slot by the copy threads
    const char *h2d_src = reinterpret_cast<const char *>(job->x) + (size_t)x_off * es;
    i64 h2d_pitch = job->x_stride;
    if (x_page && px_len > 0) {
      if (pi >= 2) B200_CUDA_CHECK(cudaEventSynchronize(pipe.ev_in[sl]));   // slot's last H2D has read it
      CopyPool::get().copy2d(pipe.pin_in[sl], es * (size_t)px_len, h2d_src, es * (size_t)job->x_stride,
                             es * (size_t)px_len, (size_t)p_rows);
      h2d_src = pipe.pin_in[sl];
      h2d_pitch = px_len;
    }
    if (!xd) {
      if (pi >= 2) B200_CUDA_CHECK(cudaStreamWaitEvent(pipe.s_in, pipe.ev_comp[sl], 0));
      if (px_len > 0 && job->x_int16) {
        // 2 bytes per sample over PCIe; widened to float on the device, into the slot the
        // kernels read (reference converter: inc/simd/arithmetic.h:231-278)
        const short *src16 = reinterpret_cast<const short *>(h2d_src);
        short *dst16 = reinterpret_cast<short *>(st->x16[sl]);
        if (p_rows == 1 || h2d_pitch == px_len) {
          B200_CUDA_CHECK(cudaMemcpyAsync(dst16, src16, sizeof(short) * px_len * p_rows,
                                          cudaMemcpyHostToDevice, pipe.s_in));
        } else {
          B200_CUDA_CHECK(cudaMemcpy2DAsync(dst16, sizeof(short) * px_len, src16,
                                            sizeof(short) * h2d_pitch, sizeof(short) * px_len,
                                            (size_t)p_rows, cudaMemcpyHostToDevice, pipe.s_in));
        }
        b200_int16_to_float_on(dst16, px_len * p_rows, st->x[sl], pipe.s_in);
      } else if (px_len > 0) {
        if (p_rows == 1 || h2d_pitch == px_len) {   // contiguous: no 2-D pitch limit
          B200_CUDA_CHECK(cudaMemcpyAsync(st->x[sl], h2d_src, sizeof(float) * px_len * p_rows,
                                          cudaMemcpyHostToDevice, pipe.s_in));
        } else {
          B200_CUDA_CHECK(cudaMemcpy2DAsync(st->x[sl], sizeof(float) * px_len, h2d_src,
                                            sizeof(float) * h2d_pitch, sizeof(float) * px_len,
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
    if (o_page) {
      // into the pinned slot; the copy threads move the previous piece to the caller's rows
      // while this one is on the GPU
      B200_CUDA_CHECK(cudaStreamWaitEvent(pipe.s_out, pipe.ev_comp[sl], 0));
      B200_CUDA_CHECK(cudaMemcpyAsync(pipe.pin_out[sl], st->out[sl], sizeof(float) * po_len * p_rows,
                                      cudaMemcpyDeviceToHost, pipe.s_out));
      B200_CUDA_CHECK(cudaEventRecord(pipe.ev_out[sl], pipe.s_out));
      unload(prev);
      prev = Unload{dst_o, po_len, p_rows, p_rows == 1 ? po_len : job->out_stride, sl, true};
    } else if (!od) {
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
  unload(prev);
  if (!od) B200_CUDA_CHECK(cudaStreamSynchronize(pipe.s_out));
  B200_CUDA_CHECK(cudaStreamSynchronize(stream));
}


// ---- several GPUs from one process ---------------------------------------------------------
// The reference has no parallelism on this path (src/convolve.c:179-180 declines even OpenMP);
// north_star: batches of independent signals, and long signals cut into pieces with an
// (h-1)-sample halo, partition across the GPUs of one box with no collective.  A host-pointer
// job is cut into one slice per device -- groups of whole signals, or for a single long signal
// contiguous output ranges, each reading its own (h-1)-sample halo straight from the caller's
// host buffer -- and every slice runs through the single-GPU path above (b200_run) on its own
// device from its own host thread: own staging buffers, own copy streams, own PCIe link.

namespace {
std::atomic<int> g_use_devices{1};

B200Plan *plan_on_device(B200Plan *plan, int dev) {
  if (!plan || plan->device == dev) return plan;
  std::lock_guard<std::mutex> lock(plan->peer_mu);
  if (!plan->peer[dev]) {
    // current device is `dev` (the worker thread selected it)
    plan->peer[dev] = b200_plan_create(plan->algorithm, plan->x_len, plan->h_len, plan->os.n_fft,
                                       plan->real_packed_arg);
  }
  return plan->peer[dev];
}
}  // namespace

extern "C" void b200_use_devices(int n) {
  int have = b200_device_count();
  // test knob: VELES_B200_OVERSUBSCRIBE=1 lets n exceed the GPUs present (slices then share
  // devices round-robin), so the partitioning can be exercised on a one-GPU box
  if (n <= 0 || (n > have && !getenv("VELES_B200_OVERSUBSCRIBE"))) n = have;
  if (n > kMaxDev) n = kMaxDev;
  g_use_devices.store(n < 1 ? 1 : n);
}
extern "C" int b200_devices_in_use(void) { return g_use_devices.load(); }

extern "C" void b200_run_multi(B200Plan *plan, int algorithm, const B200Job *job) {
  const int ndev = g_use_devices.load();
  const bool host_job = !is_device_pointer(job->x);      // the bulk of the traffic is host memory
  // slices worth a thread and a pipeline each: at least 2^20 outputs per device
  const i64 work = job->batch * job->out_len;
  int parts = ndev;
  if (work / ((i64)1 << 20) < parts) parts = (int)(work / ((i64)1 << 20));
  if (job->batch > 1 && job->batch < parts) parts = (int)job->batch;
  if (!host_job || parts <= 1) {
    b200_run(plan, algorithm, job);
    return;
  }
  const int dev0 = current_device();
  const int have = b200_device_count();
  std::vector<std::thread> workers;
  for (int i = 0; i < parts; i++) {
    B200Job sj = *job;
    if (job->batch > 1) {
      const i64 base = job->batch / parts, rem = job->batch % parts;
      const i64 b0 = i * base + (i < rem ? i : rem);
      sj.batch = base + (i < rem ? 1 : 0);
      sj.x = job->x_int16 ? reinterpret_cast<const float *>(reinterpret_cast<const short *>(job->x) + b0 * job->x_stride)
                          : job->x + b0 * job->x_stride;
      sj.out = job->out + b0 * job->out_stride;
    } else {
      // contiguous output range [o0, o1) of one long signal; the samples it needs are
      // x[o0 - pad, o1 - pad + h - 1), clipped to the signal
      const i64 per = (job->out_len + parts - 1) / parts;
      const i64 o0 = i * per;
      const i64 o1 = o0 + per < job->out_len ? o0 + per : job->out_len;
      if (o1 <= o0) continue;
      i64 lo = o0 - job->pad; if (lo < 0) lo = 0;
      i64 hi = o1 - job->pad + job->h_len - 1; if (hi > job->x_len) hi = job->x_len;
      if (hi < lo) hi = lo;
      if (hi == lo) {       // nothing of the signal under this range: zeros
        memset(job->out + o0, 0, sizeof(float) * (size_t)(o1 - o0));
        continue;
      }
      sj.x = job->x_int16 ? reinterpret_cast<const float *>(reinterpret_cast<const short *>(job->x) + lo) : job->x + lo;
      sj.x_len = hi - lo; sj.x_stride = sj.x_len;
      sj.pad = job->pad - o0 + lo;
      sj.out = job->out + o0; sj.out_len = o1 - o0; sj.out_stride = sj.out_len;
    }
    const int dev = (dev0 + i) % have;
    workers.emplace_back([plan, algorithm, sj, dev]() {
      B200_CUDA_CHECK(cudaSetDevice(dev));
      // a long signal's slices have other lengths than the handle was made for: they run
      // without it (context-owned scratch, the same kernels)
      B200Plan *mine = sj.x_len == (plan ? plan->x_len : sj.x_len) ? plan_on_device(plan, dev) : nullptr;
      b200_run(mine, algorithm, &sj);
    });
  }
  for (std::thread &w : workers) w.join();
  B200_CUDA_CHECK(cudaSetDevice(dev0));
}

// Device-resident shards of one long signal, one per GPU: shard i holds x[s_i, e_i) at
// x_shards[i] with h_len-1 writable floats in front of it.  Each shard's tail is copied into
// the front of its right neighbour (cudaMemcpyPeerAsync on the neighbour's stream: the halo
// lands in local memory before the unchanged single-GPU kernels run; no kernel reads a peer's
// memory), then every GPU filters [halo | shard] with pad = 0 (shard 0: pad = h_len-1, its
// halo is the implicit zeros).  The last shard also produces the h_len-1 tail outputs.
extern "C" void b200_filter_long_sharded(int n_shards, const int *devices, float *const *x_shards,
                                         const size_t *shard_lengths, const float *const *h_on_device,
                                         size_t h_len, int reverse, float *const *result_shards,
                                         int algorithm) {
  B200_REQUIRE(n_shards >= 1 && devices && x_shards && shard_lengths && h_on_device && result_shards,
               "filter_long_sharded: null argument");
  const int dev0 = current_device();
  const i64 halo = (i64)h_len - 1;
  for (int i = 0; i < n_shards; i++) {
    B200_REQUIRE(shard_lengths[i] > 0 && (i == 0 || (i64)shard_lengths[i - 1] >= halo),
                 "filter_long_sharded: a shard is shorter than the h-1 halo its neighbour needs");
    B200_CUDA_CHECK(cudaSetDevice(devices[i]));
    cudaStream_t stream = current_stream();
    B200Job sj = {};
    sj.h = h_on_device[i]; sj.h_len = (i64)h_len; sj.reverse = reverse; sj.batch = 1;
    sj.out = result_shards[i];
    sj.out_len = (i64)shard_lengths[i] + (i == n_shards - 1 ? halo : 0);
    sj.out_stride = sj.out_len;
    if (i == 0 || halo == 0) {
      sj.x = x_shards[i]; sj.x_len = (i64)shard_lengths[i]; sj.pad = i == 0 ? halo : 0;
    } else {
      B200_CUDA_CHECK(cudaMemcpyPeerAsync(x_shards[i] - halo, devices[i],
                                          x_shards[i - 1] + shard_lengths[i - 1] - halo, devices[i - 1],
                                          sizeof(float) * (size_t)halo, stream));
      sj.x = x_shards[i] - halo; sj.x_len = (i64)shard_lengths[i] + halo; sj.pad = 0;
    }
    sj.x_stride = sj.x_len;
    b200_run(nullptr, algorithm, &sj);       // no handle: the device context's scratch
  }
  for (int i = 0; i < n_shards; i++) {
    B200_CUDA_CHECK(cudaSetDevice(devices[i]));
    B200_CUDA_CHECK(cudaStreamSynchronize(current_stream()));
  }
  B200_CUDA_CHECK(cudaSetDevice(dev0));
}
