// tools/h2d_ceiling.cu -- what the host link of this box can carry, to put the end-to-end
// numbers of the host-pointer path (bench.py `e2e`) against a ceiling.
//
// For n = 1, 2, 4, ... GPUs at once: plain pinned cudaMemcpyAsync of 256 MiB buffers, host to
// device only, device to host only, and both directions together (PCIe is full duplex), one
// host thread and two streams per GPU, 8 copies per direction back to back.  Prints the
// aggregate GB/s per direction.  Pageable (malloc) memory is timed on one GPU as well: that is
// what a caller of the reference API hands over.
// Build: nvcc -O2 -o tools/h2d_ceiling tools/h2d_ceiling.cu -lpthread
#include <cuda_runtime.h>
#include <chrono>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <thread>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

static const size_t kBytes = 256u << 20;
static const int kReps = 8;

struct Dev {
  int id;
  char *h_in, *h_out, *d_in, *d_out;
  cudaStream_t s_in, s_out;
};

static double now() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// mode 1 = H2D, 2 = D2H, 3 = both
static double run(std::vector<Dev> &devs, int n, int mode) {
  auto enqueue = [&](Dev &d) {
    CK(cudaSetDevice(d.id));
    for (int r = 0; r < kReps; r++) {
      if (mode & 1) CK(cudaMemcpyAsync(d.d_in, d.h_in, kBytes, cudaMemcpyHostToDevice, d.s_in));
      if (mode & 2) CK(cudaMemcpyAsync(d.h_out, d.d_out, kBytes, cudaMemcpyDeviceToHost, d.s_out));
    }
  };
  auto wait = [&](Dev &d) {
    CK(cudaSetDevice(d.id));
    CK(cudaStreamSynchronize(d.s_in));
    CK(cudaStreamSynchronize(d.s_out));
  };
  for (int i = 0; i < n; i++) { enqueue(devs[i]); }       // warm-up
  for (int i = 0; i < n; i++) wait(devs[i]);
  double t0 = now();
  std::vector<std::thread> th;
  for (int i = 0; i < n; i++) th.emplace_back([&, i]() { enqueue(devs[i]); wait(devs[i]); });
  for (auto &t : th) t.join();
  double dt = now() - t0;
  return (double)kBytes * kReps * n / dt / 1e9;           // per direction
}

int main() {
  int have = 0;
  CK(cudaGetDeviceCount(&have));
  std::vector<Dev> devs(have);
  for (int i = 0; i < have; i++) {
    Dev &d = devs[i];
    d.id = i;
    CK(cudaSetDevice(i));
    CK(cudaHostAlloc(&d.h_in, kBytes, cudaHostAllocPortable));
    CK(cudaHostAlloc(&d.h_out, kBytes, cudaHostAllocPortable));
    memset(d.h_in, 1, kBytes);
    memset(d.h_out, 0, kBytes);
    CK(cudaMalloc(&d.d_in, kBytes));
    CK(cudaMalloc(&d.d_out, kBytes));
    CK(cudaStreamCreateWithFlags(&d.s_in, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&d.s_out, cudaStreamNonBlocking));
  }
  printf("{\"gpus_present\": %d, \"buffer_mib\": %zu, \"copies_per_direction\": %d, \"rows\": [\n", have, kBytes >> 20, kReps);
  bool first = true;
  for (int n = 1; n <= have; n *= 2) {
    double h2d = run(devs, n, 1), d2h = run(devs, n, 2), both = run(devs, n, 3);
    printf("%s {\"gpus\": %d, \"h2d_gbs\": %.1f, \"d2h_gbs\": %.1f, \"duplex_gbs_per_direction\": %.1f}", first ? "" : ",\n", n, h2d,
           d2h, both);
    first = false;
  }
  // pageable memory, one GPU, synchronous copies (what cudaMemcpy does with malloc'ed memory)
  {
    CK(cudaSetDevice(0));
    char *p = (char *)malloc(kBytes);
    memset(p, 2, kBytes);
    CK(cudaMemcpy(devs[0].d_in, p, kBytes, cudaMemcpyHostToDevice));
    double t0 = now();
    for (int r = 0; r < 4; r++) CK(cudaMemcpy(devs[0].d_in, p, kBytes, cudaMemcpyHostToDevice));
    double h2d = (double)kBytes * 4 / (now() - t0) / 1e9;
    t0 = now();
    for (int r = 0; r < 4; r++) CK(cudaMemcpy(p, devs[0].d_out, kBytes, cudaMemcpyDeviceToHost));
    double d2h = (double)kBytes * 4 / (now() - t0) / 1e9;
    // and a plain host memcpy of the same buffer into pinned memory (the bounce alternative)
    t0 = now();
    for (int r = 0; r < 4; r++) memcpy(devs[0].h_in, p, kBytes);
    double mc = (double)kBytes * 4 / (now() - t0) / 1e9;
    printf("],\n \"pageable_one_gpu\": {\"h2d_gbs\": %.1f, \"d2h_gbs\": %.1f, \"host_memcpy_to_pinned_gbs_one_thread\": %.1f", h2d, d2h, mc);
    // the same copy split over several threads (what the library's bounce pipeline does), and
    // both directions at once (pageable -> pinned and pinned -> pageable, half the threads each)
    char *q = (char *)malloc(kBytes);
    memset(q, 3, kBytes);
    printf(", \"host_memcpy_threads\": [");
    int counts[] = {2, 4, 8, 12, 16, 24, 32};
    for (int ci = 0; ci < 7; ci++) {
      int nt = counts[ci];
      auto run_mt = [&](bool both) {
        double t1 = now();
        for (int r = 0; r < 4; r++) {
          std::vector<std::thread> th;
          for (int i = 0; i < nt; i++) th.emplace_back([&, i]() {
            size_t per = kBytes / nt, off = per * i;
            if (both && (i & 1)) memcpy(q + off, devs[0].h_out + off, per);
            else memcpy(devs[0].h_in + off, p + off, per);
          });
          for (auto &t : th) t.join();
        }
        return (double)kBytes * 4 / (now() - t1) / 1e9;
      };
      double one = run_mt(false), two = run_mt(true);
      printf("%s{\"threads\": %d, \"to_pinned_gbs\": %.1f, \"both_ways_total_gbs\": %.1f}", ci ? ", " : "", nt, one, two);
    }
    printf("]}}\n");
    free(q);
    free(p);
  }
  return 0;
}
