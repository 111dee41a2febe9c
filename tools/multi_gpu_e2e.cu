// tools/multi_gpu_e2e.cu -- a C caller of the public API on all GPUs of the box, from ONE process.
//
// What a maintainer of the reference would write to use a multi-GPU box through the drop-in
// library: host buffers, one handle, convolve_batch; simd_b200_use_devices(n) spreads the
// rows over n GPUs (one worker thread + one staging pipeline per GPU inside the library).
// Prints one JSON object: end-to-end output GSamples/s for n = 1, 2, 4, ... GPUs with pinned
// (cudaHostAlloc, portable) and pageable (mallocf) buffers, config C4 (1024 x 262144, h = 4096)
// unless other sizes are given:  multi_gpu_e2e [batch x_len h_len]
// Build: nvcc -O2 -I include -o tools/multi_gpu_e2e tools/multi_gpu_e2e.cu -L veles/simd_b200 -lSimd \
//        -Xlinker -rpath -Xlinker '$ORIGIN/../veles/simd_b200'
#include <cuda_runtime.h>
#include <simd/b200.h>
#include <simd/convolve.h>
#include <simd/memory.h>
#include <chrono>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static double now() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int main(int argc, char **argv) {
  size_t batch = 1024, xl = 262144, hl = 4096;
  if (argc == 4) { batch = atol(argv[1]); xl = atol(argv[2]); hl = atol(argv[3]); }
  const size_t ol = xl + hl - 1;
  int have = simd_b200_device_count();
  float *x_pin, *y_pin, *h = mallocf(hl);
  if (cudaHostAlloc(&x_pin, sizeof(float) * batch * xl, cudaHostAllocPortable) != cudaSuccess ||
      cudaHostAlloc(&y_pin, sizeof(float) * batch * ol, cudaHostAllocPortable) != cudaSuccess) {
    printf("{\"error\": \"cudaHostAlloc failed\"}\n");
    return 1;
  }
  float *x_pg = mallocf(batch * xl), *y_pg = mallocf(batch * ol);
  unsigned s = 12345u;
  for (size_t i = 0; i < batch * xl; i++) { s = s * 1664525u + 1013904223u; x_pin[i] = (float)(s >> 8) / 8388608.0f - 1.0f; }
  memcpy(x_pg, x_pin, sizeof(float) * batch * xl);
  for (size_t i = 0; i < hl; i++) { s = s * 1664525u + 1013904223u; h[i] = (float)(s >> 8) / 8388608.0f - 1.0f; }
  ConvolutionHandle hd = convolve_initialize(xl, hl);
  printf("{\"batch\": %zu, \"x_length\": %zu, \"h_length\": %zu, \"gpus_present\": %d, \"rows\": [\n", batch, xl, hl, have);
  double sum1 = 0;
  for (int n = 1; n <= have; n *= 2) {
    simd_b200_use_devices(n);
    double rate[2];
    for (int kind = 0; kind < 2; kind++) {
      const float *x = kind == 0 ? x_pin : x_pg;
      float *y = kind == 0 ? y_pin : y_pg;
      convolve_batch(hd, x, xl, h, y, ol, batch);                 // warm-up: staging, per-device plans
      const int reps = 3;
      double t0 = now();
      for (int r = 0; r < reps; r++) convolve_batch(hd, x, xl, h, y, ol, batch);
      rate[kind] = (double)batch * ol * reps / (now() - t0) / 1e9;
      double sum = 0;
      for (size_t i = 0; i < batch * ol; i += 4099) sum += y[i];
      if (n == 1 && kind == 0) sum1 = sum;
      if (sum != sum1) { printf("RESULT MISMATCH n=%d kind=%d: %.9g vs %.9g\n", n, kind, sum, sum1); return 1; }
    }
    printf("%s {\"gpus\": %d, \"pinned_gsamples_s\": %.2f, \"pageable_gsamples_s\": %.2f}", n == 1 ? "" : ",\n", n, rate[0], rate[1]);
  }
  printf("]}\n");
  convolve_finalize(hd);
  return 0;
}
