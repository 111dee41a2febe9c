/* oracle/fftf_shim/fftf_shim.c -- TEST INFRASTRUCTURE, not product code.
 * Implements the four FFTF entry points the reference's convolve path calls, on top of
 * oracle/fft_pow2.h.  Real, 1-D, power-of-two lengths only (all the path needs).
 * `lengths` is read at calc time through the stored pointer, as the reference keeps it on
 * the heap for exactly that reason (src/convolve.c:125-126,245-246). */
#include "fftf/api.h"
#include <assert.h>
#include "../fft_pow2.h"

struct FFTFInstance {
  FFTFDirection direction;
  const int *lengths;
  int batch;
  const float **inputs;
  float **outputs;
  OracleFFT *fft;
};

FFTFInstance *fftf_init_batch(FFTFType type, FFTFDirection direction,
                              FFTFDimension dimension, const int *lengths,
                              FFTFOptions options, int batchSize,
                              const float *const *inputs, float *const *outputs) {
  (void)options;
  assert(type == FFTF_TYPE_REAL);
  assert(dimension == FFTF_DIMENSION_1D);
  int n = lengths[0];
  assert(n > 0 && (n & (n - 1)) == 0);
  FFTFInstance *inst = (FFTFInstance *)calloc(1, sizeof(FFTFInstance));
  inst->direction = direction;
  inst->lengths = lengths;
  inst->batch = batchSize;
  inst->inputs = (const float **)malloc(sizeof(float *) * batchSize);
  inst->outputs = (float **)malloc(sizeof(float *) * batchSize);
  for (int i = 0; i < batchSize; i++) {
    inst->inputs[i] = inputs[i];
    inst->outputs[i] = outputs[i];
  }
  inst->fft = oracle_fft_create(n);
  return inst;
}

FFTFInstance *fftf_init(FFTFType type, FFTFDirection direction,
                        FFTFDimension dimension, const int *lengths,
                        FFTFOptions options, const float *input, float *output) {
  const float *ins[1] = { input };
  float *outs[1] = { output };
  return fftf_init_batch(type, direction, dimension, lengths, options, 1, ins, outs);
}

void fftf_calc(const FFTFInstance *inst) {
  for (int i = 0; i < inst->batch; i++) {
    if (inst->direction == FFTF_DIRECTION_FORWARD) {
      oracle_rfft_forward(inst->fft, inst->inputs[i], inst->outputs[i]);
    } else {
      oracle_rfft_backward(inst->fft, inst->inputs[i], inst->outputs[i]);
    }
  }
}

void fftf_destroy(FFTFInstance *inst) {
  if (!inst) return;
  oracle_fft_destroy(inst->fft);
  free(inst->inputs);
  free(inst->outputs);
  free(inst);
}
