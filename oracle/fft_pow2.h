/* oracle/fft_pow2.h -- TEST INFRASTRUCTURE, not product code.
 *
 * A small power-of-two float32 FFT used by (a) the FFTF-API shim that lets the
 * reference's own convolve.c / correlate.c compile here (FFTF, github.com/Samsung/FFTF,
 * pkg-config "fftf >= 1.0" per /root/reference/configure.ac:47-50, is a third-party
 * dependency that is NOT vendored under /root/reference and cannot be fetched), and
 * (b) the C restatement of the reference algorithm in conv_oracle.c.
 *
 * Semantics (inferred from the reference call sites, src/convolve.c:115-143,175,198,216,
 * 253-276,309,323):
 *   - real forward:  N floats  -> N/2+1 interleaved (re,im) bins in N+2 floats, in place OK
 *   - real backward: N/2+1 bins -> N floats, UN-normalised (caller multiplies by 1/N)
 *   - N is a power of two >= 2.
 *
 * Algorithm: real FFT through a half-length complex FFT (even/odd packing) with a
 * radix-4 (+ one radix-2 step when log2 is odd) Stockham autosort kernel.  Twiddles are
 * tabulated in double precision and stored as float.
 */
#ifndef ORACLE_FFT_POW2_H_
#define ORACLE_FFT_POW2_H_

#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  int n;          /* real length */
  int nc;         /* complex length = n/2 (>=1) */
  float *tw;      /* nc complex twiddles exp(-2*pi*i*k/nc), k < nc  (interleaved) */
  float *rtw;     /* n/2+1 complex twiddles exp(-2*pi*i*k/n) for the real<->complex step */
  float *a, *b;   /* two complex work buffers of nc entries */
} OracleFFT;

static OracleFFT *oracle_fft_create(int n) {
  OracleFFT *p = (OracleFFT *)calloc(1, sizeof(OracleFFT));
  p->n = n;
  p->nc = n / 2 > 0 ? n / 2 : 1;
  int nc = p->nc;
  p->tw = (float *)malloc(sizeof(float) * 2 * nc);
  p->rtw = (float *)malloc(sizeof(float) * 2 * (n / 2 + 1));
  p->a = (float *)malloc(sizeof(float) * 2 * nc);
  p->b = (float *)malloc(sizeof(float) * 2 * nc);
  const double pi2 = 6.283185307179586476925286766559;
  for (int k = 0; k < nc; k++) {
    p->tw[2 * k] = (float)cos(pi2 * k / nc);
    p->tw[2 * k + 1] = (float)(-sin(pi2 * k / nc));
  }
  for (int k = 0; k <= n / 2; k++) {
    p->rtw[2 * k] = (float)cos(pi2 * k / n);
    p->rtw[2 * k + 1] = (float)(-sin(pi2 * k / n));
  }
  return p;
}

static void oracle_fft_destroy(OracleFFT *p) {
  if (!p) return;
  free(p->tw); free(p->rtw); free(p->a); free(p->b); free(p);
}

/* Complex Stockham FFT of nc points held in p->a; result returned as a pointer to
 * whichever work buffer holds it.  sign = -1 forward, +1 backward (un-normalised). */
static float *oracle_cfft(OracleFFT *p, int sign) {
  int nc = p->nc;
  float *x = p->a, *y = p->b;
  const float *tw = p->tw;
  const float sg = (float)sign;
  int n = nc, s = 1;
  while (n >= 4) {
    int m = n / 4;
    int tstride = nc / n;
    for (int q0 = 0; q0 < m; q0++) {
      float w1r = tw[2 * (q0 * tstride)], w1i = -sg * tw[2 * (q0 * tstride) + 1];
      float w2r = tw[2 * (2 * q0 * tstride)], w2i = -sg * tw[2 * (2 * q0 * tstride) + 1];
      float w3r = tw[2 * (3 * q0 * tstride)], w3i = -sg * tw[2 * (3 * q0 * tstride) + 1];
      const float *xa = x + 2 * s * q0, *xb = x + 2 * s * (q0 + m);
      const float *xc = x + 2 * s * (q0 + 2 * m), *xd = x + 2 * s * (q0 + 3 * m);
      float *y0 = y + 2 * s * (4 * q0), *y1 = y0 + 2 * s, *y2 = y1 + 2 * s, *y3 = y2 + 2 * s;
      for (int q = 0; q < s; q++) {
        float ar = xa[2 * q], ai = xa[2 * q + 1], br = xb[2 * q], bi = xb[2 * q + 1];
        float cr = xc[2 * q], ci = xc[2 * q + 1], dr = xd[2 * q], di = xd[2 * q + 1];
        float apcr = ar + cr, apci = ai + ci, amcr = ar - cr, amci = ai - ci;
        float bpdr = br + dr, bpdi = bi + di, bmdr = br - dr, bmdi = bi - di;
        /* j*(b-d) with j = sign*i :  (r,i) -> sign*(-i_part, r_part) */
        float jr = -sg * bmdi, ji = sg * bmdr;
        float t1r = amcr + jr, t1i = amci + ji;
        float t2r = apcr - bpdr, t2i = apci - bpdi;
        float t3r = amcr - jr, t3i = amci - ji;
        y0[2 * q] = apcr + bpdr; y0[2 * q + 1] = apci + bpdi;
        y1[2 * q] = t1r * w1r - t1i * w1i; y1[2 * q + 1] = t1r * w1i + t1i * w1r;
        y2[2 * q] = t2r * w2r - t2i * w2i; y2[2 * q + 1] = t2r * w2i + t2i * w2r;
        y3[2 * q] = t3r * w3r - t3i * w3i; y3[2 * q + 1] = t3r * w3i + t3i * w3r;
      }
    }
    float *t = x; x = y; y = t;
    n /= 4; s *= 4;
  }
  if (n == 2) {
    /* last radix-2 step: twiddle is 1 */
    const float *xa = x, *xb = x + 2 * s;
    float *y0 = y, *y1 = y + 2 * s;
    for (int q = 0; q < s; q++) {
      float ar = xa[2 * q], ai = xa[2 * q + 1], br = xb[2 * q], bi = xb[2 * q + 1];
      y0[2 * q] = ar + br; y0[2 * q + 1] = ai + bi;
      y1[2 * q] = ar - br; y1[2 * q + 1] = ai - bi;
    }
    float *t = x; x = y; y = t;
  }
  return x;
}

/* in: n floats; out: n+2 floats (n/2+1 bins).  in == out allowed. */
static void oracle_rfft_forward(OracleFFT *p, const float *in, float *out) {
  int n = p->n, nc = p->nc;
  if (n == 1) { float v = in[0]; out[0] = v; out[1] = 0.f; return; }
  memcpy(p->a, in, sizeof(float) * n);           /* z[k] = in[2k] + i in[2k+1] */
  const float *Z = oracle_cfft(p, -1);
  const float *w = p->rtw;
  for (int k = 0; k <= nc; k++) {
    int k1 = k % nc, k2 = (nc - k) % nc;
    float zr = Z[2 * k1], zi = Z[2 * k1 + 1];
    float cr = Z[2 * k2], ci = -Z[2 * k2 + 1];    /* conj Z[nc-k] */
    float er = 0.5f * (zr + cr), ei = 0.5f * (zi + ci);
    float dr = 0.5f * (zr - cr), di = 0.5f * (zi - ci);
    /* O[k] = -i * d ;  X = E + W^k * O */
    float or_ = di, oi = -dr;
    float wr = w[2 * k], wi = w[2 * k + 1];
    out[2 * k] = er + (or_ * wr - oi * wi);
    out[2 * k + 1] = ei + (or_ * wi + oi * wr);
  }
}

/* in: n+2 floats (n/2+1 bins); out: n floats, un-normalised.  in == out allowed. */
static void oracle_rfft_backward(OracleFFT *p, const float *in, float *out) {
  int n = p->n, nc = p->nc;
  if (n == 1) { out[0] = in[0]; return; }
  const float *w = p->rtw;
  float *Zp = p->a;
  for (int k = 0; k < nc; k++) {
    float xr = in[2 * k], xi = in[2 * k + 1];
    float cr = in[2 * (nc - k)], ci = -in[2 * (nc - k) + 1];   /* conj X[nc-k] */
    float er = xr + cr, ei = xi + ci;
    float dr = xr - cr, di = xi - ci;
    float wr = w[2 * k], wi = -w[2 * k + 1];                   /* W^{-k} */
    float or_ = dr * wr - di * wi, oi = dr * wi + di * wr;
    /* Z' = E' + i O' */
    Zp[2 * k] = er - oi;
    Zp[2 * k + 1] = ei + or_;
  }
  const float *z = oracle_cfft(p, +1);
  memcpy(out, z, sizeof(float) * n);
}

#endif  /* ORACLE_FFT_POW2_H_ */
