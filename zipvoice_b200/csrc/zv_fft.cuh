// 512-point complex FFT shared by the mel front-end (real FFT-1024 by even/odd packing) and the
// vocoder's iSTFT (inverse real FFT-1024).  Stockham autosort, radix 8, three passes; one
// transform is carried by 64 threads holding 8 complex values each, passes exchange data through
// a 512-entry shared-memory buffer.  The per-thread pass functions are __host__ __device__ so the
// index arithmetic is unit-tested on the CPU (tests/test_host_cpu.py builds csrc/zv_fft_selftest).
#pragma once
#include <cuda_runtime.h>

namespace zv {

#ifdef __CUDACC__
#define ZV_HD __host__ __device__ __forceinline__
#else
#define ZV_HD inline
#endif

constexpr int FFT_N = 512;      // complex points
constexpr int FFT_THREADS = 64; // threads per transform

ZV_HD float2 cmul(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
ZV_HD float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
ZV_HD float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
ZV_HD float2 cconj(float2 a) { return make_float2(a.x, -a.y); }
// multiply by -i (forward) or +i (inverse)
template <bool INV>
ZV_HD float2 crot(float2 a) {
  return INV ? make_float2(-a.y, a.x) : make_float2(a.y, -a.x);
}

// In-register 8-point DFT (decimation in frequency, natural-order output).
template <bool INV>
ZV_HD void dft8(float2 (&v)[8]) {
  const float h = 0.70710678118654752440f;
  float2 a0 = cadd(v[0], v[4]), a4 = csub(v[0], v[4]);
  float2 a1 = cadd(v[1], v[5]), a5 = csub(v[1], v[5]);
  float2 a2 = cadd(v[2], v[6]), a6 = csub(v[2], v[6]);
  float2 a3 = cadd(v[3], v[7]), a7 = csub(v[3], v[7]);
  // twiddles w8^1, w8^2, w8^3 with w8 = exp(-+ 2 pi i / 8)
  a5 = INV ? make_float2(h * (a5.x - a5.y), h * (a5.x + a5.y)) : make_float2(h * (a5.x + a5.y), h * (a5.y - a5.x));
  a6 = crot<INV>(a6);
  a7 = INV ? make_float2(-h * (a7.x + a7.y), h * (a7.x - a7.y)) : make_float2(h * (a7.y - a7.x), -h * (a7.x + a7.y));
  float2 b0 = cadd(a0, a2), b2 = csub(a0, a2), b1 = cadd(a1, a3), b3 = crot<INV>(csub(a1, a3));
  float2 c0 = cadd(a4, a6), c2 = csub(a4, a6), c1 = cadd(a5, a7), c3 = crot<INV>(csub(a5, a7));
  v[0] = cadd(b0, b1), v[4] = csub(b0, b1), v[2] = cadd(b2, b3), v[6] = csub(b2, b3);
  v[1] = cadd(c0, c1), v[5] = csub(c0, c1), v[3] = cadd(c2, c3), v[7] = csub(c2, c3);
}

// One Stockham pass for "thread" j (0..63): v[r] = in[j + 64 r] on entry.  Applies the inter-pass
// twiddles exp(-+ 2 pi i k r / (8 Ns)), k = j mod Ns, from tw512[m] = exp(-2 pi i m / 512), the
// 8-point DFT, and returns the base output index; out[base + r * Ns] = v[r].
template <bool INV>
ZV_HD int fft_pass(float2 (&v)[8], int j, int Ns, const float2* tw512) {
  const int k = j & (Ns - 1);
  if (Ns > 1) {
    const int step = k * (FFT_N / (Ns * 8));
#pragma unroll
    for (int r = 1; r < 8; ++r) {
      float2 t = tw512[(step * r) & (FFT_N - 1)];
      v[r] = cmul(v[r], INV ? cconj(t) : t);
    }
  }
  dft8<INV>(v);
  return (j - k) * 8 + k;
}

#ifdef __CUDACC__
// 64 cooperating threads (all threads of the CTA must call this in lock-step: it uses
// __syncthreads).  On entry v[r] = x[j + 64 r]; on return buf[0..511] holds the spectrum in
// natural order (after the trailing barrier).
template <bool INV>
__device__ __forceinline__ void fft512_block(float2 (&v)[8], int j, float2* buf, const float2* tw512) {
#pragma unroll
  for (int Ns = 1; Ns < FFT_N; Ns *= 8) {
    const int base = fft_pass<INV>(v, j, Ns, tw512);
    if (Ns > 1) __syncthreads();  // everyone has finished reading buf for this pass
#pragma unroll
    for (int r = 0; r < 8; ++r) buf[base + r * Ns] = v[r];
    __syncthreads();
    if (Ns * 8 < FFT_N) {
#pragma unroll
      for (int r = 0; r < 8; ++r) v[r] = buf[j + 64 * r];
    }
  }
}
#endif

// Real FFT-1024 from the packed transform Z = FFT512(x[2n] + i x[2n+1]):
// X[k] = (Z[k] + conj Z[512-k]) / 2 + exp(-2 pi i k / 1024) (Z[k] - conj Z[512-k]) / (2i),  k = 0..512
ZV_HD float2 rfft_unpack(const float2* Z, int k, float2 tw1024_k) {
  const float2 zk = Z[k & (FFT_N - 1)], zc = cconj(Z[(FFT_N - k) & (FFT_N - 1)]);
  const float2 e = cadd(zk, zc), o = csub(zk, zc);
  const float2 od = make_float2(0.5f * o.y, -0.5f * o.x);  // o / (2i)
  return cadd(make_float2(0.5f * e.x, 0.5f * e.y), cmul(tw1024_k, od));
}
// Inverse packing: Z[k] = (X[k] + conj X[512-k]) + i exp(+2 pi i k / 1024) (X[k] - conj X[512-k]),  k = 0..511;
// x[2n] + i x[2n+1] = IFFT512(Z)[n] / 1024 (unnormalised inverse).
ZV_HD float2 irfft_pack(float2 xk, float2 xnk, float2 tw1024_k) {
  const float2 xc = cconj(xnk);
  const float2 e = cadd(xk, xc), o = csub(xk, xc);
  const float2 t = cmul(cconj(tw1024_k), o);
  return cadd(e, make_float2(-t.y, t.x));
}

}  // namespace zv
