// Host-side check of the FFT index arithmetic in zv_fft.cuh: emulates the 64 cooperating threads
// sequentially and compares with a direct O(N^2) DFT in double.  Build: nvcc -o t zv_fft_selftest.cu
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "../zv_fft.cuh"
using namespace zv;

template <bool INV>
static void fft512_host(std::vector<float2>& x, const float2* tw) {
  std::vector<float2> buf = x, out(FFT_N);
  for (int Ns = 1; Ns < FFT_N; Ns *= 8) {
    for (int j = 0; j < 64; ++j) {
      float2 v[8];
      for (int r = 0; r < 8; ++r) v[r] = buf[j + 64 * r];
      const int base = fft_pass<INV>(v, j, Ns, tw);
      for (int r = 0; r < 8; ++r) out[base + r * Ns] = v[r];
    }
    buf = out;
  }
  x = buf;
}

int main() {
  std::vector<float2> tw(512), tw1024(1024);
  for (int i = 0; i < 512; ++i) tw[i] = make_float2((float)cos(-2 * M_PI * i / 512), (float)sin(-2 * M_PI * i / 512));
  for (int i = 0; i < 1024; ++i)
    tw1024[i] = make_float2((float)cos(-2 * M_PI * i / 1024), (float)sin(-2 * M_PI * i / 1024));
  srand(1);
  std::vector<double> xr(1024);
  for (auto& v : xr) v = rand() / (double)RAND_MAX - 0.5;
  // real FFT-1024
  std::vector<float2> z(512);
  for (int n = 0; n < 512; ++n) z[n] = make_float2((float)xr[2 * n], (float)xr[2 * n + 1]);
  fft512_host<false>(z, tw.data());
  double err = 0, nrm = 0;
  std::vector<float2> X(513);
  for (int k = 0; k <= 512; ++k) {
    X[k] = rfft_unpack(z.data(), k, tw1024[k]);
    double re = 0, im = 0;
    for (int n = 0; n < 1024; ++n) re += xr[n] * cos(2 * M_PI * k * n / 1024), im -= xr[n] * sin(2 * M_PI * k * n / 1024);
    err += (X[k].x - re) * (X[k].x - re) + (X[k].y - im) * (X[k].y - im), nrm += re * re + im * im;
  }
  const double e1 = sqrt(err / nrm);
  // inverse
  X[0].y = 0.3f, X[512].y = 0.7f;  // must be ignored like irfft does
  std::vector<float2> Z(512);
  for (int k = 0; k < 512; ++k) {
    float2 xk = X[k], xn = X[512 - k];
    if (k == 0) xk.y = 0.f, xn.y = 0.f;
    Z[k] = irfft_pack(xk, xn, tw1024[k]);
  }
  fft512_host<true>(Z, tw.data());
  err = 0, nrm = 0;
  for (int n = 0; n < 512; ++n) {
    const double a = Z[n].x / 1024.0 - xr[2 * n], b = Z[n].y / 1024.0 - xr[2 * n + 1];
    err += a * a + b * b, nrm += xr[2 * n] * xr[2 * n] + xr[2 * n + 1] * xr[2 * n + 1];
  }
  const double e2 = sqrt(err / nrm);
  printf("rfft rel-L2 %.3e  irfft rel-L2 %.3e\n", e1, e2);
  return (e1 < 1e-6 && e2 < 1e-6) ? 0 : 1;
}
