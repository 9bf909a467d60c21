#include <cuda_fp16.h>
#include <stdio.h>
__global__ void k(const float* x, float* y, int n) {
  int i = threadIdx.x;
  if (i >= n) return;
  __half2 h = __floats2half2_rn(x[i], x[i] - 0.25f);
  unsigned u = *reinterpret_cast<unsigned*>(&h), r;
  asm("ex2.approx.f16x2 %0, %1;" : "=r"(r) : "r"(u));
  __half2 o = *reinterpret_cast<__half2*>(&r);
  y[2 * i] = __low2float(o), y[2 * i + 1] = __high2float(o);
}
int main() {
  float hx[12] = {0.f, -1.f, -13.9f, -14.5f, -16.f, -20.f, -23.5f, -24.5f, -30.f, 15.f, 15.99f, 16.5f}, hy[24];
  float *x, *y;
  cudaMalloc(&x, sizeof(hx)), cudaMalloc(&y, sizeof(hy));
  cudaMemcpy(x, hx, sizeof(hx), cudaMemcpyHostToDevice);
  k<<<1, 32>>>(x, y, 12);
  cudaMemcpy(hy, y, sizeof(hy), cudaMemcpyDeviceToHost);
  for (int i = 0; i < 12; ++i) printf("x=%7.2f ex2=%.6e (exact %.6e) | x-0.25 -> %.6e\n", hx[i], hy[2 * i], exp2f(hx[i]), hy[2 * i + 1]);
  return 0;
}
