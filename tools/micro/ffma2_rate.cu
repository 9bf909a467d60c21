// Microbenchmark: issue rate of fma.rn.f32x2 (FFMA2) against scalar FFMA on sm_100a, per scheduler.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_rate ffma2_rate.cu && ./ffma2_rate
#include <cstdio>
#include <cuda_runtime.h>
template <int PACKED, int ILP>
__global__ void k(float* out, int iters, long long* cyc) {
  unsigned long long v[ILP];
  float w[2 * ILP];
  for (int i = 0; i < ILP; ++i) v[i] = 0ull, w[2 * i] = w[2 * i + 1] = 0.f;
  const float fa = 0.999f + threadIdx.x * 1e-9f, fb = 1e-4f;
  unsigned long long pa, pb;
  asm("mov.b64 %0, {%1, %1};" : "=l"(pa) : "f"(fa));
  asm("mov.b64 %0, {%1, %1};" : "=l"(pb) : "f"(fb));
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      if (PACKED) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v[i]) : "l"(pa), "l"(pb));
      else {
        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(w[2 * i]) : "f"(fa), "f"(fb));
        asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(w[2 * i + 1]) : "f"(fa), "f"(fb));
      }
    }
  }
  long long t1 = clock64();
  float s = 0.f;
  for (int i = 0; i < ILP; ++i) s += (float)(v[i] & 0xffff) + w[2 * i] + w[2 * i + 1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int PACKED, int ILP>
void run(int warps_per_sm) {
  float* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4);
  cudaMalloc(&cyc, 8);
  const int iters = 4096;
  for (int r = 0; r < 2; ++r) {
    k<PACKED, ILP><<<148, warps_per_sm * 32>>>(out, iters, cyc);
    cudaDeviceSynchronize();
  }
  long long c;
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  const double per = (double)c / iters / ILP / (warps_per_sm / 4.0);
  printf("%s ILP %2d, %2d warps/SM: %.2f cycles per scheduler per %s = %.1f fp32 FMA lanes / clk / SM\n",
         PACKED ? "FFMA2     " : "2 x FFMA  ", ILP, warps_per_sm, per, PACKED ? "FFMA2" : "FFMA pair", 4 * 64.0 / per);
  cudaFree(out), cudaFree(cyc);
}
int main() {
  run<1, 8>(4), run<1, 8>(8), run<1, 8>(16), run<1, 16>(16);
  run<0, 8>(4), run<0, 8>(8), run<0, 8>(16), run<0, 16>(16);
  return 0;
}
