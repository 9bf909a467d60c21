// Microbenchmark: issue rate of MUFU.EX2 (ex2.approx.ftz.f32), alone and mixed with FFMA, on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu_rate mufu_rate.cu && ./mufu_rate
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// per iteration and thread: ILP ex2 and ILP * FPM dependent-free FFMAs
template <int ILP, int FPM>
__global__ void k(float* out, int iters, long long* cyc) {
  float v[ILP], w[ILP];
  for (int i = 0; i < ILP; ++i) v[i] = threadIdx.x * 1e-3f + i, w[i] = 0.f;
  const float a = 0.999f + threadIdx.x * 1e-9f, b = 1e-4f;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      v[i] = ex2(v[i]) * 0.f + v[i];  // keeps the value bounded; the multiply-add is one FFMA
#pragma unroll
      for (int f = 0; f < FPM; ++f) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(w[i]) : "f"(a), "f"(b));
    }
  }
  long long t1 = clock64();
  float s = 0.f;
  for (int i = 0; i < ILP; ++i) s += v[i] + w[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int ILP, int FPM>
void run(int warps_per_sm) {
  float* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4);
  cudaMalloc(&cyc, 8);
  const int iters = 4096;
  k<ILP, FPM><<<148, warps_per_sm * 32>>>(out, iters, cyc);
  cudaDeviceSynchronize();
  k<ILP, FPM><<<148, warps_per_sm * 32>>>(out, iters, cyc);
  cudaDeviceSynchronize();
  long long c;
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  const double per_smsp = (double)iters * ILP * warps_per_sm / 4.0;
  printf("ILP %d, %d FFMA per MUFU, warps/SM %2d: %.2f cycles per warp-MUFU per SMSP (%.1f MUFU lanes/clk/SM)\n", ILP,
         FPM + 1, warps_per_sm, c / per_smsp, 32.0 * 4 / (c / per_smsp));
  cudaFree(out), cudaFree(cyc);
}
int main() {
  run<8, 0>(4), run<8, 0>(8), run<8, 0>(16), run<8, 3>(16), run<8, 7>(16), run<8, 9>(16), run<8, 15>(16);
  return 0;
}
