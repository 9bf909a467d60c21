// Microbenchmark: do legacy HMMA (mma.sync m16n8k16), MUFU.EX2 and FFMA overlap on one scheduler of sm_100a?
// Per iteration and warp: NH HMMAs, NM MUFUs, NF FFMAs, all independent, interleaved in program order.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_overlap pipe_overlap.cu && ./pipe_overlap
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void mma(float (&d)[4], const unsigned (&a)[4], unsigned b0, unsigned b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
template <int NH, int NM, int NF>
__global__ void k(float* out, int iters, long long* cyc) {
  unsigned a[4] = {threadIdx.x, threadIdx.x * 3u, 7u, 9u};
  float d[8][4], v[8], w[8];
  for (int i = 0; i < 8; ++i) d[i][0] = d[i][1] = d[i][2] = d[i][3] = 0.f, v[i] = threadIdx.x * 1e-3f + i, w[i] = 0.f;
  const float fa = 0.999f + threadIdx.x * 1e-9f, fb = 1e-4f;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (i < NH) mma(d[i], a, 0x3c003c00u + i, 0x3c003c00u);
#pragma unroll
      for (int m = 0; m < NM / 8; ++m) v[i] = ex2(v[i]) * 0.f + v[i];
#pragma unroll
      for (int f = 0; f < NF / 8; ++f) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(w[i]) : "f"(fa), "f"(fb));
    }
  }
  long long t1 = clock64();
  float s = 0.f;
  for (int i = 0; i < 8; ++i) s += d[i][0] + d[i][1] + d[i][2] + d[i][3] + v[i] + w[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int NH, int NM, int NF>
void run(int warps_per_sm) {
  float* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4);
  cudaMalloc(&cyc, 8);
  const int iters = 2048;
  for (int r = 0; r < 2; ++r) {
    k<NH, NM, NF><<<148, warps_per_sm * 32>>>(out, iters, cyc);
    cudaDeviceSynchronize();
  }
  long long c;
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  const double per = (double)c / iters / (warps_per_sm / 4.0);  // cycles per warp-iteration per scheduler
  printf("per warp-iteration: %d HMMA + %2d MUFU (+%2d FMA-pipe ops) + %2d FFMA, %2d warps/SM: %7.1f cycles per scheduler\n", NH,
         NM, NM, NF, warps_per_sm, per);
  cudaFree(out), cudaFree(cyc);
}
int main() {
  run<8, 0, 0>(16), run<0, 16, 0>(16), run<0, 0, 64>(16);
  run<8, 16, 0>(16), run<8, 0, 64>(16), run<0, 16, 64>(16), run<8, 16, 64>(16);
  run<8, 16, 64>(8), run<8, 16, 64>(4);
  return 0;
}
