// Microbenchmark: issue rate of legacy mma.sync.m16n8k16 (fp16 in, fp32 accumulate) on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o hmma_rate hmma_rate.cu && ./hmma_rate
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void mma(float (&d)[4], const unsigned (&a)[4], unsigned b0, unsigned b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <int ILP>
__global__ void k(float* out, int iters, long long* cyc) {
  unsigned a[4] = {threadIdx.x, threadIdx.x * 3u, 7u, 9u};
  float d[ILP][4];
  for (int i = 0; i < ILP; ++i) d[i][0] = d[i][1] = d[i][2] = d[i][3] = 0.f;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) mma(d[i], a, 0x3c003c00u + i, 0x3c003c00u);
  }
  long long t1 = clock64();
  float s = 0.f;
  for (int i = 0; i < ILP; ++i) s += d[i][0] + d[i][1] + d[i][2] + d[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int ILP>
void run(int warps_per_sm) {
  float* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4);
  cudaMalloc(&cyc, 8);
  const int iters = 4096;
  k<ILP><<<148, warps_per_sm * 32>>>(out, iters, cyc);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0), cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<ILP><<<148, warps_per_sm * 32>>>(out, iters, cyc);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  long long c;
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  const double mmas_per_smsp = (double)iters * ILP * warps_per_sm / 4.0;
  printf("ILP %d warps/SM %2d: %.3f ms, %lld cycles, %.2f cycles per HMMA per SMSP, %.1f TFLOP/s chip\n", ILP, warps_per_sm,
         ms, c, c / mmas_per_smsp, 148.0 * warps_per_sm * iters * ILP * 4096.0 / (ms * 1e-3) / 1e12);
  cudaFree(out), cudaFree(cyc);
}
int main() {
  run<1>(4), run<4>(4), run<8>(4), run<8>(8), run<8>(16), run<4>(16);
  return 0;
}
