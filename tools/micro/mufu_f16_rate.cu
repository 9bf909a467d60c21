// Microbenchmark: MUFU.EX2 throughput per warp instruction on sm_100a for fp32 (ex2.approx.ftz.f32), fp16 packed
// (ex2.approx.f16x2, which SASS splits into two MUFU.EX2.F16) and bf16 packed (ex2.approx.ftz.bf16x2), plus the
// legacy HMMA m16n8k16 (bf16 and fp16 inputs, fp32 accumulate) with independent accumulators.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mufu_f16_rate mufu_f16_rate.cu && ./mufu_f16_rate
#include <cstdio>
#include <cuda_runtime.h>
template <int KIND, int ILP>
__global__ void k(unsigned* out, int iters, long long* cyc) {
  unsigned v[ILP];
  for (int i = 0; i < ILP; ++i) v[i] = 0x30003000u + threadIdx.x + i;  // small positive halves / bf16 / a small float
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      if (KIND == 0) asm volatile("{.reg .f32 t; mov.b32 t, %0; ex2.approx.ftz.f32 t, t; mov.b32 %0, t;}" : "+r"(v[i]));
      if (KIND == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(v[i]));
      if (KIND == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(v[i]));
      v[i] &= 0x3fff3fffu;  // keep the values bounded (one LOP3 per MUFU)
    }
  }
  long long t1 = clock64();
  unsigned s = 0;
  for (int i = 0; i < ILP; ++i) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int F16>
__global__ void hmma(float* out, int iters, long long* cyc) {
  float d[4][4] = {};
  unsigned a[4] = {0x3c003c00u + threadIdx.x, 0x3c003c00u, 0x3c003c00u, 0x3c003c00u}, b0 = 0x38003800u, b1 = 0x38003800u;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (F16)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(d[i][0]), "+f"(d[i][1]), "+f"(d[i][2]), "+f"(d[i][3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
      else
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(d[i][0]), "+f"(d[i][1]), "+f"(d[i][2]), "+f"(d[i][3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
    }
  }
  long long t1 = clock64();
  float s = 0.f;
  for (int i = 0; i < 4; ++i) s += d[i][0] + d[i][1] + d[i][2] + d[i][3];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int KIND>
void run(const char* name, int warps_per_sm) {
  unsigned* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4), cudaMalloc(&cyc, 8);
  const int iters = 4096;
  for (int r = 0; r < 2; ++r) k<KIND, 8><<<148, warps_per_sm * 32>>>(out, iters, cyc), cudaDeviceSynchronize();
  long long c;
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  const double per = c / ((double)iters * 8 * warps_per_sm / 4.0);
  printf("%-28s warps/SM %2d: %6.2f cycles per warp instruction per SMSP -> %5.1f exponentials / clk / SM\n", name,
         warps_per_sm, per, (KIND == 0 ? 32.0 : 64.0) * 4 / per);
  cudaFree(out), cudaFree(cyc);
}
template <int F16>
void run_hmma(const char* name, int warps_per_sm) {
  float* out;
  long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4), cudaMalloc(&cyc, 8);
  const int iters = 4096;
  for (int r = 0; r < 2; ++r) hmma<F16><<<148, warps_per_sm * 32>>>(out, iters, cyc), cudaDeviceSynchronize();
  long long c;
  cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-28s warps/SM %2d: %6.2f cycles per HMMA.16816 per SMSP\n", name, warps_per_sm,
         c / ((double)iters * 4 * warps_per_sm / 4.0));
  cudaFree(out), cudaFree(cyc);
}
int main() {
  for (int w : {4, 8, 16}) {
    run<0>("ex2.approx.ftz.f32", w), run<1>("ex2.approx.f16x2", w), run<2>("ex2.approx.ftz.bf16x2", w);
    run_hmma<0>("mma.sync bf16", w), run_hmma<1>("mma.sync f16", w);
  }
  return 0;
}
