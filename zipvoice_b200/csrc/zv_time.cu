// Small fp32 kernels: time / guidance embedding MLP, per-batch bias projections, the projected
// relative-position table, and an fp32 matmul used only when weights are folded at finalize time.
#include "zv_common.cuh"
#include <cuda_fp16.h>

#include "zv_kernels.h"

namespace zv {

// One block per batch element.  dim = time_embed_dim (192): emb = [cos(t f_k) | sin(t f_k)]
// (+ Wg [cos(g f_k) | sin(g f_k)]), h = SwooshR(W0 emb + b0) (2*dim), e = W2 h + b2 (dim),
// a = SwooshR(e)  -- the SwooshR that opens every stack's time_emb module (zipformer.py:677-680).
__global__ void time_embed_kernel(const float* __restrict__ t, int t_stride, const float* __restrict__ g,
                                  int g_stride, int dim,
                                  const float* __restrict__ w0, const float* __restrict__ b0,
                                  const float* __restrict__ w2, const float* __restrict__ b2,
                                  const float* __restrict__ wg, float* __restrict__ a_out) {
  extern __shared__ float sm[];
  float* emb = sm;             // dim
  float* gemb = sm + dim;      // dim
  float* h = sm + 2 * dim;     // 2*dim
  const int b = blockIdx.x;
  const int half = dim / 2;
  const float tv = t[(size_t)b * t_stride];
  const float gv = g ? g[(size_t)b * g_stride] : 0.f;
  for (int i = threadIdx.x; i < half; i += blockDim.x) {
    const float f = expf(-logf(10000.0f) * (float)i / (float)half);
    emb[i] = cosf(tv * f);
    emb[half + i] = sinf(tv * f);
    gemb[i] = cosf(gv * f);
    gemb[half + i] = sinf(gv * f);
  }
  __syncthreads();
  // y[i] = W[i][:] . x: one warp per output row, 4 rows in flight per warp (coalesced row reads, shuffle sum)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  auto matvec = [&](const float* __restrict__ W, const float* x, int rows, int K, auto&& store) {
    for (int i0 = warp * 4; i0 < rows; i0 += nwarps * 4) {
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      for (int k = lane; k < K; k += 32) {
        const float xv = x[k];
#pragma unroll
        for (int r = 0; r < 4; ++r)
          if (i0 + r < rows) acc[r] = fmaf(W[(size_t)(i0 + r) * K + k], xv, acc[r]);
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float v = warp_sum(acc[r]);
        if (lane == 0 && i0 + r < rows) store(i0 + r, v);
      }
    }
  };
  if (g) {
    matvec(wg, gemb, dim, dim, [&](int i, float v) { h[i] = v; });  // temporary
    __syncthreads();
    for (int i = threadIdx.x; i < dim; i += blockDim.x) emb[i] += h[i];
    __syncthreads();
  }
  matvec(w0, emb, 2 * dim, dim, [&](int i, float v) { h[i] = swoosh_r(v + b0[i]); });
  __syncthreads();
  matvec(w2, h, dim, 2 * dim, [&](int i, float v) { a_out[(size_t)b * dim + i] = swoosh_r(v + b2[i]); });
}

int time_embed(const float* t, int t_stride, const float* g, int g_stride, int nb, int dim, const float* w0, const float* b0, const float* w2,
               const float* b2, const float* wg, float* a_out, cudaStream_t s) {
  time_embed_kernel<<<nb, 512, 4 * dim * sizeof(float), s>>>(t, t_stride, g, g_stride, dim, w0, b0, w2, b2, wg, a_out);
  return (int)cudaGetLastError();
}

// out[b][n] = W[n][:] . a[b][:] + bias[n].  One warp per n, the W row is read once and reused for all b.
__global__ void small_linear_kernel(const float* __restrict__ a, const float* __restrict__ W,
                                    const float* __restrict__ bias, int nb, int N, int K, float* __restrict__ out) {
  const int n = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (n >= N) return;
  float w[8];  // K <= 256
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int k = i * 32 + lane;
    w[i] = k < K ? W[(size_t)n * K + k] : 0.f;
  }
  const float bv = bias ? bias[n] : 0.f;
  for (int b = 0; b < nb; ++b) {
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int k = i * 32 + lane;
      if (k < K) acc += w[i] * a[(size_t)b * K + k];
    }
    acc = warp_sum(acc);
    if (lane == 0) out[(size_t)b * N + n] = acc + bv;
  }
}

int small_linear(const float* a, const float* W, const float* bias, int nb, int N, int K, float* out,
                 cudaStream_t s) {
  if (K > 256) return 920;
  const int warps = 8;
  small_linear_kernel<<<(N + warps - 1) / warps, warps * 32, 0, s>>>(a, W, bias, nb, N, K, out);
  return (int)cudaGetLastError();
}

// C = A B, fp32, 16x16 tiles.  Finalize-time only (weight folding), not on the sampling path.
__global__ void sgemm_nn_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ C,
                                int M, int N, int K) {
  __shared__ float sa[16][17], sb[16][17];
  const int row = blockIdx.y * 16 + threadIdx.y, col = blockIdx.x * 16 + threadIdx.x;
  float acc = 0.f;
  for (int k0 = 0; k0 < K; k0 += 16) {
    sa[threadIdx.y][threadIdx.x] = (row < M && k0 + threadIdx.x < K) ? A[(size_t)row * K + k0 + threadIdx.x] : 0.f;
    sb[threadIdx.y][threadIdx.x] = (col < N && k0 + threadIdx.y < K) ? B[(size_t)(k0 + threadIdx.y) * N + col] : 0.f;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) acc += sa[threadIdx.y][k] * sb[k][threadIdx.x];
    __syncthreads();
  }
  if (row < M && col < N) C[(size_t)row * N + col] = acc;
}
int sgemm_nn(const float* A, const float* B, float* C, int M, int N, int K, cudaStream_t s) {
  dim3 grid((N + 15) / 16, (M + 15) / 16), block(16, 16);
  sgemm_nn_kernel<<<grid, block, 0, s>>>(A, B, C, M, N, K);
  return (int)cudaGetLastError();
}

// E[h][r][j] = sum_c W[h*dp + j][c] * pe[r][c]; pe per zipformer.py:995-1030 (offset x = r-(L-1)).
__global__ void pos_table_kernel(const float* __restrict__ W, int H, int dp, int pos_dim, int L,
                                 float* __restrict__ E) {
  extern __shared__ float sw[];  // H*dp*pos_dim
  for (int i = threadIdx.x; i < H * dp * pos_dim; i += blockDim.x) sw[i] = W[i];
  __syncthreads();
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  const int R = 2 * L - 1;
  if (r >= R) return;
  const float x = (float)(r - (L - 1));
  const float c = sqrtf((float)pos_dim);
  const float sgn = x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f);
  const float xc = c * sgn * (logf(fabsf(x) + c) - logf(c));
  const float length_scale = (float)pos_dim / (2.0f * 3.14159265358979323846f);
  const float xa = atanf(xc / length_scale);
  float acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = 0.f;
  for (int k = 0; k < pos_dim / 2; ++k) {
    const float ang = xa * (float)(k + 1);
    const float cs = cosf(ang);
    const float sn = (2 * k + 1 == pos_dim - 1) ? 1.0f : sinf(ang);  // last column overwritten by 1 (bias)
    for (int o = 0; o < H * dp; ++o)
      acc[o] += sw[o * pos_dim + 2 * k] * cs + sw[o * pos_dim + 2 * k + 1] * sn;
  }
  for (int h = 0; h < H; ++h)
    for (int j = 0; j < dp; ++j) E[((size_t)h * R + r) * dp + j] = acc[h * dp + j];
}
// fp16 copies of E for the tensor-core position term of the SelfAttention kernel (zv_attn_sa.cu, attn_sa2_kernel):
// out = [2][H][stride][4] halves, stride = pos_table_f16_stride(L) entries per head.  Copy 0 holds POS16_PADL zero
// entries, then the R = 2L - 1 table rows, then zeros (every 200-entry window a key tile can ask for lies inside, so
// the producer issues whole, unclipped bulk copies); copy 1 is copy 0 shifted by one entry (copy1[x] = copy0[x + 1]):
// the kernel needs windows that start at odd entries too, and a bulk copy's source must be 16-byte aligned.
int pos_table_f16_stride(int L) { return 2 * L + 208; }
__global__ void pos_table_f16_kernel(const float* __restrict__ E, int H, int L, int stride, __half* __restrict__ out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, h = blockIdx.y, copy = blockIdx.z;
  if (x >= stride) return;
  const int R = 2 * L - 1;
  const int r = x + copy - POS16_PADL;
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (r >= 0 && r < R) v = reinterpret_cast<const float4*>(E)[(size_t)h * R + r];
  __half2* o = reinterpret_cast<__half2*>(out + (((size_t)copy * H + h) * stride + x) * 4);
  o[0] = __floats2half2_rn(v.x, v.y), o[1] = __floats2half2_rn(v.z, v.w);
}
int pos_table_f16(const float* E, int H, int L, void* out, cudaStream_t s) {
  const int stride = pos_table_f16_stride(L);
  pos_table_f16_kernel<<<dim3((stride + 127) / 128, H, 2), 128, 0, s>>>(E, H, L, stride, static_cast<__half*>(out));
  return (int)cudaGetLastError();
}

int pos_table(const float* W, int H, int dp, int pos_dim, int L, float* E, cudaStream_t s) {
  if (H * dp > 16) return 921;
  const int R = 2 * L - 1;
  pos_table_kernel<<<(R + 127) / 128, 128, H * dp * pos_dim * sizeof(float), s>>>(W, H, dp, pos_dim, L, E);
  return (int)cudaGetLastError();
}

}  // namespace zv
