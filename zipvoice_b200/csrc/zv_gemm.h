// GEMM interface: C[m][n] = epilogue( sum_k A[m][k] * W[n][k] ), A and W bf16 (K contiguous),
// fp32 accumulation in TMEM (tcgen05.mma), fused epilogues for the Zipformer layer.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace zv {

enum EpiMode : int {
  EPI_BIAS_BF16 = 0,  // out_bf16 = acc + bias (+ rowbias[b])
  EPI_SWOOSHL_BF16 = 1,  // out_bf16 = SwooshL(acc + bias + rowbias[b])            (FF in_proj)
                         // with gate_cols > 0 (multiple of 32): columns >= gate_cols are a second projection of the
                         // same rows, out2 = fp16(acc + bias), saturating  (attention q | k | p beside FF1's in_proj)
  EPI_GLU_BF16 = 2,      // cols come in groups of 32 = [16 x | 16 s]; out = x*sigmoid(s), masked frames -> 0
  EPI_NONLIN_BF16 = 3,   // cols < gate_cols: groups of 32 = [16 s | 16 x] -> out = x*tanh(s);
                         // cols >= gate_cols: out2 = acc + bias                     (NonlinAttention in_proj)
  EPI_RESIDUAL = 4,      // v = res + acc + bias (+ resbias[b]); optional bypass vs orig; writes fp32 + bf16
  EPI_F32 = 5,           // out_f32 = acc + bias (+ rowbias[b])
  EPI_GELU_BF16 = 6,     // out_bf16 = GELU_erf(acc + bias)                          (Vocos ConvNeXt pwconv1)
  EPI_RES24 = 7,         // EPI_RESIDUAL on the s24 residual stream (zv_common.cuh): v = res + acc + bias (+ resbias[b]),
                         // optional bypass vs orig; res may be absent (stack input).  Writes the hi and lo planes.
};

struct EpiParams {
  int mode = 0;
  int M = 0;          // valid rows
  int N = 0;          // valid accumulator columns
  int L = 1;          // rows per batch element (b = m / L)
  int gate_cols = 0;  // EPI_NONLIN_BF16: first accumulator column of the plain (y) part
  const float* bias = nullptr;     // [N]   accumulator-column order
  const float* rowbias = nullptr;  // [B][bias_ld] optional, added before the activation
  int bias_ld = 0;                 // row pitch of rowbias / resbias (per batch element)
  __nv_bfloat16* out_bf16 = nullptr;
  int ld_bf16 = 0;
  int out_f16 = 0;    // EPI_BIAS_BF16 only: store IEEE fp16 instead of bf16 (attention q / k / p: 11-bit mantissa)
  __nv_bfloat16* out2_bf16 = nullptr;
  int ld2_bf16 = 0;
  float* out_f32 = nullptr;
  int ld_f32 = 0;
  const float* res = nullptr;  // EPI_RESIDUAL input (may alias out_f32), same ld as out_f32
  const float* resbias = nullptr;  // [B][bias_ld] optional (time embedding folded into the residual)
  const float* orig = nullptr;     // optional bypass: v = orig + (v - orig) * bypass_scale[n]
  const float* bypass_scale = nullptr;
  // EPI_RES24: s24 planes, all with row pitch ld_bf16 (elements); the output hi plane is out_bf16
  const __nv_bfloat16* res_hi = nullptr;   // may alias out_bf16 / out_lo (in-place update)
  const uint8_t* res_lo = nullptr;
  const __nv_bfloat16* orig_hi = nullptr;  // with bypass_scale, like orig
  const uint8_t* orig_lo = nullptr;
  uint8_t* out_lo = nullptr;
  const uint8_t* mask = nullptr;   // [B][L] 1 = padded frame (EPI_GLU_BF16)
  int f32_direct = 0;              // EPI_F32: set by the launcher when the output cannot be a TMA tensor (pitch % 4)
};

struct GemmPlan {
  int M, N_pad, K;      // N_pad = rows of the packed weight (multiple of block_n)
  int lda;              // A row pitch in elements (multiple of 8)
  int ldw;              // W row pitch in elements (multiple of 8, >= K)
  int block_n;          // multiple of 32, <= 256
  // Optional second A operand: K columns >= k_split (multiple of 64) come from A2 (row pitch lda2), so that
  // [x | y] @ W^T runs without concatenating x and y in memory.
  const __nv_bfloat16* A2 = nullptr;
  int lda2 = 0, k_split = 0;
};

// Launch on `stream`.  impl: 0 = tcgen05/TMA kernel, 1 = SIMT checker kernel (tests / debugging only).
// Returns cudaError_t-like code (0 = ok).
int gemm_launch(const __nv_bfloat16* A, const __nv_bfloat16* W, const GemmPlan& plan, const EpiParams& ep,
                int impl, cudaStream_t stream);

int gemm_pick_block_n(int n_valid, int mode, int gate_cols);

// Cached cuTensorMapEncodeTiled for a row-major [rows][cols] tensor with row pitch ld (elements):
// dtype 0 = 16-bit, 1 = fp32, 2 = bytes; box [box_rows][box_cols]; swizzle 0 (none) / 32 / 64 / 128 bytes.  `tm` is a CUtensorMap*.
int make_tmap(void* tm, const void* ptr, int dtype, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
              uint32_t box_cols, int swizzle);

// 16-bit tensor [n2][n1][n0] (n0 contiguous, pitches ld1 / ld2 in elements), box [1][box1][box0], no swizzle,
// zero fill outside the tensor.  Not cached by the GEMM map cache key: separate small cache.
int make_tmap3d_16(void* tm, const void* ptr, uint64_t n0, uint64_t n1, uint64_t n2, uint64_t ld1, uint64_t ld2,
                   uint32_t box0, uint32_t box1);

}  // namespace zv
