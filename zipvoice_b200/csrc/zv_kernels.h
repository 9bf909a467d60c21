// Launchers of the bandwidth-bound / small kernels and of the fused attention.
// All tensors are token-major: [B][L][C] with C contiguous.  Return 0 or a cudaError_t.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace zv {

// x,text,speech fp32 [B][T][F] -> bf16 [nb][T][ldo] = [x | text | speech | 0-pad]  (zipvoice.py:163).
// cfg != 0 doubles the batch (solver.py:83-98): rows [0,B) unconditional (text = 0, speech = 0 when
// zero_speech_uncond), rows [B,2B) conditional.  Also writes the doubled key-padding mask.
int pack_input(const float* x, const float* text, const float* speech, const uint8_t* mask, int B, int T, int Fx,
               int Ft, int cfg, int zero_speech_uncond, __nv_bfloat16* out, int ldo, uint8_t* mask_out, cudaStream_t s);

// x += dt * v  (plain, nb == B)   or   x += dt * ((1+g_b) v[B+b] - g_b v[b])  (cfg; solver.py:100-110,239)
int cfg_euler_update(float* x, const float* v, const float* guidance, int guidance_len, float g_mult, int B, int T,
                     int F, int cfg, float dt, float* v_out, cudaStream_t s);

// Depthwise conv over time ("same" zero padding per sequence) + bias + SwooshR (zipformer.py:1672-1678).
int dwconv_swooshr(const __nv_bfloat16* in, const float* w, const float* bias, int nb, int L, int C, int k,
                   __nv_bfloat16* out, int ld_out, cudaStream_t s);

// The next three read and write the residual stream in its s24 storage format (zv_common.cuh): a bf16 `hi` plane
// (the GEMM A operand) and a byte `lo` plane of the same [rows][D] shape.
// out = orig + (BiasNorm(src) - orig) * scale   (scaling.py:345-348, zipformer.py:635-637); out may alias orig.
int biasnorm_bypass(const __nv_bfloat16* src_hi, const uint8_t* src_lo, const __nv_bfloat16* orig_hi,
                    const uint8_t* orig_lo, const float* nbias, float log_scale, const float* bscale, int M, int D,
                    __nv_bfloat16* out_hi, uint8_t* out_lo, cudaStream_t s);

// SimpleDownsample (zipformer.py:887-913): out[b][j] = sum_k w[k] * src[b][min(j*ds+k, L-1)]
int downsample(const __nv_bfloat16* src_hi, const uint8_t* src_lo, const float* w, int nb, int L, int D, int ds,
               __nv_bfloat16* out_hi, uint8_t* out_lo, uint8_t* mask_out, const uint8_t* mask_in, cudaStream_t s);

// SimpleUpsample + out_combiner (zipformer.py:866-870): x = orig + (y[t/ds] - orig) * scale, in place on orig.
int upsample_combine(__nv_bfloat16* orig_hi, uint8_t* orig_lo, const __nv_bfloat16* y_hi, const uint8_t* y_lo,
                     const float* bscale, int nb, int L, int L2, int D, int ds, cudaStream_t s);

int f32_to_bf16(const float* in, __nv_bfloat16* out, size_t n, cudaStream_t s);

// Time embedding (zipformer.py:47-69,267-278): a[b] = SwooshR(time_embed(sincos(t_b) [+ Wg sincos(g_b)])).
// t value of sequence b is t[b * t_stride] (stride 0 = one scalar for all), same for the guidance g.
int time_embed(const float* t, int t_stride, const float* g, int g_stride, int nb, int dim, const float* w0, const float* b0, const float* w2,
               const float* b2, const float* wg, float* a_out, cudaStream_t s);
// out[b][n] = W[n][:] . a[b][:] + bias[n]   (fp32, W [N][K]); one warp per n.
int small_linear(const float* a, const float* W, const float* bias, int nb, int N, int K, float* out, cudaStream_t s);
// C[M][N] = A[M][K] B[K][N] fp32 (finalize-time weight folding only).
int sgemm_nn(const float* A, const float* B, float* C, int M, int N, int K, cudaStream_t s);

// Projected relative-position table E[h][r][dp] = sum_c Wpos[h*dp+?][c] pe[r][c], r in [0, 2L-1)
// (zipformer.py:983-1056,1215-1219).  W [H*dp][pos_dim].
int pos_table(const float* W, int H, int dp, int pos_dim, int L, float* E, cudaStream_t s);
// fp16 copies of E for the tensor-core position term of the SelfAttention kernel: out = [2][H][stride][4] halves
// (2 * H * stride * 8 bytes), stride = pos_table_f16_stride(L); see zv_time.cu.
static constexpr int POS16_PADL = 128;
int pos_table_f16_stride(int L);
int pos_table_f16(const float* E, int H, int L, void* out, cudaStream_t s);

// The scores arrive in LOG2 units: the producer scales q (the rows of the attention in_proj that make q) and the
// relative-position table E by log2(e), so that softmax = exp2(S - max) without a multiply per score and the
// reference of a speculative tile can be folded into the MMA accumulator initialisation.  Key padding = ATTN_MASKED.
static constexpr float ATTN_LOG2E = 1.4426950408889634f;
static constexpr float ATTN_MASKED = -1000.0f * 1.4426950408889634f;  // masked_fill(-1000) of zipformer.py:1286-1289
struct AttnParams {
  const __nv_bfloat16* qkp;  // IEEE fp16 bit patterns (not bf16) [nb][L][ldq]: q at h*dq, k at H*dq + h*dq, p at 2*H*dq + h*dp
  int ldq;
  const float* E;            // [H][2L-1][4], times log2(e)
  const void* E16 = nullptr; // pos_table_f16 of E (SelfAttention, tensor-core position term); null: CUDA-core kernel
  int e16_stride = 0;        // entries per head of E16
  // Key split of the SelfAttention kernel for grids smaller than the machine (single utterances, the 60 s dialog):
  // ksplit CTAs share a query tile, each takes a contiguous range of key tiles and writes its unnormalised state
  // (16 fp32 columns of O -- column dv is the row sum -- and the row's reference maximum) to `part`; a merge
  // kernel combines them.  part: [nb][H][ksplit][L][17] floats; 0 / null: chosen by the launcher from part_cap.
  float* part = nullptr;
  size_t part_cap = 0;       // bytes available at `part`
  int ksplit = 0;            // 0: let the launcher choose; tests force a value
  const uint8_t* mask;       // [nb][L] 1 = padded key (may be null)
  const __nv_bfloat16* V;    // [nb][L][ldv]
  int ldv;
  int vhs = 0;               // column stride between the heads of V (0 = dv; SelfAttention uses 16: 12 + 4 zeros)
  const __nv_bfloat16* gate; // optional [nb][L][ldg]: out *= gate (NonlinAttention y)
  int ldg;
  __nv_bfloat16* out;        // [nb][L][ldo]
  int ldo;
  int nb, L, H;
  int heads;                 // heads processed: H (SelfAttention) or 1 (NonlinAttention, head 0)
  int dv;                    // value width per head (12 or hidden)
  int dbg = 0;               // profiling experiments only: bit 0 skip value TMA loads, bit 1 skip the P.X MMAs,
                             // bit 2 force the exact-maxima repeat of the NonlinAttention kernel
};
int attention(const AttnParams& p, cudaStream_t s);

}  // namespace zv
