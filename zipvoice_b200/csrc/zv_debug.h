// PRIVATE test / bench hooks of libzipvoice_b200.so -- not part of the reference-facing C ABI
// (include/zipvoice_b200.h).  Bound by zipvoice_b200/_lib.py (DEBUG_EXPORTS) for tests/ and tools/ only.
#pragma once
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

int zv_debug_gemm(int mode, int impl, int M, int N, int K, int L, int gate_cols, const float* A_f32,
                  const float* W_f32, const float* bias, const float* rowbias, const float* res, const float* resbias,
                  const float* orig, const float* bypass_scale, const uint8_t* mask, float* out_f32, float* out2_f32,
                  void* stream);
int zv_debug_attention(int nb, int L, int H, int heads, int dv, const float* qkp_f32, int ldq, const float* E,
                       const uint8_t* mask, const float* V_f32, int ldv, const float* gate_f32, float* out_f32,
                       void* stream);
int zv_debug_set_attn_impl(int impl);
/* force the key split of the SelfAttention kernel in zv_debug_attention (0 = let the launcher choose) */
int zv_debug_attention_ksplit(int ksplit);
int zv_debug_attention_bench(int nb, int L, int H, int heads, int dv, int use_mask, int dbg, int iters, float* ms_out);
int zv_debug_gemm_bench(int mode, int impl, int M, int N, int K, int L, int gate_cols, int block_n, int iters,
                        float* ms_out);
int zv_debug_pos_table(const float* W, int H, int dp, int pos_dim, int L, float* E_out, void* stream);

#ifdef __cplusplus
}
#endif
