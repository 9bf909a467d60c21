// tcgen05 / TMA GEMM with fused Zipformer epilogues (sm_100a).
//   warp 0      : TMA producer (one lane) -- A tile 128x64, W tile block_n x 64, SWIZZLE_128B
//   warp 1      : MMA issuer (one lane)   -- tcgen05.mma kind::f16, M=128, N=block_n, K=16 x4 per stage
//   warp 2      : TMEM allocator (512 columns = 2 accumulator stages of <=256 fp32 columns)
//   warps 4..11 : epilogue, two column groups x four TMEM lane quarters.  A group walks the 32-column
//                 chunks of the accumulator tile (chunk c -> group c % 2): tcgen05.ld 32 lanes x 32 columns
//                 -> registers -> fused math (MUFU ex2/lg2/tanh/rcp) -> swizzled shared-memory staging ->
//                 one elected thread issues a TMA store (rows / columns outside the tensor are clipped by
//                 the tensor map).  EPI_RESIDUAL also TMA-loads the fp32 residual chunk one chunk ahead and
//                 updates it in place in shared memory.
// Persistent: grid = min(#tiles, #SMs); tile t -> (m_blk = t / n_tiles, n_blk = t % n_tiles); the tiles of a partial
// last wave are cut into narrower column slices so that every CTA gets one (TileSched).
#include <cuda.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <unordered_map>

#include "zv_common.cuh"
#include "zv_gemm.h"

namespace zv {

static constexpr int BM = 128;
static constexpr int BK = 64;
static constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KB
static constexpr int ACC_STAGE_COLS = 256;

// One row, 32 consecutive accumulator columns starting at n0 (n0 % 32 == 0).
template <int MODE>
ZV_DEVINL void epi_apply(const EpiParams& ep, int m, int b, int n0, const float (&acc)[32]) {
  if (m >= ep.M || n0 >= ep.N) return;
  const int nvalid = min(32, ep.N - n0);
  float v[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    float x = acc[i];
    if (i < nvalid) {
      x += __ldg(ep.bias + n0 + i);
      if (MODE != EPI_RESIDUAL && MODE != EPI_RES24 && ep.rowbias &&
          !(MODE == EPI_SWOOSHL_BF16 && ep.gate_cols > 0 && n0 >= ep.gate_cols))
        x += __ldg(ep.rowbias + (size_t)b * ep.bias_ld + n0 + i);
    }
    v[i] = x;
  }
  if (MODE == EPI_SWOOSHL_BF16 && ep.gate_cols > 0 && n0 >= ep.gate_cols) {
    // plain fp16 tail of a split launch (see the tcgen05 epilogue): no row bias, no activation
    __half* oh = reinterpret_cast<__half*>(ep.out2_bf16) + (size_t)m * ep.ld2_bf16 + (n0 - ep.gate_cols);
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (i < nvalid) oh[i] = f2h_sat(acc[i] + __ldg(ep.bias + n0 + i));
  } else if (MODE == EPI_BIAS_BF16 || MODE == EPI_SWOOSHL_BF16 || MODE == EPI_GELU_BF16) {
    __nv_bfloat16* o = ep.out_bf16 + (size_t)m * ep.ld_bf16 + n0;
    if (MODE == EPI_SWOOSHL_BF16) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = swoosh_l(v[i]);
    }
    if (MODE == EPI_GELU_BF16) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = gelu_erf(v[i]);
    }
    if (MODE == EPI_BIAS_BF16 && ep.out_f16) {
      __half* oh = reinterpret_cast<__half*>(o);
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i < nvalid) oh[i] = f2h_sat(v[i]);
    } else if (nvalid == 32) {
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        uint4 p = make_uint4(pack_bf16x2(v[i], v[i + 1]), pack_bf16x2(v[i + 2], v[i + 3]),
                             pack_bf16x2(v[i + 4], v[i + 5]), pack_bf16x2(v[i + 6], v[i + 7]));
        *reinterpret_cast<uint4*>(o + i) = p;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i < nvalid) o[i] = __float2bfloat16(v[i]);
    }
  } else if (MODE == EPI_GLU_BF16) {
    // [16 x | 16 s] -> 16 outputs at column n0/2
    const bool masked = ep.mask && ep.mask[(size_t)b * ep.L + (m - b * ep.L)];
    float o16[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) o16[i] = masked ? 0.0f : v[i] * sigmoidf_(v[16 + i]);
    __nv_bfloat16* o = ep.out_bf16 + (size_t)m * ep.ld_bf16 + (n0 >> 1);
#pragma unroll
    for (int i = 0; i < 16; i += 8) {
      uint4 p = make_uint4(pack_bf16x2(o16[i], o16[i + 1]), pack_bf16x2(o16[i + 2], o16[i + 3]),
                           pack_bf16x2(o16[i + 4], o16[i + 5]), pack_bf16x2(o16[i + 6], o16[i + 7]));
      *reinterpret_cast<uint4*>(o + i) = p;
    }
  } else if (MODE == EPI_NONLIN_BF16) {
    if (n0 < ep.gate_cols) {
      float o16[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) o16[i] = v[16 + i] * tanhf_(v[i]);
      __nv_bfloat16* o = ep.out_bf16 + (size_t)m * ep.ld_bf16 + (n0 >> 1);
#pragma unroll
      for (int i = 0; i < 16; i += 8) {
        uint4 p = make_uint4(pack_bf16x2(o16[i], o16[i + 1]), pack_bf16x2(o16[i + 2], o16[i + 3]),
                             pack_bf16x2(o16[i + 4], o16[i + 5]), pack_bf16x2(o16[i + 6], o16[i + 7]));
        *reinterpret_cast<uint4*>(o + i) = p;
      }
    } else {
      __nv_bfloat16* o = ep.out2_bf16 + (size_t)m * ep.ld2_bf16 + (n0 - ep.gate_cols);
      if (nvalid == 32) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 p = make_uint4(pack_bf16x2(v[i], v[i + 1]), pack_bf16x2(v[i + 2], v[i + 3]),
                               pack_bf16x2(v[i + 4], v[i + 5]), pack_bf16x2(v[i + 6], v[i + 7]));
          *reinterpret_cast<uint4*>(o + i) = p;
        }
      } else {
  #pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i < nvalid) o[i] = __float2bfloat16(v[i]);
      }
    }
  } else if (MODE == EPI_RESIDUAL) {
    const float* r = ep.res + (size_t)m * ep.ld_f32 + n0;
    float* o = ep.out_f32 + (size_t)m * ep.ld_f32 + n0;
    __nv_bfloat16* ob = ep.out_bf16 ? ep.out_bf16 + (size_t)m * ep.ld_bf16 + n0 : nullptr;
    if (nvalid == 32) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        float4 rr = *reinterpret_cast<const float4*>(r + i);
        v[i] += rr.x, v[i + 1] += rr.y, v[i + 2] += rr.z, v[i + 3] += rr.w;
      }
      if (ep.resbias) {
        const float* rb = ep.resbias + (size_t)b * ep.bias_ld + n0;
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] += __ldg(rb + i);
      }
      if (ep.orig) {
        const float* og = ep.orig + (size_t)m * ep.ld_f32 + n0;
#pragma unroll
        for (int i = 0; i < 32; i += 4) {
          float4 oo = *reinterpret_cast<const float4*>(og + i);
          v[i] = oo.x + (v[i] - oo.x) * __ldg(ep.bypass_scale + n0 + i);
          v[i + 1] = oo.y + (v[i + 1] - oo.y) * __ldg(ep.bypass_scale + n0 + i + 1);
          v[i + 2] = oo.z + (v[i + 2] - oo.z) * __ldg(ep.bypass_scale + n0 + i + 2);
          v[i + 3] = oo.w + (v[i + 3] - oo.w) * __ldg(ep.bypass_scale + n0 + i + 3);
        }
      }
#pragma unroll
      for (int i = 0; i < 32; i += 4)
        *reinterpret_cast<float4*>(o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
      if (ob) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 p = make_uint4(pack_bf16x2(v[i], v[i + 1]), pack_bf16x2(v[i + 2], v[i + 3]),
                               pack_bf16x2(v[i + 4], v[i + 5]), pack_bf16x2(v[i + 6], v[i + 7]));
          *reinterpret_cast<uint4*>(ob + i) = p;
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        if (i >= nvalid) continue;
        float x = v[i] + r[i];
        if (ep.resbias) x += ep.resbias[(size_t)b * ep.bias_ld + n0 + i];
        if (ep.orig) {
          float oo = ep.orig[(size_t)m * ep.ld_f32 + n0 + i];
          x = oo + (x - oo) * ep.bypass_scale[n0 + i];
        }
        o[i] = x;
        if (ob) ob[i] = __float2bfloat16(x);
      }
    }
  } else if (MODE == EPI_RES24) {
    // checker path: element by element through the scalar forms of the s24 format
    const size_t off = (size_t)m * ep.ld_bf16 + n0;
    const uint16_t* rh = reinterpret_cast<const uint16_t*>(ep.res_hi);
    const uint16_t* oh = reinterpret_cast<const uint16_t*>(ep.orig_hi);
    uint16_t* wh = reinterpret_cast<uint16_t*>(ep.out_bf16);
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      if (i >= nvalid) continue;
      float x = v[i];
      if (rh) x += s24_dec(rh[off + i], ep.res_lo[off + i]);
      if (ep.resbias) x += ep.resbias[(size_t)b * ep.bias_ld + n0 + i];
      if (oh) {
        const float oo = s24_dec(oh[off + i], ep.orig_lo[off + i]);
        x = oo + (x - oo) * ep.bypass_scale[n0 + i];
      }
      const uint32_t w = s24_w(x);
      wh[off + i] = (uint16_t)(w >> 16);
      ep.out_lo[off + i] = (uint8_t)(w >> 8);
    }
  } else {  // EPI_F32
    float* o = ep.out_f32 + (size_t)m * ep.ld_f32 + n0;
    if (nvalid == 32 && (ep.ld_f32 & 3) == 0) {
#pragma unroll
      for (int i = 0; i < 32; i += 4)
        *reinterpret_cast<float4*>(o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i < nvalid) o[i] = v[i];
    }
    if (ep.out_bf16) {  // optional bf16 shadow (stack input)
      __nv_bfloat16* ob = ep.out_bf16 + (size_t)m * ep.ld_bf16 + n0;
#pragma unroll
      for (int i = 0; i < 32; i += 2)
        if (i < nvalid) *reinterpret_cast<uint32_t*>(ob + i) = pack_bf16x2(v[i], v[i + 1]);
    }
  }
}

// Epilogue column groups (4 warps each, one per TMEM lane quarter).  The modes whose epilogue is
// transcendental-heavy (SwooshL / GELU / gates) are bound by the epilogue's issue rate when K is
// short, so they get 4 groups = 4 warps per scheduler; the HBM-bound modes keep 2.
static constexpr int MAX_NCG = 4;
__host__ __device__ constexpr int ncg_for(int mode) {
  return (mode == EPI_SWOOSHL_BF16 || mode == EPI_GELU_BF16 || mode == EPI_GLU_BF16) ? 4 : 2;
}
static constexpr int STG_BF16_BYTES = BM * 64;   // 128 rows x 32 bf16
static constexpr int STG_F32_BYTES = BM * 128;   // 128 rows x 32 fp32
static constexpr int STG_LO_BYTES = BM * 32;     // 128 rows x 32 bytes (lo plane of the s24 stream)
static constexpr int STG_S24_BYTES = STG_BF16_BYTES + STG_LO_BYTES;  // one s24 staging buffer: hi rows, then lo rows
// s24 staging buffers per column group: while chunk n is updated the residual rows of chunks n + 1 .. n + S24_BUFS - 1
// are in flight.  Measured on B200 (tools/res24_bench.py, M = 156 k): 3 or 4 buffers change nothing for K = 384 / 576
// (121 us either way: the residual prefetch is not what holds the short-K shapes at 0.77 of the HBM roofline) and
// cost the K >= 1152 shapes a stage of the operand ring (181 -> 200 us at K = 1536), so two it is.
#ifndef ZV_S24_BUFS
#define ZV_S24_BUFS 2
#endif
static constexpr int S24_BUFS = ZV_S24_BUFS;

struct GemmMaps {
  CUtensorMap A, W;
  CUtensorMap Wt;  // W with a box of block_n / split rows: the narrow units of the last wave (TileSched)
  CUtensorMap A2;  // optional second A operand for k-blocks >= kb_split
  CUtensorMap O0;  // bf16 output (box 32 cols, SWIZZLE_64B; GLU / gated NONLIN: box 16 cols, SWIZZLE_32B)
  CUtensorMap O1;  // EPI_NONLIN plain part: bf16 (box 32, SWIZZLE_64B); EPI_RESIDUAL: fp32 out (box 32, SWIZZLE_128B)
  CUtensorMap R;   // EPI_RESIDUAL: fp32 residual input (box 32, SWIZZLE_128B)
  // EPI_RES24: O0 / O1 = hi / lo plane of the output (bf16 box 32, SWIZZLE_64B / byte box 32, SWIZZLE_32B),
  //            R / R2 = hi / lo plane of the residual input, same boxes
  CUtensorMap R2;
};

// Work schedule of one launch.  Units [0, main_units) are full tiles (128 x block_n; 256 x block_n for CTA pairs)
// in the order unit -> (unit / n_tiles, unit % n_tiles).  A tile count that is not a multiple of the CTA count
// would leave most SMs idle for one whole tile time at the end, so the launcher cuts the remaining tiles of the
// last wave into `split` column slices each: units >= main_units are 128 x (block_n / split) slices, again one
// per CTA.  split = 1: no narrow units.
struct TileSched {
  int main_units, total_units, split;
};

static constexpr int MAX_STAGES = 10;
struct GemmSmemTail {
  uint64_t full[MAX_STAGES];
  uint64_t empty[MAX_STAGES];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint64_t rbar[MAX_NCG * 4][4];  // per epilogue warp (EPI_RESIDUAL: 2 in use, EPI_RES24: S24_BUFS)
  uint32_t tmem_base;
};

// bytes of epilogue staging per column group
__host__ __device__ constexpr int stg_group_bytes(int mode) {
  return (mode == EPI_RESIDUAL || mode == EPI_F32) ? 2 * STG_F32_BYTES + STG_BF16_BYTES
         : mode == EPI_RES24                        ? S24_BUFS * STG_S24_BYTES
                                                    : 2 * STG_BF16_BYTES;
}

// 32 fp32 values of one row -> 32 bf16 in a 64-byte staging row (SWIZZLE_64B pattern of the TMA store)
ZV_DEVINL void stage_row_bf16x32(uint8_t* stg, int r, const float (&v)[32]) {
  const int x = (r >> 1) & 3;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint4 p = make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                         pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
    *reinterpret_cast<uint4*>(stg + r * 64 + ((j ^ x) << 4)) = p;
  }
}
ZV_DEVINL void stage_row_f16x32(uint8_t* stg, int r, const float (&v)[32]) {
  const int x = (r >> 1) & 3;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint4 p = make_uint4(pack_f16x2_sat(v[8 * j], v[8 * j + 1]), pack_f16x2_sat(v[8 * j + 2], v[8 * j + 3]),
                         pack_f16x2_sat(v[8 * j + 4], v[8 * j + 5]), pack_f16x2_sat(v[8 * j + 6], v[8 * j + 7]));
    *reinterpret_cast<uint4*>(stg + r * 64 + ((j ^ x) << 4)) = p;
  }
}
// 16 fp32 values -> 16 bf16 in a 32-byte staging row (SWIZZLE_32B)
ZV_DEVINL void stage_row_bf16x16(uint8_t* stg, int r, const float (&v)[16]) {
  const int x = (r >> 2) & 1;
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    uint4 p = make_uint4(pack_bf16x2(v[8 * j], v[8 * j + 1]), pack_bf16x2(v[8 * j + 2], v[8 * j + 3]),
                         pack_bf16x2(v[8 * j + 4], v[8 * j + 5]), pack_bf16x2(v[8 * j + 6], v[8 * j + 7]));
    *reinterpret_cast<uint4*>(stg + r * 32 + ((j ^ x) << 4)) = p;
  }
}

// acc + bias (+ per-batch row bias) for one 32-column chunk; vector loads when the chunk is full
ZV_DEVINL void add_bias32(float (&v)[32], const float* __restrict__ bias, const float* __restrict__ rowb, int n0,
                          int N) {
  if (n0 + 32 <= N) {
#pragma unroll
    for (int i = 0; i < 32; i += 4) {
      const float4 bb = __ldg(reinterpret_cast<const float4*>(bias + n0 + i));
      f2_unpack(f2_add(f2_pack(v[i], v[i + 1]), f2_pack(bb.x, bb.y)), v[i], v[i + 1]);
      f2_unpack(f2_add(f2_pack(v[i + 2], v[i + 3]), f2_pack(bb.z, bb.w)), v[i + 2], v[i + 3]);
    }
    if (rowb) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        const float4 bb = __ldg(reinterpret_cast<const float4*>(rowb + n0 + i));
        f2_unpack(f2_add(f2_pack(v[i], v[i + 1]), f2_pack(bb.x, bb.y)), v[i], v[i + 1]);
        f2_unpack(f2_add(f2_pack(v[i + 2], v[i + 3]), f2_pack(bb.z, bb.w)), v[i + 2], v[i + 3]);
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i)
      if (n0 + i < N) v[i] += __ldg(bias + n0 + i) + (rowb ? __ldg(rowb + n0 + i) : 0.f);
  }
}

// PAIR: cluster of two CTAs (one TPC) driving `tcgen05.mma.cta_group::2`: the pair owns a 256 x block_n
// tile, each CTA loads its own 128 rows of A and HALF of the W tile, the leader (cluster rank 0) issues
// the MMAs for both, accumulators land in each CTA's own TMEM, each CTA runs the epilogue of its rows.
// Per output element this halves the W bytes pulled from L2, which is what bounds the K = 512 GEMMs.
// DUAL (own instantiation, so that the large-grid kernels keep their register allocation): every CTA owns one
// tile and two threads issue the K loop into two accumulators (see warp 3 below).
template <int MODE, bool PAIR, bool DUAL = false>
__global__ void __launch_bounds__(128 + 128 * ncg_for(MODE), 1)
gemm_tc_kernel(const __grid_constant__ GemmMaps maps, int m_tiles, int n_tiles, int k_blocks, int kb_split,
               int block_n, int stages, EpiParams ep, TileSched sched) {
  constexpr int NCG = ncg_for(MODE);
  constexpr int EPI_WARPS = 4 * NCG;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: [stages x A 16 KB][stages x W block_n*128 B][NCG x epilogue staging][tail]
  // align inside the shared window (pointer + integer keeps the shared address space: LDS/STS, not generic LD/ST)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
  const int w_stage_bytes = (PAIR ? (block_n >> 1) : block_n) * BK * 2;  // W rows this CTA holds per stage
  uint8_t* sA = smem;
  uint8_t* sW = smem + stages * A_STAGE_BYTES;
  uint8_t* sStg = sW + stages * w_stage_bytes;
  GemmSmemTail* tail = reinterpret_cast<GemmSmemTail*>(sStg + NCG * stg_group_bytes(MODE));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // work units: tiles (single CTA) or pair tiles of 256 rows; unit u -> (u / n_tiles, u % n_tiles)
  const int unit0 = PAIR ? (blockIdx.x >> 1) : blockIdx.x;
  const int nunits = PAIR ? (gridDim.x >> 1) : gridDim.x;
  const int total_tiles = sched.total_units;
  const int bn_tail = block_n / sched.split;
  // unit -> first row block, first accumulator column and width of the unit
  auto decode = [&](int t, int& m_blk, int& n0, int& bn) {
    int tile = t, sub = 0;
    bn = block_n;
    if (t >= sched.main_units) {
      const int u = t - sched.main_units;
      tile = sched.main_units + u / sched.split, sub = u % sched.split, bn = bn_tail;
    }
    const int mu = tile / n_tiles;
    m_blk = PAIR ? 2 * mu + (int)rank : mu;
    n0 = (tile - mu * n_tiles) * block_n + sub * bn;
  };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.A);
    tma_prefetch_desc(&maps.W);
    if (sched.split > 1) tma_prefetch_desc(&maps.Wt);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < stages; ++i) {
      mbar_init(&tail->full[i], 1);
      mbar_init(&tail->empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tail->tmem_full[i], DUAL ? 2 : 1);
      mbar_init(&tail->tmem_empty[i], PAIR ? 2 * EPI_WARPS : EPI_WARPS);
    }
    for (int g = 0; g < NCG * 4; ++g)
      for (int i = 0; i < 4; ++i) mbar_init(&tail->rbar[g][i], 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    if (PAIR) tmem_alloc_pair(&tail->tmem_base, 512);
    else tmem_alloc(&tail->tmem_base, 512);
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // the peer's barriers are initialised before anything signals them
  tc_fence_after();
  const uint32_t tmem_base = tail->tmem_base;
  // programmatic dependent launch: the set-up above overlaps the tail of the previous kernel; nothing below reads
  // global memory before the predecessor has completed
  griddep_wait();
  griddep_launch();

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
#ifdef ZV_TRACE
      long long tr_pw = 0, tr_pi = 0;
      const long long tr_t0 = clock64();
#endif
      for (int t = unit0; t < total_tiles; t += nunits) {
        int m_blk, n0t, bn;
        decode(t, m_blk, n0t, bn);
        const int wn = PAIR ? (bn >> 1) : bn;
        const uint32_t tx = A_STAGE_BYTES + wn * BK * 2;
        const CUtensorMap* mapW = bn == block_n ? &maps.W : &maps.Wt;
        for (int kb = 0; kb < k_blocks; ++kb) {
#ifdef ZV_TRACE
          const long long c0 = clock64();
#endif
          mbar_wait(&tail->empty[stage], phase ^ 1);
#ifdef ZV_TRACE
          const long long c1 = clock64();
          tr_pw += c1 - c0;
#endif
          const CUtensorMap* mapA = kb < kb_split ? &maps.A : &maps.A2;
          const int ka = (kb < kb_split ? kb : kb - kb_split) * BK;
          if (PAIR) {
            // both CTAs' bytes complete on the leader's full barrier
            const uint32_t fb = mapa_shared(smem_u32(&tail->full[stage]), 0);
            if (rank == 0) mbar_arrive_expect_tx(&tail->full[stage], 2 * tx);
            tma_load_2d_pair(sA + stage * A_STAGE_BYTES, mapA, fb, ka, m_blk * BM);
            tma_load_2d_pair(sW + stage * w_stage_bytes, mapW, fb, kb * BK, n0t + (int)rank * wn);
          } else {
            mbar_arrive_expect_tx(&tail->full[stage], tx);
            tma_load_2d(sA + stage * A_STAGE_BYTES, mapA, &tail->full[stage], ka, m_blk * BM);
            tma_load_2d(sW + stage * w_stage_bytes, mapW, &tail->full[stage], kb * BK, n0t);
          }
#ifdef ZV_TRACE
          tr_pi += clock64() - c1;
#endif
          if (++stage == stages) stage = 0, phase ^= 1;
        }
      }
#ifdef ZV_TRACE
      if (blockIdx.x == 0)
        printf("[trace] producer: k_blocks %d wait_empty %lld issue %lld total %lld\n", k_blocks, tr_pw, tr_pi,
               clock64() - tr_t0);
#endif
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0 && rank == 0) {
      const uint32_t idesc_main = make_idesc_bf16(PAIR ? 2 * BM : BM, block_n);
      const uint32_t idesc_tail = make_idesc_bf16(PAIR ? 2 * BM : BM, bn_tail);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
#ifdef ZV_TRACE
      long long tr_mw = 0, tr_mi = 0, tr_mc = 0, tr_me = 0;
      const long long tr_m0 = clock64();
#endif
      for (int t = unit0; t < total_tiles; t += nunits, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
#ifdef ZV_TRACE
        const long long e0 = clock64();
#endif
        mbar_wait(&tail->tmem_empty[as], aphase ^ 1);
        tc_fence_after();
#ifdef ZV_TRACE
        tr_me += clock64() - e0;
#endif
        const uint32_t d_tmem = tmem_base + as * ACC_STAGE_COLS;
        const uint32_t idesc = t < sched.main_units ? idesc_main : idesc_tail;
        for (int kb = 0; kb < k_blocks; ++kb) {
          if (DUAL && (kb & 1)) {  // odd K steps belong to the second issuing thread (warp 3)
            if (++stage == stages) stage = 0, phase ^= 1;
            continue;
          }
#ifdef ZV_TRACE
          const long long m0 = clock64();
#endif
          mbar_wait(&tail->full[stage], phase);
          tc_fence_after();
#ifdef ZV_TRACE
          const long long m1 = clock64();
          tr_mw += m1 - m0;
#endif
          const uint64_t da = make_sdesc_sw128(smem_u32(sA + stage * A_STAGE_BYTES));
          const uint64_t dw = make_sdesc_sw128(smem_u32(sW + stage * w_stage_bytes));
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            if (PAIR) umma_bf16_pair(d_tmem, da + 2 * k, dw + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
            else umma_bf16(d_tmem, da + 2 * k, dw + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
          }
#ifdef ZV_TRACE
          const long long m2 = clock64();
          tr_mi += m2 - m1;
#endif
          if (PAIR) {
            umma_commit_pair(&tail->empty[stage], 3);
            if (kb == k_blocks - 1) umma_commit_pair(&tail->tmem_full[as], 3);
          } else {
            umma_commit(&tail->empty[stage]);
            if (!DUAL && kb == k_blocks - 1) umma_commit(&tail->tmem_full[as]);
          }
#ifdef ZV_TRACE
          tr_mc += clock64() - m2;
#endif
          if (++stage == stages) stage = 0, phase ^= 1;
        }
        if (!PAIR && DUAL) umma_commit(&tail->tmem_full[as]);  // one of the two arrivals (warp 3: the other)
      }
#ifdef ZV_TRACE
      if (blockIdx.x == 0)
        printf("[trace] mma: wait_acc_free %lld wait_full %lld issue %lld commit %lld total %lld\n", tr_me, tr_mw, tr_mi,
               tr_mc, clock64() - tr_m0);
#endif
    }
    __syncwarp();
  } else if (warp == 3) {
    // Second MMA-issuing thread (DUAL): one thread needs ~340 cycles per K step for four tcgen05.mma and the
    // commit whatever the tile holds, which is all of the K loop when a CTA owns a single narrow tile.  Odd K
    // steps accumulate into the second (otherwise unused) accumulator stage; the epilogue adds the two.
    if (DUAL && lane == 0 && rank == 0 && unit0 < total_tiles) {
      const uint32_t d_tmem = tmem_base + ACC_STAGE_COLS;
      const uint32_t idesc = make_idesc_bf16(BM, block_n);
      for (int kb = 1; kb < k_blocks; kb += 2) {
        const int stage = kb % stages;
        mbar_wait(&tail->full[stage], (kb / stages) & 1);
        tc_fence_after();
        const uint64_t da = make_sdesc_sw128(smem_u32(sA + stage * A_STAGE_BYTES));
        const uint64_t dw = make_sdesc_sw128(smem_u32(sW + stage * w_stage_bytes));
#pragma unroll
        for (int k = 0; k < BK / 16; ++k)
          umma_bf16(d_tmem, da + 2 * k, dw + 2 * k, idesc, (kb == 1 && k == 0) ? 0u : 1u);
        umma_commit(&tail->empty[stage]);
      }
      umma_commit(&tail->tmem_full[0]);  // arrives once all MMAs of this thread have completed
    }
    __syncwarp();
  } else if (warp >= 4) {
    const int q = warp & 3;            // TMEM lane quarter == warp % 4
    const int cg = (warp - 4) >> 2;    // column group
    const int r = q * 32 + lane;       // row within the tile
    uint8_t* stg = sStg + cg * stg_group_bytes(MODE);
    uint32_t nq = 0;  // chunks processed by this group (selects the staging buffer / residual barrier phase)
#ifdef ZV_TRACE  // cycles of one epilogue warp per phase (bf16-output modes)
    long long et_acc = 0, et_ld = 0, et_math = 0, et_stw = 0, et_stg = 0;
    const long long et_0 = clock64();
#endif

    // EPI_RESIDUAL: the leader prefetches the first residual chunk of this group
    // tmem_empty lives in the leader CTA of a pair
    auto release_acc = [&](int as) {
      if (PAIR) mbar_arrive_cluster_relaxed(mapa_shared(smem_u32(&tail->tmem_empty[as]), 0));
      else mbar_arrive(&tail->tmem_empty[as]);
    };
    // Every epilogue warp is its own pipeline: it stages, loads and stores only its 32 rows (TMA boxes of
    // 32 rows), so the only synchronisation inside the epilogue is __syncwarp.
    uint64_t* rbar = &tail->rbar[cg * 4 + q][0];
    // (every unit of an EPI_RESIDUAL launch is at least 32 * NCG columns wide: each group owns a chunk of it)
    if (MODE == EPI_RESIDUAL && lane == 0 && unit0 < total_tiles) {
      int m_blk, n0t, bn;
      decode(unit0, m_blk, n0t, bn);
      mbar_arrive_expect_tx(&rbar[0], STG_F32_BYTES / 4);
      tma_load_2d(stg + q * (STG_F32_BYTES / 4), &maps.R, &rbar[0], n0t + cg * 32, m_blk * BM + q * 32);
    }
    // EPI_RES24: prefetch cursor of this warp -- the next chunk whose residual rows are to be requested (lane 0)
    int pf_c = cg, pf_t = unit0;
    uint32_t pf_n = 0;
    auto pf_issue = [&]() {
      if (pf_t >= total_tiles) return;
      int mb2, n02, bn2;
      decode(pf_t, mb2, n02, bn2);
      const uint32_t sb2 = pf_n % S24_BUFS;
      uint8_t* nb2 = stg + sb2 * STG_S24_BYTES;
      mbar_arrive_expect_tx(&rbar[sb2], STG_S24_BYTES / 4);
      tma_load_2d(nb2 + q * (STG_BF16_BYTES / 4), &maps.R, &rbar[sb2], n02 + pf_c * 32, mb2 * BM + q * 32);
      tma_load_2d(nb2 + STG_BF16_BYTES + q * (STG_LO_BYTES / 4), &maps.R2, &rbar[sb2], n02 + pf_c * 32,
                  mb2 * BM + q * 32);
      pf_c += NCG, ++pf_n;
      if (pf_c >= (bn2 >> 5)) pf_c = cg, pf_t += nunits;
    };
    if (MODE == EPI_RES24 && ep.res_hi && lane == 0) {
#pragma unroll
      for (int i = 0; i < S24_BUFS - 1; ++i) pf_issue();
    }
    int it = 0;
    for (int t = unit0; t < total_tiles; t += nunits, ++it) {
      int m_blk, n0t, bn;
      decode(t, m_blk, n0t, bn);
      const int nchunks_tile = bn >> 5;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      const int m = m_blk * BM + r;
      const int mm = min(m, ep.M - 1);
      const int b = mm / ep.L;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * ACC_STAGE_COLS;
      const float* rowb =
          (MODE != EPI_RESIDUAL && MODE != EPI_RES24 && ep.rowbias) ? ep.rowbias + (size_t)b * ep.bias_ld : nullptr;
      // chunks that lie entirely in the padded columns of the last n-tile are skipped (never for EPI_RESIDUAL,
      // whose prefetch chain assumes full tiles: its N is a multiple of block_n)
      const int nchunks = (MODE == EPI_RESIDUAL || MODE == EPI_RES24) ? nchunks_tile
                                                                      : min(nchunks_tile, (ep.N - n0t + 31) >> 5);
#ifdef ZV_TRACE
      long long et_c = clock64();
#define ZV_ET(acc) { const long long c_ = clock64(); acc += c_ - et_c; et_c = c_; }
#else
#define ZV_ET(acc)
#endif
      mbar_wait(&tail->tmem_full[as], aphase);
      tc_fence_after();
      ZV_ET(et_acc)
      if (cg >= nchunks) {  // this group owns no chunk of the tile: release the accumulator right away
        tc_fence_before();
        __syncwarp();
        if (lane == 0) release_acc(as);
      }
      for (int c = cg; c < nchunks; c += NCG, ++nq) {
        const int n0 = n0t + c * 32;
        uint32_t raw[32];
        tmem_ld32(taddr + c * 32, raw);
        tmem_ld_wait();
        if (DUAL) {  // + the odd K steps (second accumulator), in two halves to bound the live registers
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            uint32_t odd[16];
            tmem_ld16(taddr + ACC_STAGE_COLS + c * 32 + h * 16, odd);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i)
              raw[h * 16 + i] = __float_as_uint(__uint_as_float(raw[h * 16 + i]) + __uint_as_float(odd[i]));
          }
        }
        if (c + NCG >= nchunks) {  // last accumulator read of this tile by this warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) release_acc(as);
        }
        float v[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[i]);

        if (MODE == EPI_F32) {
          if (ep.f32_direct) {  // output pitch not TMA-addressable: plain stores
            epi_apply<EPI_F32>(ep, m, b, n0, v);
            continue;
          }
          // fp32 rows (+ optional bf16 shadow) staged per warp and written by TMA, like EPI_RESIDUAL without the load
          float* fb = reinterpret_cast<float*>(stg + (nq & 1) * STG_F32_BYTES);
          uint8_t* shadow = stg + 2 * STG_F32_BYTES;
          add_bias32(v, ep.bias, rowb, n0, ep.N);
          if (lane == 0) tma_store_wait_read<0>();
          __syncwarp();
          const int x = r & 7;
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(fb + r * 32 + ((j ^ x) << 2)) =
                make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          if (ep.out_bf16) stage_row_bf16x32(shadow, r, v);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&maps.O1, fb + q * 32 * 32, n0, m_blk * BM + q * 32);
            if (ep.out_bf16) tma_store_2d(&maps.O0, shadow + q * 32 * 64, n0, m_blk * BM + q * 32);
            tma_store_commit();
          }
          continue;
        }
        if (MODE == EPI_RESIDUAL) {
          const int sb = nq & 1;
          float* fb = reinterpret_cast<float*>(stg + sb * STG_F32_BYTES);
          uint8_t* shadow = stg + 2 * STG_F32_BYTES;
          if (lane == 0) {
            // the previous chunk's stores must have finished reading buffer sb^1 and the shadow buffer
            tma_store_wait_read<0>();
            // prefetch the residual of the next chunk of this group (possibly in the next tile)
            int c2 = c + NCG, t2 = t;
            if (c2 >= nchunks) c2 = cg, t2 = t + nunits;
            if (t2 < total_tiles) {
              int mb2 = m_blk, n02 = n0t, bn2;
              if (t2 != t) decode(t2, mb2, n02, bn2);
              mbar_arrive_expect_tx(&rbar[sb ^ 1], STG_F32_BYTES / 4);
              tma_load_2d(stg + (sb ^ 1) * STG_F32_BYTES + q * (STG_F32_BYTES / 4), &maps.R, &rbar[sb ^ 1],
                          n02 + c2 * 32, mb2 * BM + q * 32);
            }
          }
          add_bias32(v, ep.bias, ep.resbias ? ep.resbias + (size_t)b * ep.bias_ld : nullptr, n0, ep.N);
          mbar_wait(&rbar[sb], (nq >> 1) & 1);
          const int x = r & 7;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 rr = *reinterpret_cast<const float4*>(fb + r * 32 + ((j ^ x) << 2));
            v[4 * j] += rr.x, v[4 * j + 1] += rr.y, v[4 * j + 2] += rr.z, v[4 * j + 3] += rr.w;
          }
          if (ep.orig) {
            const float* og = ep.orig + (size_t)mm * ep.ld_f32 + n0;
            if (n0 + 32 <= ep.N) {
#pragma unroll
              for (int i = 0; i < 32; i += 4) {
                const float4 oo = *reinterpret_cast<const float4*>(og + i);
                const float4 bs = __ldg(reinterpret_cast<const float4*>(ep.bypass_scale + n0 + i));
                v[i] = oo.x + (v[i] - oo.x) * bs.x, v[i + 1] = oo.y + (v[i + 1] - oo.y) * bs.y;
                v[i + 2] = oo.z + (v[i + 2] - oo.z) * bs.z, v[i + 3] = oo.w + (v[i + 3] - oo.w) * bs.w;
              }
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i)
                if (n0 + i < ep.N) v[i] = og[i] + (v[i] - og[i]) * ep.bypass_scale[n0 + i];
            }
          }
          __syncwarp();  // shadow buffer free (lane 0 waited), every lane has read its row
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(fb + r * 32 + ((j ^ x) << 2)) =
                make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
          if (ep.out_bf16) stage_row_bf16x32(shadow, r, v);
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&maps.O1, fb + q * 32 * 32, n0, m_blk * BM + q * 32);
            if (ep.out_bf16) tma_store_2d(&maps.O0, shadow + q * 32 * 64, n0, m_blk * BM + q * 32);
            tma_store_commit();
          }
          continue;
        }
        if (MODE == EPI_RES24) {
          // s24 stream: the hi / lo rows of the residual chunk arrive by TMA one chunk ahead, are decoded, updated and
          // re-encoded in registers and leave through the same staging rows
          const uint32_t sb = nq % S24_BUFS;
          uint8_t* hb = stg + sb * STG_S24_BYTES;
          uint8_t* lb = hb + STG_BF16_BYTES;
          if (lane == 0) {
            // the previous chunk's stores have read their buffer, (nq - 1) % S24_BUFS: the one the next prefetch fills
            tma_store_wait_read<0>();
            if (ep.res_hi) pf_issue();
          }
          add_bias32(v, ep.bias, ep.resbias ? ep.resbias + (size_t)b * ep.bias_ld : nullptr, n0, ep.N);
          const int xh = (r >> 1) & 3, xl = (r >> 2) & 1;
          if (ep.res_hi) {
            mbar_wait(&rbar[sb], (nq / S24_BUFS) & 1);
            const uint4 l0 = *reinterpret_cast<const uint4*>(lb + r * 32 + ((0 ^ xl) << 4));
            const uint4 l1 = *reinterpret_cast<const uint4*>(lb + r * 32 + ((1 ^ xl) << 4));
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint4 hh = *reinterpret_cast<const uint4*>(hb + r * 64 + ((j ^ xh) << 4));
              const uint4 lw = j < 2 ? l0 : l1;
              float d[8];
              s24_dec8(hh, (j & 1) ? make_uint2(lw.z, lw.w) : make_uint2(lw.x, lw.y), d);
#pragma unroll
              for (int i = 0; i < 8; ++i) v[8 * j + i] += d[i];
            }
          }
          if (ep.orig_hi) {
            const uint4* oh = reinterpret_cast<const uint4*>(ep.orig_hi + (size_t)mm * ep.ld_bf16 + n0);
            const uint4* ol = reinterpret_cast<const uint4*>(ep.orig_lo + (size_t)mm * ep.ld_bf16 + n0);
            const uint4 l0 = ol[0], l1 = ol[1];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint4 lw = j < 2 ? l0 : l1;
              float d[8];
              s24_dec8(oh[j], (j & 1) ? make_uint2(lw.z, lw.w) : make_uint2(lw.x, lw.y), d);
#pragma unroll
              for (int i = 0; i < 8; i += 4) {
                const float4 bs = __ldg(reinterpret_cast<const float4*>(ep.bypass_scale + n0 + 8 * j + i));
                v[8 * j + i] = d[i] + (v[8 * j + i] - d[i]) * bs.x;
                v[8 * j + i + 1] = d[i + 1] + (v[8 * j + i + 1] - d[i + 1]) * bs.y;
                v[8 * j + i + 2] = d[i + 2] + (v[8 * j + i + 2] - d[i + 2]) * bs.z;
                v[8 * j + i + 3] = d[i + 3] + (v[8 * j + i + 3] - d[i + 3]) * bs.w;
              }
            }
          }
          __syncwarp();  // lane 0 has waited for the stores that read this buffer; every lane has read its row
          uint4 lo4[2];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            float d[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) d[i] = v[8 * j + i];
            uint4 hh;
            uint2 l2;
            s24_enc8(d, hh, l2);
            *reinterpret_cast<uint4*>(hb + r * 64 + ((j ^ xh) << 4)) = hh;
            if (j & 1) lo4[j >> 1].z = l2.x, lo4[j >> 1].w = l2.y;
            else lo4[j >> 1].x = l2.x, lo4[j >> 1].y = l2.y;
          }
          *reinterpret_cast<uint4*>(lb + r * 32 + ((0 ^ xl) << 4)) = lo4[0];
          *reinterpret_cast<uint4*>(lb + r * 32 + ((1 ^ xl) << 4)) = lo4[1];
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&maps.O0, hb + q * (STG_BF16_BYTES / 4), n0, m_blk * BM + q * 32);
            tma_store_2d(&maps.O1, lb + q * (STG_LO_BYTES / 4), n0, m_blk * BM + q * 32);
            tma_store_commit();
          }
          continue;
        }
        // ---- bf16 outputs
        ZV_ET(et_ld)
        // EPI_SWOOSHL_BF16 with gate_cols > 0: columns >= gate_cols are a second, plain projection of the same A rows
        // (the attention q | k | p of the layer beside the first feed-forward's in_proj): bias only, saturating fp16
        const bool tail16 = MODE == EPI_SWOOSHL_BF16 && ep.gate_cols > 0 && n0 >= ep.gate_cols;
        add_bias32(v, ep.bias, tail16 ? nullptr : rowb, n0, ep.N);
        uint8_t* sbuf = stg + (nq & 1) * STG_BF16_BYTES;
        const bool gated = (MODE == EPI_GLU_BF16) || (MODE == EPI_NONLIN_BF16 && n0 < ep.gate_cols);
        float o16[16];
        if (MODE == EPI_SWOOSHL_BF16) {
          if (!tail16) {
#pragma unroll
            for (int i = 0; i < 32; i += 2) swoosh_l_fast2(v[i], v[i + 1]);
          }
        } else if (MODE == EPI_GELU_BF16) {
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = gelu_erf(v[i]);
        } else if (MODE == EPI_GLU_BF16) {
          const bool masked = ep.mask && ep.mask[mm];
#pragma unroll
          for (int i = 0; i < 16; ++i) o16[i] = masked ? 0.0f : v[i] * sigmoid_fast(v[16 + i]);
        } else if (MODE == EPI_NONLIN_BF16) {
          if (gated) {
#pragma unroll
            for (int i = 0; i < 16; ++i) o16[i] = v[16 + i] * tanh_approx(v[i]);
          }
        }
        ZV_ET(et_math)
        if (lane == 0) tma_store_wait_read<1>();  // the store that last used this staging buffer has read it
        __syncwarp();
        ZV_ET(et_stw)
        // every warp owns bytes [q * 2048, q * 2048 + 2048) of the buffer for both row widths (gated rows use half)
        if (gated) stage_row_bf16x16(sbuf + q * 2048, lane, o16);
        else if ((MODE == EPI_BIAS_BF16 && ep.out_f16) || tail16) stage_row_f16x32(sbuf, r, v);
        else stage_row_bf16x32(sbuf, r, v);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          const uint8_t* src = sbuf + q * 2048;
          if ((MODE == EPI_NONLIN_BF16 && !gated) || tail16)
            tma_store_2d(&maps.O1, src, n0 - ep.gate_cols, m_blk * BM + q * 32);
          else tma_store_2d(&maps.O0, src, gated ? (n0 >> 1) : n0, m_blk * BM + q * 32);
          tma_store_commit();
        }
        ZV_ET(et_stg)
      }
    }
    if (lane == 0) tma_store_wait_all<0>();
#ifdef ZV_TRACE
    if (blockIdx.x == 0 && lane == 0 && (warp == 4 || warp == 13))
      printf("[trace] epilogue warp %d: chunks %u wait_acc %lld tmem_ld+release %lld math %lld wait_store %lld "
             "stage+store %lld total %lld\n", warp, nq, et_acc, et_ld, et_math, et_stw, et_stg, clock64() - et_0);
#endif
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // no CTA leaves while its peer may still signal it or be read by the leader's MMAs
  if (warp == 2) {
    if (PAIR) tmem_dealloc_pair(tmem_base, 512);
    else tmem_dealloc(tmem_base, 512);
  }
}

// SIMT checker: one thread per (row, 32-column group); same epilogue.  Tests / debugging only.
template <int MODE>
__global__ void gemm_simt_kernel(const __nv_bfloat16* __restrict__ A, const __nv_bfloat16* __restrict__ W, int lda,
                                 int ldw, int K, int n_groups, EpiParams ep, const __nv_bfloat16* __restrict__ A2,
                                 int lda2, int k_split) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int m = (int)(gid / n_groups);
  const int n0 = (int)(gid % n_groups) * 32;
  if (m >= ep.M) return;
  float acc[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) acc[i] = 0.f;
  for (int k = 0; k < K; ++k) {
    const float a = __bfloat162float(k < k_split ? A[(size_t)m * lda + k] : A2[(size_t)m * lda2 + (k - k_split)]);
#pragma unroll
    for (int i = 0; i < 32; ++i) acc[i] += a * __bfloat162float(W[(size_t)(n0 + i) * ldw + k]);
  }
  epi_apply<MODE>(ep, m, m / ep.L, n0, acc);
}

// ------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    fn = reinterpret_cast<PFN_encodeTiled>(p);
  }
  return fn;
}

// Row-major [rows][cols] tensor with row pitch ld (elements) -> tiled map with box [box_rows][box_cols].
// Maps are cached by their full description: the workspace pointers are stable, so after the first
// NFE no driver call is made on the launch path.
struct TmapKey {
  const void* ptr;
  uint64_t rows, cols, ld;
  uint32_t box_rows, box_cols;
  int dtype, swizzle;
  bool operator==(const TmapKey& o) const {
    return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows &&
           box_cols == o.box_cols && dtype == o.dtype && swizzle == o.swizzle;
  }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    uint64_t h = reinterpret_cast<uint64_t>(k.ptr) * 0x9E3779B97F4A7C15ull;
    h ^= (k.rows + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2));
    h ^= (k.cols * 31 + k.ld * 131 + k.box_rows * 7 + k.box_cols * 3 + k.dtype * 1009 + k.swizzle * 9176 +
          (h << 6) + (h >> 2));
    return (size_t)h;
  }
};
static std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> g_tmap_cache;

static int make_tmap_impl(CUtensorMap* tm, const void* ptr, int dtype /*0 bf16, 1 fp32, 2 u8*/, uint64_t rows, uint64_t cols,
                     uint64_t ld, uint32_t box_rows, uint32_t box_cols, int swizzle /*32, 64, 128*/) {
  const TmapKey key{ptr, rows, cols, ld, box_rows, box_cols, dtype, swizzle};
  auto it = g_tmap_cache.find(key);
  if (it != g_tmap_cache.end()) {
    *tm = it->second;
    return 0;
  }
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) return 901;
  const uint64_t esz = dtype == 0 ? 2 : (dtype == 2 ? 1 : 4);
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || ((ld * esz) & 15)) return 906;  // TMA alignment rules
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * esz};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  const CUtensorMapSwizzle sw = swizzle == 128  ? CU_TENSOR_MAP_SWIZZLE_128B
                                : swizzle == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                : swizzle == 32 ? CU_TENSOR_MAP_SWIZZLE_32B
                                                : CU_TENSOR_MAP_SWIZZLE_NONE;
  const CUtensorMapDataType dt = dtype == 0   ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                 : dtype == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                                              : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  CUresult r = enc(tm, dt, 2,
                   const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return 902;
  if (g_tmap_cache.size() > 4096) g_tmap_cache.clear();
  g_tmap_cache.emplace(key, *tm);
  return 0;
}

int make_tmap(void* tm, const void* ptr, int dtype, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows,
              uint32_t box_cols, int swizzle) {
  return make_tmap_impl(static_cast<CUtensorMap*>(tm), ptr, dtype, rows, cols, ld, box_rows, box_cols, swizzle);
}

struct Tmap3Key {
  const void* ptr;
  uint64_t n0, n1, n2, ld1, ld2;
  uint32_t box0, box1;
  bool operator==(const Tmap3Key& o) const {
    return ptr == o.ptr && n0 == o.n0 && n1 == o.n1 && n2 == o.n2 && ld1 == o.ld1 && ld2 == o.ld2 && box0 == o.box0 &&
           box1 == o.box1;
  }
};
struct Tmap3KeyHash {
  size_t operator()(const Tmap3Key& k) const {
    uint64_t h = reinterpret_cast<uint64_t>(k.ptr) * 0x9E3779B97F4A7C15ull;
    h ^= k.n0 * 31 + k.n1 * 131 + k.n2 * 1009 + k.ld1 * 7 + k.ld2 * 3 + k.box0 * 9176 + k.box1 * 77 + (h << 6) + (h >> 2);
    return (size_t)h;
  }
};
static std::unordered_map<Tmap3Key, CUtensorMap, Tmap3KeyHash> g_tmap3_cache;

int make_tmap3d_16(void* tm, const void* ptr, uint64_t n0, uint64_t n1, uint64_t n2, uint64_t ld1, uint64_t ld2,
                   uint32_t box0, uint32_t box1) {
  const Tmap3Key key{ptr, n0, n1, n2, ld1, ld2, box0, box1};
  auto it = g_tmap3_cache.find(key);
  if (it != g_tmap3_cache.end()) {
    *static_cast<CUtensorMap*>(tm) = it->second;
    return 0;
  }
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) return 901;
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || ((ld1 * 2) & 15) || ((ld2 * 2) & 15) || box0 > 256 || box1 > 256)
    return 906;
  cuuint64_t dims[3] = {n0, n1, n2};
  cuuint64_t strides[2] = {ld1 * 2, ld2 * 2};
  cuuint32_t box[3] = {box0, box1, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(static_cast<CUtensorMap*>(tm), CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims,
                   strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return 902;
  if (g_tmap3_cache.size() > 1024) g_tmap3_cache.clear();
  g_tmap3_cache.emplace(key, *static_cast<CUtensorMap*>(tm));
  return 0;
}

int gemm_pick_block_n(int n_valid, int mode, int gate_cols) {
  // block_n is a multiple of 32 (<= 256: two accumulator stages in 512 TMEM columns); the gated part of
  // EPI_NONLIN_BF16 must end on a tile boundary.  Measured on B200 (tools/gemm_bench.py, K = 512) the time of
  // a 128 x bn tile is ~ (160 + 0.6 bn): the A tile is re-read from L2 for every n-tile, so wide tiles win
  // even with padded columns (N = 1152 -> 5 x 256 beats 6 x 192, N = 1920 -> 8 x 256 beats 10 x 192).
  int best = 32;
  double best_cost = 1e30;
  for (int bn = 32; bn <= 256; bn += 32) {
    if (mode == EPI_NONLIN_BF16 && gate_cols % bn != 0) continue;
    const int tiles = (n_valid + bn - 1) / bn;
    const double cost = tiles * (160.0 + 0.6 * bn);
    if (cost <= best_cost) best = bn, best_cost = cost;
  }
  return best;
}

static int g_num_sms = 0;

template <int MODE, bool PAIR>
static int launch_tc(const GemmMaps& maps_in, const __nv_bfloat16* W, const GemmPlan& p, const EpiParams& ep_in,
                     cudaStream_t stream) {
  constexpr int NCG = ncg_for(MODE);
  constexpr int threads = 128 + 128 * NCG;
  static unsigned long long attr_set = 0;
  static int max_clusters = 0;
  if (first_use_on_device(attr_set)) {
    cudaFuncSetAttribute(gemm_tc_kernel<MODE, PAIR>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (PAIR) {
      cudaFuncSetAttribute(gemm_tc_kernel<MODE, PAIR>, cudaFuncAttributeNonPortableClusterSizeAllowed, 0);
      cudaLaunchConfig_t q = {};
      q.gridDim = dim3(g_num_sms & ~1), q.blockDim = dim3(threads), q.dynamicSmemBytes = 227 * 1024;
      cudaLaunchAttribute qa[1];
      qa[0].id = cudaLaunchAttributeClusterDimension;
      qa[0].val.clusterDim.x = 2, qa[0].val.clusterDim.y = 1, qa[0].val.clusterDim.z = 1;
      q.attrs = qa, q.numAttrs = 1;
      if (cudaOccupancyMaxActiveClusters(&max_clusters, gemm_tc_kernel<MODE, PAIR>, &q) != cudaSuccess ||
          max_clusters < 1) {
        cudaGetLastError();
        max_clusters = g_num_sms / 2;
      }
      if (getenv("ZV_VERBOSE"))
        fprintf(stderr, "[zv] gemm mode %d: %d SMs, %d co-resident CTA pairs\n", MODE, g_num_sms, max_clusters);
    }
  }
  const EpiParams& ep = ep_in;
  GemmMaps maps = maps_in;
  const int m_tiles = (p.M + BM - 1) / BM;
  const int m_units = PAIR ? (m_tiles + 1) / 2 : m_tiles;
  const int max_units = PAIR ? min(max_clusters, g_num_sms / 2) : g_num_sms;  // co-resident work units
  // Column slices of a tile must keep every epilogue column group busy (>= 32 NCG columns), be a legal UMMA N
  // and end the gated part of EPI_NONLIN on a slice boundary; at most `room` of them fit beside each other.
  auto pick_split = [&](int bn_full, int tiles, int room) {
    int split = 1;
    for (int sp = 2; sp <= 8; sp *= 2) {
      const int bn = bn_full / sp;
      if (bn_full % sp || bn % 32 || bn < 32 * NCG || tiles * sp > room) break;
      if (MODE == EPI_NONLIN_BF16 && ep.gate_cols % bn != 0) break;
      split = sp;
    }
    return split;
  };
  // ZV_GEMM_TAILSPLIT: 0 = off, 1 = last wave of large grids only, 2 (default) = small grids too.
  static const int tail_split = getenv("ZV_GEMM_TAILSPLIT") ? atoi(getenv("ZV_GEMM_TAILSPLIT")) : 2;
  // A grid smaller than the machine (single utterances): every tile is "last wave".  Run the whole GEMM on
  // narrower tiles (any multiple of 32 that divides N_pad) -- more SMs pull W (which streams from HBM: the weights
  // of all layers exceed L2) and each K step moves fewer bytes through one SM's ~75 GB/s TMA feed.
  int block_n = p.block_n;
  if (!PAIR && tail_split >= 2 && m_units * (p.N_pad / p.block_n) < max_units) {
    for (int bn = 32; bn < p.block_n; bn += 32) {  // narrowest width whose grid still fits the machine
      if (p.N_pad % bn || m_units * (p.N_pad / bn) > max_units) continue;
      if (MODE == EPI_NONLIN_BF16 && ep.gate_cols % bn != 0) continue;
      block_n = bn;
      break;
    }
    if (block_n != p.block_n) {
      const int rc = make_tmap_impl(&maps.W, W, 0, (uint64_t)p.N_pad, (uint64_t)p.K, (uint64_t)p.ldw,
                                    (uint32_t)block_n, 64, 128);
      if (rc) return rc;
    }
  }
  const int w_stage = (PAIR ? block_n / 2 : block_n) * BK * 2;
  const int fixed = 1024 + (int)sizeof(GemmSmemTail) + NCG * stg_group_bytes(MODE);
  int stages = (227 * 1024 - fixed) / (A_STAGE_BYTES + w_stage);
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  static const int stages_cap = getenv("ZV_GEMM_STAGES") ? atoi(getenv("ZV_GEMM_STAGES")) : 0;  // measurement aid
  if (stages_cap >= 2 && stages > stages_cap) stages = stages_cap;
  if (stages < 2) return 903;
  const int smem = stages * (A_STAGE_BYTES + w_stage) + fixed;
  const int n_tiles = p.N_pad / block_n;
  const int k_blocks = (p.K + BK - 1) / BK;
  const int kb_split = p.A2 ? p.k_split / BK : k_blocks;
  const int units = m_units * n_tiles;
  // Last wave of a large grid: R = units % max_units tiles would keep R CTAs busy for a whole tile time while the
  // others idle.  Cut them into `split` column slices (one slice per CTA, R * split <= max_units).
  TileSched sched = {units, units, 1};
  const int rem = units > max_units ? units % max_units : 0;
  if (tail_split && rem != 0) {
    const int split = pick_split(block_n, rem, max_units);
    if (split > 1) {
      const int bn = block_n / split;
      const int rc = make_tmap_impl(&maps.Wt, W, 0, (uint64_t)p.N_pad, (uint64_t)p.K, (uint64_t)p.ldw,
                                    (uint32_t)(PAIR ? bn / 2 : bn), 64, 128);
      if (rc) return rc;
      sched.main_units = units - rem, sched.split = split, sched.total_units = units - rem + rem * split;
    }
  }
  const int nunits = min(sched.total_units, max_units);
  // one tile per CTA and at least two K steps: split the K loop over two issuing threads / accumulators
  static const bool dual_on = !(getenv("ZV_GEMM_DUAL") && atoi(getenv("ZV_GEMM_DUAL")) == 0);  // A/B measurements
  const bool dual = dual_on && !PAIR && tail_split >= 2 && sched.total_units <= max_units && k_blocks >= 2;
  if (dual) {
    static unsigned long long attr_dual = 0;
    if (first_use_on_device(attr_dual)) {
      cudaFuncSetAttribute(gemm_tc_kernel<MODE, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    }
  }
  static const bool verbose = getenv("ZV_VERBOSE") != nullptr;
  if (verbose)
    fprintf(stderr, "[zv] gemm mode %d M %d N_pad %d K %d bn %d (plan %d), %d stages: %d units, tail split %d -> %d "
            "work items on %d CTAs\n", MODE, p.M, p.N_pad, p.K, block_n, p.block_n, stages, units, sched.split,
            sched.total_units, nunits);
  // Programmatic dependent launch (ZV_PDL=0 switches it off), also as programmatic edges of a captured graph
  // (ZV_PDL=1: plain launches only).
  static const int pdl_on = getenv("ZV_PDL") ? atoi(getenv("ZV_PDL")) : 2;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (pdl_on == 1) cudaStreamIsCapturing(stream, &cap);
  const bool pdl = pdl_on && cap == cudaStreamCaptureStatusNone;  // ZV_PDL=1: not inside graph capture
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(PAIR ? 2 * nunits : nunits);
  cfg.blockDim = dim3(threads), cfg.dynamicSmemBytes = smem, cfg.stream = stream;
  cudaLaunchAttribute at[2];
  int na = 0;
  if (PAIR) {
    at[na].id = cudaLaunchAttributeClusterDimension;
    at[na].val.clusterDim.x = 2, at[na].val.clusterDim.y = 1, at[na].val.clusterDim.z = 1;
    ++na;
  }
  if (pdl) {
    at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  cfg.attrs = at, cfg.numAttrs = na;
  if (dual)
    return (int)cudaLaunchKernelEx(&cfg, gemm_tc_kernel<MODE, false, true>, maps, m_tiles, n_tiles, k_blocks,
                                   kb_split, block_n, stages, ep, sched);
  return (int)cudaLaunchKernelEx(&cfg, gemm_tc_kernel<MODE, PAIR>, maps, m_tiles, n_tiles, k_blocks, kb_split,
                                 block_n, stages, ep, sched);
}

// impl: 0 = tcgen05, CTA pairs (cta_group::2) for the shapes where they measure faster; 1 = SIMT checker;
//       2 = tcgen05 single-CTA tiles only; 3 = tcgen05 CTA pairs only.
template <int MODE>
static int launch_mode(const __nv_bfloat16* A, const __nv_bfloat16* W, const GemmPlan& p, const EpiParams& ep,
                       int impl, cudaStream_t stream) {
  if (impl == 1) {
    const int n_groups = p.N_pad / 32;
    const long long total = (long long)p.M * n_groups;
    const int threads = 128;
    gemm_simt_kernel<MODE><<<(unsigned)((total + threads - 1) / threads), threads, 0, stream>>>(
        A, W, p.lda, p.ldw, p.K, n_groups, ep, p.A2, p.lda2, p.A2 ? p.k_split : p.K);
    return (int)cudaGetLastError();
  }
  if (!g_num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
  }
  const int m_tiles_all = (p.M + BM - 1) / BM;
  // CTA pairs (cta_group::2) halve the W bytes each SM pulls from L2 (a single-CTA 128 x 256 tile at K = 512 pulls
  // 384 KB through its SM's L2 port and its MMA thread waits for operands 22-27 % of the time) and halve the number of
  // tcgen05.mma / commit issues per FLOP.  Measured per shape at M = 156 k (tools/gemm_bench.py, after the accumulator
  // release of a pair stopped using a cluster-scope RELEASE arrive -- ~1000 cycles each inside a busy epilogue warp):
  // pairs against single-CTA tiles, SwooshL N = 1152 / 1536 / 1920 165 / 200 / 265 against 174 / 210 / 271 us, GLU 118
  // against 132, tanh gate 140 against 143, s24 residual K = 1152 / 1536 / 1920 149 / 181 / 214 against 177 / 218 /
  // 261; a tie for the short-K residual shapes (K = 384 / 576) and the narrow q | k | p projection, which stay on
  // single-CTA tiles.
  const bool big = (long long)m_tiles_all * (p.N_pad / p.block_n) >= 2LL * g_num_sms;
  // ZV_GEMM_PAIR_INPROJ=0: single-CTA tiles for the in-proj epilogues (A/B measurements)
  static const bool pair_inproj = !(getenv("ZV_GEMM_PAIR_INPROJ") && atoi(getenv("ZV_GEMM_PAIR_INPROJ")) == 0);
  const bool pair = impl == 3 ||
                    (impl == 0 && big &&
                     (((MODE == EPI_RESIDUAL || MODE == EPI_RES24) && p.K >= 1024) ||
                      (pair_inproj && (MODE == EPI_SWOOSHL_BF16 || MODE == EPI_GLU_BF16 || MODE == EPI_NONLIN_BF16 ||
                                      MODE == EPI_GELU_BF16))));
  GemmMaps maps;
  memset(&maps, 0, sizeof(maps));
  const int k_a = p.A2 ? p.k_split : p.K;
  if (p.A2 && (p.k_split % BK || p.k_split <= 0 || p.k_split >= p.K || p.lda2 % 8)) return 907;
  int rc = make_tmap_impl(&maps.A, A, 0, (uint64_t)p.M, (uint64_t)k_a, (uint64_t)p.lda, BM, 64, 128);
  if (!rc && p.A2)
    rc = make_tmap_impl(&maps.A2, p.A2, 0, (uint64_t)p.M, (uint64_t)(p.K - p.k_split), (uint64_t)p.lda2, BM, 64, 128);
  if (rc) return rc;
  rc = make_tmap_impl(&maps.W, W, 0, (uint64_t)p.N_pad, (uint64_t)p.K, (uint64_t)p.ldw,
                      (uint32_t)(pair ? p.block_n / 2 : p.block_n), 64, 128);
  if (rc) return rc;
  const uint64_t Mu = (uint64_t)ep.M;
  if (MODE == EPI_SWOOSHL_BF16 && ep.gate_cols > 0) {
    if (ep.gate_cols % 32 || ep.gate_cols >= ep.N || !ep.out2_bf16) return 909;
    rc = make_tmap_impl(&maps.O0, ep.out_bf16, 0, Mu, (uint64_t)ep.gate_cols, (uint64_t)ep.ld_bf16, 32, 32, 64);
    if (!rc)
      rc = make_tmap_impl(&maps.O1, ep.out2_bf16, 0, Mu, (uint64_t)(ep.N - ep.gate_cols), (uint64_t)ep.ld2_bf16, 32, 32, 64);
  } else if (MODE == EPI_BIAS_BF16 || MODE == EPI_SWOOSHL_BF16 || MODE == EPI_GELU_BF16) {
    rc = make_tmap_impl(&maps.O0, ep.out_bf16, 0, Mu, (uint64_t)ep.N, (uint64_t)ep.ld_bf16, 32, 32, 64);
  } else if (MODE == EPI_GLU_BF16) {
    rc = make_tmap_impl(&maps.O0, ep.out_bf16, 0, Mu, (uint64_t)ep.N / 2, (uint64_t)ep.ld_bf16, 32, 16, 32);
  } else if (MODE == EPI_NONLIN_BF16) {
    rc = make_tmap_impl(&maps.O0, ep.out_bf16, 0, Mu, (uint64_t)ep.gate_cols / 2, (uint64_t)ep.ld_bf16, 32, 16, 32);
    if (!rc && ep.N > ep.gate_cols)
      rc = make_tmap_impl(&maps.O1, ep.out2_bf16, 0, Mu, (uint64_t)(ep.N - ep.gate_cols), (uint64_t)ep.ld2_bf16, 32, 32, 64);
  } else if (MODE == EPI_RES24) {
    if (!ep.out_bf16 || !ep.out_lo || (ep.res_hi && !ep.res_lo) || (ep.orig_hi && !(ep.orig_lo && ep.bypass_scale)) ||
        ep.N % 32 || ep.ld_bf16 % 16)
      return 908;
    const uint64_t ld = (uint64_t)ep.ld_bf16;
    rc = make_tmap_impl(&maps.O0, ep.out_bf16, 0, Mu, (uint64_t)ep.N, ld, 32, 32, 64);
    if (!rc) rc = make_tmap_impl(&maps.O1, ep.out_lo, 2, Mu, (uint64_t)ep.N, ld, 32, 32, 32);
    if (!rc && ep.res_hi) rc = make_tmap_impl(&maps.R, ep.res_hi, 0, Mu, (uint64_t)ep.N, ld, 32, 32, 64);
    if (!rc && ep.res_hi) rc = make_tmap_impl(&maps.R2, ep.res_lo, 2, Mu, (uint64_t)ep.N, ld, 32, 32, 32);
  } else if (MODE == EPI_RESIDUAL) {
    rc = make_tmap_impl(&maps.O1, ep.out_f32, 1, Mu, (uint64_t)ep.N, (uint64_t)ep.ld_f32, 32, 32, 128);
    if (!rc) rc = make_tmap_impl(&maps.R, ep.res, 1, Mu, (uint64_t)ep.N, (uint64_t)ep.ld_f32, 32, 32, 128);
    if (!rc && ep.out_bf16)
      rc = make_tmap_impl(&maps.O0, ep.out_bf16, 0, Mu, (uint64_t)ep.N, (uint64_t)ep.ld_bf16, 32, 32, 64);
  }
  if (rc) return rc;
  if (MODE == EPI_F32) {
    EpiParams ep2 = ep;
    const bool tma_ok = (ep.ld_f32 % 4 == 0) && ((reinterpret_cast<uintptr_t>(ep.out_f32) & 15) == 0) &&
                        (!ep.out_bf16 || (ep.ld_bf16 % 8 == 0 && (reinterpret_cast<uintptr_t>(ep.out_bf16) & 15) == 0));
    if (tma_ok) {
      rc = make_tmap_impl(&maps.O1, ep.out_f32, 1, Mu, (uint64_t)ep.N, (uint64_t)ep.ld_f32, 32, 32, 128);
      if (!rc && ep.out_bf16)
        rc = make_tmap_impl(&maps.O0, ep.out_bf16, 0, Mu, (uint64_t)ep.N, (uint64_t)ep.ld_bf16, 32, 32, 64);
      if (rc) return rc;
    } else {
      ep2.f32_direct = 1;
    }
    return pair ? launch_tc<MODE, true>(maps, W, p, ep2, stream) : launch_tc<MODE, false>(maps, W, p, ep2, stream);
  }
  return pair ? launch_tc<MODE, true>(maps, W, p, ep, stream) : launch_tc<MODE, false>(maps, W, p, ep, stream);
}

int gemm_launch(const __nv_bfloat16* A, const __nv_bfloat16* W, const GemmPlan& plan, const EpiParams& ep, int impl,
                cudaStream_t stream) {
  if (plan.block_n % 32 || plan.block_n > 256 || plan.N_pad % plan.block_n || plan.lda % 8 || plan.ldw % 8)
    return 904;
  switch (ep.mode) {
    case EPI_BIAS_BF16: return launch_mode<EPI_BIAS_BF16>(A, W, plan, ep, impl, stream);
    case EPI_SWOOSHL_BF16: return launch_mode<EPI_SWOOSHL_BF16>(A, W, plan, ep, impl, stream);
    case EPI_GLU_BF16: return launch_mode<EPI_GLU_BF16>(A, W, plan, ep, impl, stream);
    case EPI_NONLIN_BF16: return launch_mode<EPI_NONLIN_BF16>(A, W, plan, ep, impl, stream);
    case EPI_RESIDUAL: return launch_mode<EPI_RESIDUAL>(A, W, plan, ep, impl, stream);
    case EPI_F32: return launch_mode<EPI_F32>(A, W, plan, ep, impl, stream);
    case EPI_GELU_BF16: return launch_mode<EPI_GELU_BF16>(A, W, plan, ep, impl, stream);
    case EPI_RES24: return launch_mode<EPI_RES24>(A, W, plan, ep, impl, stream);
  }
  return 905;
}

}  // namespace zv
