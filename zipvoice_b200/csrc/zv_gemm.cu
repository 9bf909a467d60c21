// tcgen05 / TMA GEMM with fused Zipformer epilogues (sm_100a).
//   warp 0      : TMA producer (one lane) -- A tile 128x64, W tile block_n x 64, SWIZZLE_128B
//   warp 1      : MMA issuer (one lane)   -- tcgen05.mma kind::f16, M=128, N=block_n, K=16 x4 per stage
//   warp 2      : TMEM allocator (512 columns = 2 accumulator stages of <=256 fp32 columns)
//   warps 4..7  : epilogue: tcgen05.ld 32 lanes x 32 columns -> registers -> fused math -> global
// Persistent: grid = min(#tiles, #SMs); tile t -> (m_blk = t / n_tiles, n_blk = t % n_tiles).
#include <cuda.h>
#include <stdio.h>

#include "zv_common.cuh"
#include "zv_gemm.h"

namespace zv {

static constexpr int BM = 128;
static constexpr int BK = 64;
static constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KB
static constexpr int ACC_STAGE_COLS = 256;
static constexpr int GEMM_THREADS = 256;

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
      if (MODE != EPI_RESIDUAL && ep.rowbias) x += __ldg(ep.rowbias + (size_t)b * ep.bias_ld + n0 + i);
    }
    v[i] = x;
  }
  if (MODE == EPI_BIAS_BF16 || MODE == EPI_SWOOSHL_BF16) {
    __nv_bfloat16* o = ep.out_bf16 + (size_t)m * ep.ld_bf16 + n0;
    if (MODE == EPI_SWOOSHL_BF16) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = swoosh_l(v[i]);
    }
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

struct GemmSmemTail {
  uint64_t full[8];
  uint64_t empty[8];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};

template <int MODE>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, int m_tiles,
               int n_tiles, int k_blocks, int block_n, int stages, EpiParams ep) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: [stages x A 16 KB][stages x W block_n*128 B][tail]
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int w_stage_bytes = block_n * BK * 2;
  uint8_t* sA = smem;
  uint8_t* sW = smem + stages * A_STAGE_BYTES;
  GemmSmemTail* tail = reinterpret_cast<GemmSmemTail*>(sW + stages * w_stage_bytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_tiles = m_tiles * n_tiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < stages; ++i) {
      mbar_init(&tail->full[i], 1);
      mbar_init(&tail->empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tail->tmem_full[i], 1);
      mbar_init(&tail->tmem_empty[i], 128);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(&tail->tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tail->tmem_base;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      const uint32_t tx = A_STAGE_BYTES + w_stage_bytes;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int m_blk = t / n_tiles, n_blk = t % n_tiles;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&tail->empty[stage], phase ^ 1);
          mbar_arrive_expect_tx(&tail->full[stage], tx);
          tma_load_2d(sA + stage * A_STAGE_BYTES, &tmA, &tail->full[stage], kb * BK, m_blk * BM);
          tma_load_2d(sW + stage * w_stage_bytes, &tmW, &tail->full[stage], kb * BK, n_blk * block_n);
          if (++stage == stages) stage = 0, phase ^= 1;
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(BM, block_n);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tail->tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * ACC_STAGE_COLS;
        for (int kb = 0; kb < k_blocks; ++kb) {
          mbar_wait(&tail->full[stage], phase);
          tc_fence_after();
          const uint64_t da = make_sdesc_sw128(smem_u32(sA + stage * A_STAGE_BYTES));
          const uint64_t dw = make_sdesc_sw128(smem_u32(sW + stage * w_stage_bytes));
#pragma unroll
          for (int k = 0; k < BK / 16; ++k)
            umma_bf16(d_tmem, da + 2 * k, dw + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
          umma_commit(&tail->empty[stage]);
          if (kb == k_blocks - 1) umma_commit(&tail->tmem_full[as]);
          if (++stage == stages) stage = 0, phase ^= 1;
        }
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    const int q = warp - 4;  // TMEM lane quarter == warp % 4
    int it = 0;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
      const int m_blk = t / n_tiles, n_blk = t % n_tiles;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      mbar_wait(&tail->tmem_full[as], aphase);
      tc_fence_after();
      const int m = m_blk * BM + q * 32 + lane;
      const int b = m / ep.L;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * ACC_STAGE_COLS;
      for (int c = 0; c < block_n; c += 32) {
        uint32_t r0[16], r1[16];
        tmem_ld16(taddr + c, r0);
        tmem_ld16(taddr + c + 16, r1);
        tmem_ld_wait();
        float acc[32];
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = __uint_as_float(r0[i]), acc[16 + i] = __uint_as_float(r1[i]);
        epi_apply<MODE>(ep, m, b, n_blk * block_n + c, acc);
      }
      tc_fence_before();
      mbar_arrive(&tail->tmem_empty[as]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

// SIMT checker: one thread per (row, 32-column group); same epilogue.  Tests / debugging only.
template <int MODE>
__global__ void gemm_simt_kernel(const __nv_bfloat16* __restrict__ A, const __nv_bfloat16* __restrict__ W, int lda,
                                 int ldw, int K, int n_groups, EpiParams ep) {
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int m = (int)(gid / n_groups);
  const int n0 = (int)(gid % n_groups) * 32;
  if (m >= ep.M) return;
  float acc[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) acc[i] = 0.f;
  for (int k = 0; k < K; ++k) {
    const float a = __bfloat162float(A[(size_t)m * lda + k]);
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

// bf16 row-major [rows][cols] with row pitch ld (elements); box = [box_rows][64], SWIZZLE_128B.
static int make_tmap_bf16(CUtensorMap* tm, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld,
                          uint32_t box_rows) {
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) return 901;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {64, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : 902;
}

int gemm_pick_block_n(int n_valid, int mode, int gate_cols) {
  // block_n is a multiple of 32 (<= 256); the gated part of EPI_NONLIN_BF16 must end on a tile boundary.
  // Preference: largest block_n >= 128 with no padding; else largest with <= 12.5 % padding; else least padding.
  auto ok = [&](int bn) { return !(mode == EPI_NONLIN_BF16 && gate_cols % bn != 0); };
  auto waste = [&](int bn) { return (n_valid + bn - 1) / bn * bn - n_valid; };
  for (int bn = 256; bn >= 128; bn -= 32)
    if (ok(bn) && waste(bn) == 0) return bn;
  for (int bn = 256; bn >= 32; bn -= 32)
    if (ok(bn) && waste(bn) * 8 <= n_valid) return bn;
  int best = 32, best_waste = 1 << 30;
  for (int bn = 256; bn >= 32; bn -= 32)
    if (ok(bn) && waste(bn) < best_waste) best = bn, best_waste = waste(bn);
  return best;
}

static int g_num_sms = 0;

template <int MODE>
static int launch_mode(const __nv_bfloat16* A, const __nv_bfloat16* W, const GemmPlan& p, const EpiParams& ep,
                       int impl, cudaStream_t stream) {
  if (impl == 1) {
    const int n_groups = p.N_pad / 32;
    const long long total = (long long)p.M * n_groups;
    const int threads = 128;
    gemm_simt_kernel<MODE><<<(unsigned)((total + threads - 1) / threads), threads, 0, stream>>>(A, W, p.lda, p.ldw,
                                                                                                 p.K, n_groups, ep);
    return (int)cudaGetLastError();
  }
  if (!g_num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
  }
  CUtensorMap tmA, tmW;
  int rc = make_tmap_bf16(&tmA, A, (uint64_t)p.M, (uint64_t)p.K, (uint64_t)p.lda, BM);
  if (rc) return rc;
  rc = make_tmap_bf16(&tmW, W, (uint64_t)p.N_pad, (uint64_t)p.K, (uint64_t)p.ldw, (uint32_t)p.block_n);
  if (rc) return rc;
  const int w_stage = p.block_n * BK * 2;
  const int budget = 227 * 1024 - 1024 - (int)sizeof(GemmSmemTail);
  int stages = budget / (A_STAGE_BYTES + w_stage);
  if (stages > 8) stages = 8;
  if (stages < 2) return 903;
  const int smem = stages * (A_STAGE_BYTES + w_stage) + (int)sizeof(GemmSmemTail) + 1024;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(gemm_tc_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    attr_set = true;
  }
  const int m_tiles = (p.M + BM - 1) / BM;
  const int n_tiles = p.N_pad / p.block_n;
  const int k_blocks = (p.K + BK - 1) / BK;
  const int grid = min(m_tiles * n_tiles, g_num_sms);
  gemm_tc_kernel<MODE><<<grid, GEMM_THREADS, smem, stream>>>(tmA, tmW, m_tiles, n_tiles, k_blocks, p.block_n, stages,
                                                             ep);
  return (int)cudaGetLastError();
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
  }
  return 905;
}

}  // namespace zv
