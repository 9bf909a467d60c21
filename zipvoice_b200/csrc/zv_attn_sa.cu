// SelfAttention value product (zipformer.py:1359-1396 on the weights of :1149-1306), dv = 12 per head:
//   O_h = softmax_j(q_h k_h^T + p_h . E_h[L-1-i+j], key-padded -> -1000) V_h        (scores in log2 units)
// Same flash-style mathematics as attn_kernel<16> (zv_attn.cu) with a Blackwell data path:
//   producer warp : feeds a 4-stage shared-memory ring with TMA (K tile: SWIZZLE_64B tensor box of the fp16 q|k|p
//                   rows; V tile: 16-column box at column 16 h of the padded value rows; relative-position window:
//                   one 1-D bulk copy);
//   MMA warp      : S = Q K^T of every key tile with tcgen05.mma kind::f16 on those fp16 tiles as they lie,
//                   S[128 x 64] fp32 double-buffered in 128 TMEM columns;
//   8 compute warps (16 query rows each): tcgen05.ld 16x256b -> the mma.sync accumulator layout -> position term
//                   on CUDA cores, exp2, P V on mma.sync.  Interior key tiles are exponentiated against the running
//                   reference maximum (key_tile, SPEC); tile 0, the last tile and repeats are exact.
// The roles synchronise only through mbarriers -- no CTA-wide barrier in the key loop.
#include <cuda.h>
#include <stdlib.h>

#include <type_traits>

#include <stdio.h>

#include "zv_common.cuh"
#include "zv_gemm.h"
#include "zv_kernels.h"

namespace zv {
namespace sa {

static constexpr int BQ = 128, BKV = 64, DQ = 32, EWIN = 192, STAGES = 4, DVP = 16;
static constexpr int COMPUTE_THREADS = 256, THREADS = COMPUTE_THREADS + 64;  // + producer warp + MMA warp
static constexpr int PRODUCER_WARP = COMPUTE_THREADS / 32, MMA_WARP = PRODUCER_WARP + 1;
static constexpr int S_COLS = 2 * BKV;  // TMEM columns: two S[128 x 64] fp32 accumulators
// speculative tiles: exponent bound (log2) of P against the running reference (2^64 * 1219 keys << fp32 max)
static constexpr float SPEC_LOG2 = 64.0f;

struct Stage {
  __half k[BKV][DQ];    // 4 KB, SWIZZLE_64B
  bf16 v[BKV][DVP];     // 2 KB, no swizzle
  float4 e[EWIN];       // 3 KB
};
struct alignas(16) Smem {  // the key mask that follows is read with 16-byte loads
  __half q[BQ][DQ];     // 8 KB, SWIZZLE_64B (two 64-row boxes)
  Stage st[STAGES];
  uint64_t full[STAGES], empty[STAGES], qbar, pad_;
  uint64_t sfull[2], sfree[2];  // S accumulator in TMEM: written by the MMAs of a tile / read by all compute warps
  uint32_t tmem_base, pad2_[3];
  // followed by: uint8_t mask[n_tiles * BKV], uint8_t flag[n_tiles]
};

ZV_DEVINL void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
// K-major SWIZZLE_64B operand with 64-byte rows (32 fp16): 8-row swizzle atoms 512 B apart (SBO), LBO unused;
// version 1 (sm_100), layout_type 4 = SWIZZLE_64B (cute/arch/mma_sm100_desc.hpp)
ZV_DEVINL uint64_t make_sdesc_sw64(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(1) << 16;
  d |= static_cast<uint64_t>(512 >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(4) << 61;
  return d;
}
// kind::f16 instruction descriptor: A, B = IEEE fp16 (format 0) K-major, D = fp32
ZV_DEVINL uint32_t make_idesc_f16(uint32_t M, uint32_t N) { return (1u << 4) | ((N >> 3) << 17) | ((M >> 4) << 24); }
// 16 TMEM lanes x 64 fp32 columns in the mma.sync accumulator layout: r[4 nt + {0,1}] = lane g, columns
// 8 nt + 2 qd + {0,1}; r[4 nt + {2,3}] = lane g + 8 (cute SM100_TMEM_LOAD_16dp256b8x)
ZV_DEVINL void tmem_ld_16x256b_x8(uint32_t taddr, float (&s)[8][4]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.16x256b.x8.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < 32; ++i) s[i >> 2][i & 3] = __uint_as_float(r[i]);
}
// D[tmem] (+)= A[smem] * B[smem]^T, fp16 inputs (format chosen by the instruction descriptor); ONE thread issues
ZV_DEVINL void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  umma_bf16(tmem_d, desc_a, desc_b, idesc, accumulate);  // same tcgen05.mma.kind::f16 instruction
}
ZV_DEVINL void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// 1-D bulk copy global -> shared, completion counted on an mbarrier (bytes multiple of 16, 16-byte aligned)
ZV_DEVINL void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

}  // namespace sa

using namespace sa;

__global__ void __launch_bounds__(sa::THREADS, 2)
attn_sa_kernel(const __grid_constant__ CUtensorMap tmQK, const __grid_constant__ CUtensorMap tmV, AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // align inside the shared window (pointer + integer keeps the shared address space: LDS/STS, not generic LD/ST)
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.z, head = blockIdx.y, i0 = blockIdx.x * BQ;
  const int L = p.L, H = p.H, R = 2 * L - 1;
  const int n_tiles = (L + BKV - 1) / BKV;
  uint8_t* smask = reinterpret_cast<uint8_t*>(&sm) + sizeof(Smem);
  uint8_t* sflag = smask + n_tiles * BKV;
  const uint8_t* mask = p.mask ? p.mask + (size_t)b * L : nullptr;

  if (tid == 0) {
    for (int i = 0; i < STAGES; ++i) mbar_init(&sm.full[i], 1), mbar_init(&sm.empty[i], COMPUTE_THREADS / 32);
    mbar_init(&sm.qbar, 1);
    for (int i = 0; i < 2; ++i) mbar_init(&sm.sfull[i], 1), mbar_init(&sm.sfree[i], COMPUTE_THREADS / 32);
    fence_barrier_init();
  }
  if (warp == MMA_WARP) tmem_alloc(&sm.tmem_base, S_COLS);
  // window entries outside the table are never copied: keep them finite
  for (int i = tid; i < STAGES * EWIN; i += THREADS) sm.st[i / EWIN].e[i % EWIN] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int j = tid; j < n_tiles * BKV; j += THREADS) smask[j] = (j < L) ? (mask ? mask[j] : 0) : 2;
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  for (int t = tid; t < n_tiles; t += THREADS) {
    const uint4* mp = reinterpret_cast<const uint4*>(smask + t * BKV);
    const uint4 a = mp[0], b4 = mp[1], c = mp[2], d = mp[3];
    sflag[t] =
        ((a.x | a.y | a.z | a.w | b4.x | b4.y | b4.z | b4.w | c.x | c.y | c.z | c.w | d.x | d.y | d.z | d.w) != 0);
  }
  __syncthreads();

  const uint32_t tmem_base = sm.tmem_base;
  if (warp == MMA_WARP) {
    // ---------------- S = Q K^T of every key tile on tcgen05: fp16 operands straight from the TMA-written
    // SWIZZLE_64B tiles, fp32 accumulator S[128 x 64] in TMEM, double buffered against the compute warps
    if (lane == 0) {
      const uint32_t idesc = make_idesc_f16(BQ, BKV);
      mbar_wait(&sm.qbar, 0);
      const uint64_t dq = make_sdesc_sw64(smem_u32(&sm.q[0][0]));
      for (int kt = 0; kt < n_tiles; ++kt) {
        const int st = kt % STAGES, sb = kt & 1;
        mbar_wait(&sm.full[st], (kt / STAGES) & 1);
        if (kt >= 2) mbar_wait(&sm.sfree[sb], ((kt >> 1) - 1) & 1);
        tc_fence_after();
        const uint64_t dk = make_sdesc_sw64(smem_u32(&sm.st[st].k[0][0]));
#pragma unroll
        for (int k = 0; k < DQ / 16; ++k) umma_f16(tmem_base + sb * BKV, dq + 2 * k, dk + 2 * k, idesc, k);
        umma_commit(&sm.sfull[sb]);
      }
    }
    __syncwarp();
  } else if (warp == PRODUCER_WARP) {
    // ---------------- producer
    if (lane == 0) {
      tma_prefetch_desc(&tmQK);
      tma_prefetch_desc(&tmV);
      if (p.dbg & 8) {
        mbar_arrive(&sm.qbar);
      } else {
        mbar_arrive_expect_tx(&sm.qbar, BQ * DQ * 2);
        tma_load_2d(&sm.q[0][0], &tmQK, &sm.qbar, head * DQ, b * L + i0);
        tma_load_2d(&sm.q[64][0], &tmQK, &sm.qbar, head * DQ, b * L + i0 + 64);
      }
      const float4* E = reinterpret_cast<const float4*>(p.E) + (size_t)head * R;
      for (int kt = 0; kt < n_tiles; ++kt) {
        const int s = kt % STAGES;
        if (kt >= STAGES) mbar_wait_backoff(&sm.empty[s], ((kt / STAGES) - 1) & 1);
        const int j0 = kt * BKV;
        const int rr0 = (j0 - i0 - (BQ - 1)) + (L - 1);  // table row of window entry 0
        const int lo = max(rr0, 0), hi = min(rr0 + EWIN, R);
        const uint32_t ebytes = (hi > lo && !(p.dbg & 1)) ? (uint32_t)(hi - lo) * 16u : 0u;
        const uint32_t kbytes = (p.dbg & 4) ? 0u : BKV * DQ * 2, vbytes = (p.dbg & 2) ? 0u : BKV * DVP * 2;
        mbar_arrive_expect_tx(&sm.full[s], kbytes + vbytes + ebytes);
        if (kbytes) tma_load_2d(&sm.st[s].k[0][0], &tmQK, &sm.full[s], H * DQ + head * DQ, b * L + j0);
        if (vbytes) tma_load_2d(&sm.st[s].v[0][0], &tmV, &sm.full[s], head * DVP, b * L + j0);
        if (ebytes) bulk_load_1d(&sm.st[s].e[lo - rr0], E + lo, ebytes, &sm.full[s]);
      }
    }
    __syncwarp();
  } else {
    // ---------------- compute warps: 16 query rows each
    const int g = lane >> 2, qd = lane & 3;
    // a warp reads TMEM lanes of its own quarter only (warp % 4): rows 32 (warp % 4) + 16 (warp / 4) + ...
    const int row0 = 32 * (warp & 3) + 16 * (warp >> 2);
    const int rl0 = row0 + g, rl1 = rl0 + 8;
    const __half* qkp = reinterpret_cast<const __half*>(p.qkp) + (size_t)b * L * p.ldq;
    float pq0[4], pq1[4];
    {
      const __half* pa = qkp + (size_t)min(i0 + rl0, L - 1) * p.ldq + 2 * H * DQ + head * 4;
      const __half* pb = qkp + (size_t)min(i0 + rl1, L - 1) * p.ldq + 2 * H * DQ + head * 4;
#pragma unroll
      for (int c = 0; c < 4; ++c) pq0[c] = __half2float(pa[c]), pq1[c] = __half2float(pb[c]);
    }
    float o[2][4];
#pragma unroll
    for (int n = 0; n < 2; ++n) o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY;  // the row sums live in column dv of O (V carries a ones column)
    const uint32_t s_taddr = tmem_base + (static_cast<uint32_t>(row0) << 16);
    // One key tile, as ONE straight-line block per variant so that the scheduler can overlap the tensor, FMA and
    // MUFU work of a warp:
    //   SPEC  -- interior tiles: P = exp(S - m) against the running reference m WITHOUT waiting for the tile's own
    //            maximum (softmax is shift invariant; P is bf16 and the sums fp32, so any shift whose exponent
    //            stays below 2^SPEC_LOG2 is exact to rounding).  The tile maximum is reduced on the side; if some
    //            score exceeds its reference by more than SPEC_LOG2 the warp returns false before touching O and
    //            the caller repeats the tile the exact way.
    //   exact -- tile 0 (sets the reference), the last tile (only its first nbv 8-key blocks hold keys: the MMAs,
    //            position terms and exponentials of the absent blocks are skipped) and repeats: classic online
    //            softmax with the maximum first and O rescaled when it grows.
    // mode_tag: 0 = SPEC, 1 = exact, 2 = exact with the block count of the last tile
    auto key_tile = [&](const int kt, auto mode_tag, auto masked_tag) -> bool {
      constexpr bool SPEC = decltype(mode_tag)::value == 0;
      constexpr bool TAIL = decltype(mode_tag)::value == 2;
      constexpr bool MASKED = decltype(masked_tag)::value;
      const int nbv = TAIL ? min(8, (L - kt * BKV + 7) >> 3) : 8;
      const int st = kt % STAGES;
      mbar_wait(&sm.full[st], (kt / STAGES) & 1);
      const Stage& S = sm.st[st];
      const uint32_t vbase = smem_u32(&S.v[0][0]);
      const uint8_t* mrow = smask + kt * BKV;
      // ---- S = Q K^T from TMEM (this warp's 16 rows, all 64 keys, mma.sync accumulator layout)
      const int sb = kt & 1;
      float s[8][4];
      mbar_wait(&sm.sfull[sb], (kt >> 1) & 1);
      tc_fence_after();
      tmem_ld_16x256b_x8(s_taddr + sb * BKV, s);
      tc_fence_before();
      // the buffer is handed back once this warp cannot need the raw scores again: a speculative tile may still be
      // repeated the exact way (the repeat reads S again), so it releases after its vote
      auto release_s = [&] {
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.sfree[sb]);
      };
      if (!SPEC) release_s();
      if (TAIL) {  // blocks beyond the sequence: P = 0
#pragma unroll
        for (int nt = 0; nt < 8; ++nt)
          if (nt >= nbv) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
      }
      // ---- position term (rows g and g + 8 share table entries 8 columns apart), masks, tile maximum;
      //      SPEC: P right away (mx0 / mx1 then hold the largest P of the two rows: the vote below depends on the
      //      exponentials, which keeps them next to the FMAs in the schedule instead of behind the branch)
      float mx0 = -INFINITY, mx1 = -INFINITY;
      {
        const float4* ew = &S.e[2 * qd - rl0 + (BQ - 1)];
        float4 ep0 = ew[-8], ep1 = ew[-7];
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          if (TAIL && nt >= nbv) break;
          const float4 ec0 = ew[nt * 8], ec1 = ew[nt * 8 + 1];
          float a0 = s[nt][0] + pq0[0] * ec0.x + pq0[1] * ec0.y + pq0[2] * ec0.z + pq0[3] * ec0.w;
          float a1 = s[nt][1] + pq0[0] * ec1.x + pq0[1] * ec1.y + pq0[2] * ec1.z + pq0[3] * ec1.w;
          float b0 = s[nt][2] + pq1[0] * ep0.x + pq1[1] * ep0.y + pq1[2] * ep0.z + pq1[3] * ep0.w;
          float b1 = s[nt][3] + pq1[0] * ep1.x + pq1[1] * ep1.y + pq1[2] * ep1.z + pq1[3] * ep1.w;
          if (MASKED) {
            const uint8_t k0 = mrow[nt * 8 + 2 * qd], k1 = mrow[nt * 8 + 2 * qd + 1];
            if (k0) a0 = b0 = (k0 == 1 ? ATTN_MASKED : -INFINITY);
            if (k1) a1 = b1 = (k1 == 1 ? ATTN_MASKED : -INFINITY);
          }
          if (SPEC) a0 = ex2_approx(a0 - m0), a1 = ex2_approx(a1 - m0), b0 = ex2_approx(b0 - m1), b1 = ex2_approx(b1 - m1);
          s[nt][0] = a0, s[nt][1] = a1, s[nt][2] = b0, s[nt][3] = b1;
          mx0 = fmaxf(mx0, fmaxf(a0, a1)), mx1 = fmaxf(mx1, fmaxf(b0, b1));
          ep0 = ec0, ep1 = ec1;
        }
      }
      float mn0 = m0, mn1 = m1;
      if (!SPEC) {
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
      }
      const float ms0 = mn0, ms1 = mn1;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        if (SPEC || (TAIL && nt >= nbv)) break;  // s stays 0 = P of the absent keys
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const float pa = ex2_approx(s[nt][e] - ms0), pb = ex2_approx(s[nt][2 + e] - ms1);
          s[nt][e] = pa, s[nt][2 + e] = pb;
        }
      }
      if (SPEC) {
        // a lane's own maxima suffice: any lane over the bound sends the whole warp to the exact repeat
        const float bound = 18446744073709551616.0f;  // 2^SPEC_LOG2
        if (__any_sync(0xffffffffu, !(fmaxf(mx0, mx1) <= bound))) return false;
        release_s();
      } else if (__any_sync(0xffffffffu, (mn0 != m0) || (mn1 != m1))) {
        const float sc0 = ex2_approx(m0 - mn0), sc1 = ex2_approx(m1 - mn1);
#pragma unroll
        for (int n = 0; n < 2; ++n) o[n][0] *= sc0, o[n][1] *= sc0, o[n][2] *= sc1, o[n][3] *= sc1;
        m0 = mn0, m1 = mn1;
      }
      // ---- O += P V
#pragma unroll
      for (int kk = 0; kk < BKV / 16; ++kk) {
        if (TAIL && 2 * kk >= nbv) break;
        uint32_t pa[4];
        pa[0] = pack_bf16x2(s[2 * kk][0], s[2 * kk][1]);
        pa[1] = pack_bf16x2(s[2 * kk][2], s[2 * kk][3]);
        pa[2] = pack_bf16x2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
        pa[3] = pack_bf16x2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
        uint32_t vb[4];
        ldsm_x4_t(vb, vbase + (kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * (DVP * 2) + (lane >> 4) * 16);
        mma_bf16(o[0], pa, vb[0], vb[1]);
        mma_bf16(o[1], pa, vb[2], vb[3]);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.empty[st]);
      return true;
    };
    const std::true_type masked{};
    const std::false_type clear{};
    const std::integral_constant<int, 0> spec{};
    const std::integral_constant<int, 1> exact{};
    const std::integral_constant<int, 2> exact_tail{};
    bool flag = sflag[0] != 0;
    for (int kt = 0; kt + 1 < n_tiles; ++kt) {
      const bool mk = flag;
      flag = sflag[kt + 1] != 0;  // fetched a tile ahead of its use
      bool done = false;
      if (kt != 0 && !(p.dbg & 16)) done = mk ? key_tile(kt, spec, masked) : key_tile(kt, spec, clear);
      if (!done) {
        if (mk) key_tile(kt, exact, masked);
        else key_tile(kt, exact, clear);
      }
    }
    key_tile(n_tiles - 1, exact_tail, masked);
    // ---- normalise, store (columns >= dv belong to the next head / padding and are dropped)
    // column dv = 12 of O (n-tile 1, columns 4..5 -> the lane with qd == 2) is sum_k P[i][k] * 1
    const float l0 = __shfl_sync(0xffffffffu, o[1][0], (lane & ~3) | 2);
    const float l1 = __shfl_sync(0xffffffffu, o[1][2], (lane & ~3) | 2);
    const float inv0 = 1.f / l0, inv1 = 1.f / l1;
    const int ia = i0 + rl0, ib = i0 + rl1;
    const int vcol0 = head * p.dv;
#pragma unroll
    for (int n = 0; n < 2; ++n) {
      const int col = n * 8 + 2 * qd;
      if (col >= p.dv) continue;
      if (ia < L)
        *reinterpret_cast<uint32_t*>(p.out + ((size_t)b * L + ia) * p.ldo + vcol0 + col) =
            pack_bf16x2(o[n][0] * inv0, o[n][1] * inv0);
      if (ib < L)
        *reinterpret_cast<uint32_t*>(p.out + ((size_t)b * L + ib) * p.ldo + vcol0 + col) =
            pack_bf16x2(o[n][2] * inv1, o[n][3] * inv1);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) tmem_dealloc(tmem_base, S_COLS);
}

// =====================================================================================================================
// attn_sa2_kernel: the same attention with the relative-position term ON THE TENSOR CORES.
//
//   T[m][n] = p_m . W[127 - m + n]          (m: query row of the 128-row tile, n: key of the 64-key tile, W: the
//                                            window of the projected position table that the tile pair can see)
// is a Toeplitz form, not a product of a row factor and a column factor.  Write m = 8 a + s and u = n + 120 - 8 a:
//   T[m][n] = p_m . W[u + s'],  s' = 7 - s   =   sum_{s', c} A[m][4 s' + c] * B[u][4 s' + c]
//   A[m][4 s' + c] = p_m[c] if s' == 7 - (m % 8) else 0          (128 x 32 fp16, built once per CTA)
//   B[u][4 s' + c] = W[u + s'][c]  =  Wflat[4 u + k],  k = 4 s' + c
// so row u of B is nothing but 32 consecutive halves of the flat fp16 window starting at half 4 u: an operand with
// OVERLAPPING rows, 8 bytes apart.  The no-swizzle K-major UMMA layout wants the 8 rows of a core matrix 16 bytes
// apart, so the even rows (u = 2 v) are read from the window as it lies (row v at byte 16 v: core matrix = 128
// contiguous bytes, SBO = 128, LBO = 16) and the odd rows from a second copy shifted by one entry; nothing is
// materialised.  D_even / D_odd [128 x 96] land in TMEM beside S = Q K^T.  What is left of the skew, 8 a, is the
// same for the 8 rows a warp reads with one TMEM lane group, i.e. a per-warp COLUMN BASE of tcgen05.ld:
// shape 16x128b hands thread (g, qd) the columns base + 4 i + qd of rows g and g + 8 -- with the even / odd split
// exactly the keys 8 nt + 2 qd (+ 1) of the mma.sync fragment that S arrives in.  Per score: one FADD instead of four
// FMAs and a shared-memory table load (128 of ~300 instructions per thread and tile before).
// The table is rounded to fp16 for this (q, k, p already are); accumulation is fp32.
// TMEM: S 64 + D_even 96 + D_odd 96 = 256 columns, two CTAs per SM; one accumulator set, handed back to the MMA warp as
// soon as the 8 compute warps hold their scores in registers (the speculative / exact decision is taken on the score
// maxima before any exponential, so a tile never has to be read twice).
namespace sa2 {

// tile sizes, warp roles and SPEC_LOG2 are those of namespace sa above (two CTAs of 8 compute warps per SM; a layout
// with one CTA of 16 compute warps, two accumulator sets and two MMA-issuing warps measured 0.62 ms against 0.44 ms)
static constexpr int EW16 = 200;   // fp16 window entries (8 bytes) per copy: rows u < 96 x 2 read up to entry 197
static constexpr int D_N = 96;     // columns of D_even / D_odd (u < 184 are used)
static constexpr int COL_S = 0, COL_DE = 64, COL_DO = 160, TM_COLS = 256;

struct alignas(1024) Smem {
  __half q[BQ][DQ];            // 8 KB, SWIZZLE_64B (TMA)
  __half ap[BQ][DQ];           // 8 KB, SWIZZLE_64B: one-hot position operand A
  __half k[STAGES][BKV][DQ];   // 4 KB each, SWIZZLE_64B (TMA)
  bf16 v[STAGES][BKV][DVP];    // 2 KB each, SWIZZLE_32B (TMA)
  uint2 we[STAGES][EW16];      // position window, entries x0 .. x0 + 199
  uint2 wo[STAGES][EW16];      // ... shifted by one entry
  uint64_t full[STAGES], empty[STAGES], qbar, sfull, sfree, pad_;
  uint32_t tmem_base, pad2_[3];
  // followed by: uint8_t mask[n_tiles * BKV], uint8_t flag[n_tiles]
};

// K-major operand without swizzle: core matrices of 8 rows x 16 bytes (128 contiguous bytes), `sbo` bytes between
// core matrices along M / N, `lbo` bytes between the two 16-byte K chunks of one K = 16 instruction
ZV_DEVINL uint64_t make_sdesc_kmajor_plain(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(lbo >> 4) << 16;
  d |= static_cast<uint64_t>(sbo >> 4) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  return d;
}
// 16 TMEM lanes x (4 N) fp32 columns: r[2 i] = lane g, column 4 i + qd; r[2 i + 1] = lane g + 8, same column
// (cute SM100_TMEM_LOAD_16dp128b8x / 1x); NO wait inside
ZV_DEVINL void tmem_ld_16x128b_x8(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.16x128b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
ZV_DEVINL void tmem_ld_16x128b_x1(uint32_t taddr, uint32_t (&r)[2]) {
  asm volatile("tcgen05.ld.sync.aligned.16x128b.x1.b32 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "r"(taddr) : "memory");
}
ZV_DEVINL float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
// 2^x on the FMA / ALU pipes for x <= 64 (x below -126, -inf included, gives 2^-126): Cody-Waite split at the nearest
// integer (magic-number rounding), a cubic for 2^f on [-0.5, 0.5] (max relative error 7.5e-5, P is rounded to bf16
// at 2e-3 afterwards) and the integer part added to the exponent field.  MUFU.EX2 runs at 16 results / clk / SM for
// fp32, f16x2 and bf16x2 alike (tools/micro/mufu_f16_rate.cu).  Measurement aid (ZV_SA_SOFT, see the launcher): the
// default kernel computes every exponential on MUFU.
ZV_DEVINL float ex2_soft(float x) {
  x = fmaxf(x, -126.0f);
  const float xr = x + 12582912.0f;          // 1.5 * 2^23: the low mantissa bits now hold round(x)
  const float f = x - (xr - 12582912.0f);
  float pl = fmaf(0.055171654f, f, 0.24261113f);
  pl = fmaf(pl, f, 0.69326097f);
  pl = fmaf(pl, f, 0.99992806f);
  return __int_as_float(__float_as_int(pl) + (__float_as_int(xr) << 23));
}
// element e (0, 1: row g; 2, 3: row g + 8) of every 8-key block goes to the FMA pipe when e < SOFT
template <int SOFT, int E>
ZV_DEVINL float ex2_sel(float x) {
  return E < SOFT ? ex2_soft(x) : ex2_approx(x);
}

}  // namespace sa2

template <int SOFT>
__global__ void __launch_bounds__(sa::THREADS, 2)
attn_sa2_kernel(const __grid_constant__ CUtensorMap tmQK, const __grid_constant__ CUtensorMap tmV, AttnParams p,
                int ksplit) {
  using namespace sa2;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  sa2::Smem& sm = *reinterpret_cast<sa2::Smem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.z, head = blockIdx.y;
  const int qt = blockIdx.x / ksplit, sp = blockIdx.x - qt * ksplit, i0 = qt * BQ;
  const int L = p.L, H = p.H;
  const int n_tiles = (L + BKV - 1) / BKV;
  // this CTA's key tiles [kt0, kt1): never empty (the launcher keeps (ksplit - 1) * tiles_per < n_tiles)
  const int tiles_per = (n_tiles + ksplit - 1) / ksplit;
  const int kt0 = sp * tiles_per, kt1 = min(n_tiles, kt0 + tiles_per);
  uint8_t* smask = reinterpret_cast<uint8_t*>(&sm) + sizeof(sa2::Smem);
  uint8_t* sflag = smask + n_tiles * BKV;
  const uint8_t* mask = p.mask ? p.mask + (size_t)b * L : nullptr;
  const __half* qkp = reinterpret_cast<const __half*>(p.qkp) + (size_t)b * L * p.ldq;

  if (tid == 0) {
    for (int i = 0; i < STAGES; ++i) mbar_init(&sm.full[i], 1), mbar_init(&sm.empty[i], COMPUTE_THREADS / 32);
    mbar_init(&sm.qbar, 1);
    mbar_init(&sm.sfull, 1), mbar_init(&sm.sfree, COMPUTE_THREADS / 32);
    fence_barrier_init();
  }
  if (warp == MMA_WARP) tmem_alloc(&sm.tmem_base, TM_COLS);
  if (tid < BQ) {
    // row m of A: zeros except the 4 halves of p_m at K position 4 (7 - m % 8); SWIZZLE_64B rows of 64 bytes
    const int m = tid;
    uint8_t* row = reinterpret_cast<uint8_t*>(&sm.ap[m][0]);
#pragma unroll
    for (int c = 0; c < 4; ++c) *reinterpret_cast<uint4*>(row + c * 16) = make_uint4(0u, 0u, 0u, 0u);
    const uint2 pm = *reinterpret_cast<const uint2*>(qkp + (size_t)min(i0 + m, L - 1) * p.ldq + 2 * H * DQ + head * 4);
    const int sp = 7 - (m & 7);
    *reinterpret_cast<uint2*>(row + (((sp >> 1) ^ ((m >> 1) & 3)) << 4) + (sp & 1) * 8) = pm;
  }
  for (int j = tid; j < n_tiles * BKV; j += THREADS) smask[j] = (j < L) ? (mask ? mask[j] : 0) : 2;
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  for (int t = tid; t < n_tiles; t += THREADS) {
    const uint4* mp = reinterpret_cast<const uint4*>(smask + t * BKV);
    const uint4 a = mp[0], b4 = mp[1], c = mp[2], d = mp[3];
    sflag[t] =
        ((a.x | a.y | a.z | a.w | b4.x | b4.y | b4.z | b4.w | c.x | c.y | c.z | c.w | d.x | d.y | d.z | d.w) != 0);
  }
  __syncthreads();

  const uint32_t tmem_base = sm.tmem_base;
  if (warp == MMA_WARP) {
    const bool issuer = elect_one();  // uniform loop, one elected lane issues: operands stay in uniform registers
    {
      const uint32_t idesc_s = sa::make_idesc_f16(BQ, BKV), idesc_d = sa::make_idesc_f16(BQ, D_N);
      mbar_wait(&sm.qbar, 0);
      const uint64_t dq = sa::make_sdesc_sw64(smem_u32(&sm.q[0][0]));
      const uint64_t dap = sa::make_sdesc_sw64(smem_u32(&sm.ap[0][0]));
      for (int it = 0; it < kt1 - kt0; ++it) {
        const int st = it % STAGES;
        mbar_wait(&sm.full[st], (it / STAGES) & 1);
        if (it >= 1) mbar_wait(&sm.sfree, (it - 1) & 1);
        tc_fence_after();
        const uint64_t dk = sa::make_sdesc_sw64(smem_u32(&sm.k[st][0][0]));
        const uint64_t dwe = make_sdesc_kmajor_plain(smem_u32(&sm.we[st][0]), 16, 128);
        const uint64_t dwo = make_sdesc_kmajor_plain(smem_u32(&sm.wo[st][0]), 16, 128);
        if (issuer) {
#pragma unroll
          for (int k = 0; k < DQ / 16; ++k) sa::umma_f16(tmem_base + COL_S, dq + 2 * k, dk + 2 * k, idesc_s, k);
#pragma unroll
          for (int k = 0; k < 2; ++k) sa::umma_f16(tmem_base + COL_DE, dap + 2 * k, dwe + 2 * k, idesc_d, k);
#pragma unroll
          for (int k = 0; k < 2; ++k) sa::umma_f16(tmem_base + COL_DO, dap + 2 * k, dwo + 2 * k, idesc_d, k);
          umma_commit(&sm.sfull);
        }
        __syncwarp();
      }
    }
    __syncwarp();
  } else if (warp == PRODUCER_WARP) {
    const bool issuer = elect_one();
    {
      if (issuer) {
        tma_prefetch_desc(&tmQK);
        tma_prefetch_desc(&tmV);
        mbar_arrive_expect_tx(&sm.qbar, BQ * DQ * 2);
        tma_load_2d(&sm.q[0][0], &tmQK, &sm.qbar, head * DQ, b * L + i0);
        tma_load_2d(&sm.q[64][0], &tmQK, &sm.qbar, head * DQ, b * L + i0 + 64);
      }
      const uint2* Ea = reinterpret_cast<const uint2*>(p.E16) + (size_t)head * p.e16_stride;
      const uint2* Eb = Ea + (size_t)H * p.e16_stride;
      for (int kt = kt0; kt < kt1; ++kt) {
        const int it = kt - kt0, s = it % STAGES;
        if (it >= STAGES) mbar_wait_backoff(&sm.empty[s], ((it / STAGES) - 1) & 1);
        const int j0 = kt * BKV;
        const int x0 = (j0 - i0 - (BQ - 1)) + (L - 1) + POS16_PADL;  // padded-table entry of window entry 0 (>= 1)
        const uint2* src_e = (x0 & 1) ? Eb + (x0 - 1) : Ea + x0;       // both sources start at an even entry:
        const uint2* src_o = (x0 & 1) ? Ea + (x0 + 1) : Eb + x0;       // 16-byte aligned
        if (issuer) {
          mbar_arrive_expect_tx(&sm.full[s], BKV * DQ * 2 + BKV * DVP * 2 + 2 * EW16 * 8);
          tma_load_2d(&sm.k[s][0][0], &tmQK, &sm.full[s], H * DQ + head * DQ, b * L + j0);
          tma_load_2d(&sm.v[s][0][0], &tmV, &sm.full[s], head * DVP, b * L + j0);
          sa::bulk_load_1d(&sm.we[s][0], src_e, EW16 * 8, &sm.full[s]);
          sa::bulk_load_1d(&sm.wo[s][0], src_o, EW16 * 8, &sm.full[s]);
        }
        __syncwarp();
      }
    }
    __syncwarp();
  } else {
    // ---------------- compute warps: 16 query rows each (TMEM lanes row0 + g and row0 + 8 + g)
    const int g = lane >> 2, qd = lane & 3;
    const int row0 = 32 * (warp & 3) + 16 * (warp >> 2);
    const int rl0 = row0 + g, rl1 = rl0 + 8;
    float o[2][4];
#pragma unroll
    for (int n = 0; n < 2; ++n) o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY;
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(row0) << 16);
    // rows g + 8 (a = row0 / 8 + 1) sit at v = 56 - row0 / 2 + 4 nt + qd, rows g four columns further
    const uint32_t dcol = 56 - (row0 >> 1);
#ifdef ZV_SA_TRACE  // measurement build: cycles of one warp per phase of the key-tile loop
    long long tr_w = 0, tr_l = 0, tr_m = 0, tr_v = 0, tr_e = 0;
    const long long tr_0 = clock64();
#define ZV_SA_TR(acc) { const long long c_ = clock64(); acc += c_ - tr_c; tr_c = c_; }
#else
#define ZV_SA_TR(acc)
#endif
    for (int kt = kt0; kt < kt1; ++kt) {
      const int it = kt - kt0, st = it % STAGES;
      const bool mk = sflag[kt] != 0;
#ifdef ZV_SA_TRACE
      long long tr_c = clock64();
#endif
      // ---- scores: S + D_even / D_odd from TMEM, then the accumulators go back to the MMA warp
      float s[8][4];
      mbar_wait(&sm.sfull, it & 1);
      tc_fence_after();
      ZV_SA_TR(tr_w)
      {
        uint32_t r[32];
#if defined(ZV_SA_ABL) && (ZV_SA_ABL & 2)  // measurement build: no score load (wrong results)
#pragma unroll
        for (int i = 0; i < 32; ++i) r[i] = __float_as_uint((float)((lane + i + kt) & 7));
        if (false)
#endif
        asm volatile(
            "tcgen05.ld.sync.aligned.16x256b.x8.b32 "
            "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
            "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
              "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
              "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
              "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
            : "r"(lane_addr + COL_S)
            : "memory");
        uint32_t te[16], te1[2];
#if defined(ZV_SA_ABL) && (ZV_SA_ABL & 1)  // measurement build: no position-term loads (wrong results)
#pragma unroll
        for (int i = 0; i < 16; ++i) te[i] = r[i];
        te1[0] = te1[1] = r[5];
#else
        tmem_ld_16x128b_x8(lane_addr + COL_DE + dcol, te);
        tmem_ld_16x128b_x1(lane_addr + COL_DE + dcol + 32, te1);
#endif
        tmem_ld_wait();
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          s[nt][0] = __uint_as_float(r[4 * nt]) + __uint_as_float(nt < 7 ? te[2 * (nt + 1)] : te1[0]);
          s[nt][2] = __uint_as_float(r[4 * nt + 2]) + __uint_as_float(te[2 * nt + 1]);
          s[nt][1] = __uint_as_float(r[4 * nt + 1]);
          s[nt][3] = __uint_as_float(r[4 * nt + 3]);
        }
        uint32_t to[16], to1[2];
#if defined(ZV_SA_ABL) && (ZV_SA_ABL & 1)
#pragma unroll
        for (int i = 0; i < 16; ++i) to[i] = r[16 + i];
        to1[0] = to1[1] = r[7];
#else
        tmem_ld_16x128b_x8(lane_addr + COL_DO + dcol, to);
        tmem_ld_16x128b_x1(lane_addr + COL_DO + dcol + 32, to1);
        tmem_ld_wait();
#endif
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          s[nt][1] += __uint_as_float(nt < 7 ? to[2 * (nt + 1)] : to1[0]);
          s[nt][3] += __uint_as_float(to[2 * nt + 1]);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.sfree);
      ZV_SA_TR(tr_l)
      // ---- key mask (padded keys -> -1000 log2 e, keys beyond the sequence -> -inf) and the row maxima
      if (mk) {
        const uint8_t* mrow = smask + kt * BKV;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          const uint8_t k0 = mrow[nt * 8 + 2 * qd], k1 = mrow[nt * 8 + 2 * qd + 1];
          if (k0) s[nt][0] = s[nt][2] = (k0 == 1 ? ATTN_MASKED : -INFINITY);
          if (k1) s[nt][1] = s[nt][3] = (k1 == 1 ? ATTN_MASKED : -INFINITY);
        }
      }
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) mx0 = fmax3(mx0, s[nt][0], s[nt][1]), mx1 = fmax3(mx1, s[nt][2], s[nt][3]);
      // Softmax is shift invariant and P is bf16 with fp32 sums: while no score exceeds the running reference by more
      // than 2^SPEC_LOG2 the reference stays as it is (no maximum reduction, no rescaling of O); otherwise -- always
      // for tile 0, whose reference is -inf -- the classic online-softmax step.
      if (!__all_sync(0xffffffffu, (mx0 <= m0 + SPEC_LOG2) && (mx1 <= m1 + SPEC_LOG2))) {
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
        const float sc0 = ex2_approx(m0 - mn0), sc1 = ex2_approx(m1 - mn1);
#pragma unroll
        for (int n = 0; n < 2; ++n) o[n][0] *= sc0, o[n][1] *= sc0, o[n][2] *= sc1, o[n][3] *= sc1;
        m0 = mn0, m1 = mn1;
      }
      // ---- P = exp2(S - m) (SOFT of every 4 on the FMA pipe, the rest on MUFU), O += P V  (V tile of this stage;
      //      its ones column carries the row sums)
      ZV_SA_TR(tr_m)
      mbar_wait(&sm.full[st], (it / STAGES) & 1);  // complete long ago (the MMAs waited for it); orders the V reads
      ZV_SA_TR(tr_v)
      const uint32_t vbase = smem_u32(&sm.v[st][0][0]);
#pragma unroll
      for (int kk = 0; kk < BKV / 16; ++kk) {
        uint32_t pa[4];
        pa[0] = pack_bf16x2(ex2_sel<SOFT, 0>(s[2 * kk][0] - m0), ex2_sel<SOFT, 1>(s[2 * kk][1] - m0));
        pa[1] = pack_bf16x2(ex2_sel<SOFT, 2>(s[2 * kk][2] - m1), ex2_sel<SOFT, 3>(s[2 * kk][3] - m1));
        pa[2] = pack_bf16x2(ex2_sel<SOFT, 0>(s[2 * kk + 1][0] - m0), ex2_sel<SOFT, 1>(s[2 * kk + 1][1] - m0));
        pa[3] = pack_bf16x2(ex2_sel<SOFT, 2>(s[2 * kk + 1][2] - m1), ex2_sel<SOFT, 3>(s[2 * kk + 1][3] - m1));
        uint32_t vb[4];
        // V rows are 32 bytes apart: SWIZZLE_32B (16-byte chunk ^= bit 2 of the row) keeps the 8 rows of an
        // ldmatrix 8 x 8 tile in 8 different bank groups
        const int vrow = kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
        sa::ldsm_x4_t(vb, vbase + vrow * (DVP * 2) + ((((lane >> 4) ^ (vrow >> 2)) & 1) << 4));
        sa::mma_bf16(o[0], pa, vb[0], vb[1]);
        sa::mma_bf16(o[1], pa, vb[2], vb[3]);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.empty[st]);
      ZV_SA_TR(tr_e)
    }
#ifdef ZV_SA_TRACE
    if (lane == 0 && (warp == 0 || warp == 5) && blockIdx.x == 1 && blockIdx.y == 1 && (blockIdx.z == 3 || blockIdx.z == 100))
      printf("[sa trace] z %d warp %d tiles %d: wait_S %lld  ld %lld  max %lld  wait_V %lld  exp_pv %lld  loop %lld  since_start %lld\n",
             (int)blockIdx.z, warp, kt1 - kt0, tr_w, tr_l, tr_m, tr_v, tr_e, tr_w + tr_l + tr_m + tr_v + tr_e,
             clock64() - tr_0);
#endif
    const int ia = i0 + rl0, ib = i0 + rl1;
    if (ksplit > 1) {
      // ---- key split: the unnormalised state of this key range (O incl. the row-sum column, reference maximum)
      float* pa = p.part + ((((size_t)b * H + head) * ksplit + sp) * L) * 17;
#pragma unroll
      for (int n = 0; n < 2; ++n) {
        const int col = n * 8 + 2 * qd;
        if (ia < L) pa[(size_t)ia * 17 + 1 + col] = o[n][0], pa[(size_t)ia * 17 + 2 + col] = o[n][1];
        if (ib < L) pa[(size_t)ib * 17 + 1 + col] = o[n][2], pa[(size_t)ib * 17 + 2 + col] = o[n][3];
      }
      if (qd == 0) {
        if (ia < L) pa[(size_t)ia * 17] = m0;
        if (ib < L) pa[(size_t)ib * 17] = m1;
      }
    } else {
      // ---- normalise, store: column dv = 12 of O (n-tile 1, columns 4..5 -> the lane with qd == 2) is the row sum
      const float l0 = __shfl_sync(0xffffffffu, o[1][0], (lane & ~3) | 2);
      const float l1 = __shfl_sync(0xffffffffu, o[1][2], (lane & ~3) | 2);
      const float inv0 = 1.f / l0, inv1 = 1.f / l1;
      const int vcol0 = head * p.dv;
#pragma unroll
      for (int n = 0; n < 2; ++n) {
        const int col = n * 8 + 2 * qd;
        if (col >= p.dv) continue;
        if (ia < L)
          *reinterpret_cast<uint32_t*>(p.out + ((size_t)b * L + ia) * p.ldo + vcol0 + col) =
              pack_bf16x2(o[n][0] * inv0, o[n][1] * inv0);
        if (ib < L)
          *reinterpret_cast<uint32_t*>(p.out + ((size_t)b * L + ib) * p.ldo + vcol0 + col) =
              pack_bf16x2(o[n][2] * inv1, o[n][3] * inv1);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) tmem_dealloc(tmem_base, sa2::TM_COLS);
}

// Merge of the key-split partial states: one thread per (sequence, head, query row).
__global__ void attn_sa_merge_kernel(const float* __restrict__ part, int nb, int H, int L, int ksplit, int dv,
                                     __nv_bfloat16* __restrict__ out, int ldo) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)nb * H * L) return;
  const int row = (int)(gid % L), head = (int)((gid / L) % H), b = (int)(gid / ((size_t)L * H));
  const float* pa = part + (((size_t)b * H + head) * ksplit * L + row) * 17;
  const size_t sstride = (size_t)L * 17;
  float m = -INFINITY;
  for (int s = 0; s < ksplit; ++s) m = fmaxf(m, pa[s * sstride]);
  float acc[13];
#pragma unroll
  for (int c = 0; c < 13; ++c) acc[c] = 0.f;
  for (int s = 0; s < ksplit; ++s) {
    const float w = ex2_approx(pa[s * sstride] - m);
#pragma unroll
    for (int c = 0; c < 13; ++c) acc[c] += w * pa[s * sstride + 1 + c];
  }
  const float inv = 1.f / acc[12];  // column dv = 12 of the padded V is the constant 1: the softmax row sum
  __nv_bfloat16* o = out + ((size_t)b * L + row) * ldo + head * dv;
  for (int c = 0; c < dv; ++c) o[c] = __float2bfloat16(acc[c] * inv);
}

// Host launcher; returns -1 if the shape is not supported (caller falls back to attn_kernel).
int attention_sa_tma(const AttnParams& p, cudaStream_t s) {
  // V must be in the padded layout (16 columns per head: 12 values, a column of ones, zeros): the TMA box start
  // has to be 16-byte aligned, and the ones column yields the softmax row sums from the P.V product
  if (p.dv != 12 || p.vhs != DVP || p.gate || p.heads != p.H || p.ldv % 8 || p.ldq % 8 || (p.ldo & 1)) return -1;
  const int n_tiles = (p.L + BKV - 1) / BKV;
  const int smem = (int)sizeof(Smem) + 1024 + n_tiles * BKV + ((n_tiles + 15) & ~15);
  if (smem > 100 * 1024) return -1;
  CUtensorMap tmQK, tmV;
  int rc = make_tmap(&tmQK, p.qkp, 0, (uint64_t)p.nb * p.L, (uint64_t)(2 * p.H * DQ + 4 * p.H), (uint64_t)p.ldq, BKV,
                     DQ, 64);
  if (rc) return rc;
  rc = make_tmap(&tmV, p.V, 0, (uint64_t)p.nb * p.L, (uint64_t)p.heads * DVP, (uint64_t)p.ldv, BKV, DVP, 0);
  if (rc) return rc;
  CUtensorMap tmV32;  // the same box with SWIZZLE_32B (attn_sa2_kernel)
  rc = make_tmap(&tmV32, p.V, 0, (uint64_t)p.nb * p.L, (uint64_t)p.heads * DVP, (uint64_t)p.ldv, BKV, DVP, 32);
  if (rc) return rc;
  static unsigned long long attr = 0;
  if (first_use_on_device(attr)) {
    cudaFuncSetAttribute(attn_sa_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    cudaFuncSetAttribute(attn_sa2_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    cudaFuncSetAttribute(attn_sa2_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  }
  dim3 grid((p.L + BQ - 1) / BQ, p.heads, p.nb);
  // ZV_SA_IMPL=1: the CUDA-core position term (A/B measurements); default: position term on the tensor cores.
  // ZV_SA_SOFT=2: half of the exponentials on the FMA pipe instead of MUFU -- measured SLOWER on B200 (nb = 128,
  // L = 1219: 0.441 ms with none, 0.455 / 0.500 / 0.521 ms with 1 / 2 / 3 of every 4; identical accuracy): the kernel
  // is bound by issue slots and the per-tile MMA -> TMEM -> release chain, not by the MUFU pipe.  Kept for A/B runs.
  static const int sa_impl = getenv("ZV_SA_IMPL") ? atoi(getenv("ZV_SA_IMPL")) : 2;
  static const int sa_soft = getenv("ZV_SA_SOFT") ? atoi(getenv("ZV_SA_SOFT")) : 0;
  const int smem2 = (int)sizeof(sa2::Smem) + 1024 + n_tiles * BKV + ((n_tiles + 15) & ~15);
  if (sa_impl == 2 && p.E16 && smem2 <= 100 * 1024) {
    // Key split for grids that leave SMs idle or end in a mostly empty wave.  Cost of a launch in key-tile times:
    // waves x (key tiles per CTA + 4 for a CTA's prologue: mask copy, operand A, TMEM, Q load) + 2 for the partial
    // stores and the merge kernel.  Large grids (B = 64) stay unsplit: halving a 5-tile CTA doubles its prologue share
    // (measured: nb = 128, L = 305 0.052 -> 0.109 ms when split).
    static int slots = 0;
    if (!slots) {
      int dev = 0, sms = 0;
      cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
      slots = 2 * sms;
    }
    static const int ks_env = getenv("ZV_SA_KSPLIT") ? atoi(getenv("ZV_SA_KSPLIT")) : 0;
    int ksplit = p.ksplit > 0 ? p.ksplit : ks_env;
    const long long base = (long long)grid.x * grid.y * grid.z;
    auto fits = [&](int k) {
      const int per = (n_tiles + k - 1) / k;
      return k == 1 || (p.part && per >= 2 && (long long)(k - 1) * per < n_tiles &&
                        (size_t)p.nb * p.H * k * p.L * 17 * sizeof(float) <= p.part_cap);
    };
    if (ksplit <= 0) {
      // automatic: long sequences only (every CTA keeps >= 8 key tiles: a single utterance of 8 s is a 15 us launch and
      // the extra merge launch costs more than the split saves -- measured) and a predicted gain of >= 15 %
      auto cost = [&](int k) {
        return (double)((base * k + slots - 1) / slots) * ((n_tiles + k - 1) / k + 4) + (k > 1 ? 2 : 0);
      };
      double best = 0.85 * cost(1);
      ksplit = 1;
      for (int k = 2; k <= 8; ++k) {
        if (!fits(k) || (n_tiles + k - 1) / k < 8) continue;
        if (cost(k) < best - 1e-9) best = cost(k), ksplit = k;
      }
    } else if (!fits(ksplit)) {
      ksplit = 1;
    }
    grid.x *= ksplit;
    if (sa_soft == 2) attn_sa2_kernel<2><<<grid, THREADS, smem2, s>>>(tmQK, tmV32, p, ksplit);
    else attn_sa2_kernel<0><<<grid, THREADS, smem2, s>>>(tmQK, tmV32, p, ksplit);
    int lrc = (int)cudaGetLastError();
    if (!lrc && ksplit > 1) {
      const size_t n = (size_t)p.nb * p.H * p.L;
      attn_sa_merge_kernel<<<(unsigned)((n + 127) / 128), 128, 0, s>>>(p.part, p.nb, p.H, p.L, ksplit, p.dv, p.out,
                                                                      p.ldo);
      lrc = (int)cudaGetLastError();
    }
    return lrc;
  }
  attn_sa_kernel<<<grid, THREADS, smem, s>>>(tmQK, tmV, p);
  return (int)cudaGetLastError();
}

}  // namespace zv
