// NonlinAttention value product on tcgen05 (zipformer.py:1499-1544 with the head-0 weights of
// RelPositionMultiheadAttentionWeights, :542,1529-1536):  O = softmax(q k^T + p.E) X * y,  X up to 384 wide.
//
// CTA = 128 queries of one sequence.  The scores are cheap (dq = 32) but every one needs the
// relative-position term and an exponential on the CUDA cores, while A.X is a 128 x 384 x L product:
//   reference maxima: S tiles (mma.sync fp16 + position term) of key tile 0 and of the tile under the
//            diagonal only -> a per-row reference m (nothing stored).  Softmax is shift invariant and P is bf16
//            (fp32 exponent range), so ANY finite m works as long as exp(S - m) cannot overflow;
//   value sweep: S of every tile, P = exp(S - m) as bf16 into a SWIZZLE_128B K-major shared-memory tile, and ONE
//            thread issues tcgen05.mma  O[128 x 384] += P[128 x 64] X[64 x 384]  with X tiles TMA-loaded as
//            MN-major operands (the row-major [key][channel] layout is used as is); O stays in TMEM for the
//            whole key loop -- no online-softmax rescaling -- and is normalised by the row sums at the end.
//   Every thread checks S - m <= 2^OVF_LOG2 on the way; if any score breaks that bound (66 nats above the
//   reference: not seen with real weights, covered by a unit test) the CTA repeats the work the exact way:
//   a full maxima sweep, then the value sweep again with the exact row maxima.
// 16 compute warps (8 row groups x 2 key halves of each 64-key tile) + an MMA warp (X tiles by TMA,
// tcgen05.mma issue) + a producer warp that streams the K tiles (SWIZZLE_64B boxes of the fp16 q|k|p rows)
// and the relative-position windows (1-D bulk copies) of all sweeps through a 4-stage full / empty mbarrier
// ring: the compute warps never meet at a CTA barrier inside the key loops.
#include <cuda.h>

#include "zv_common.cuh"
#include "zv_gemm.h"
#include "zv_kernels.h"

namespace zv {

namespace natc {

static constexpr int BQ = 128, BKV = 64, DQ = 32, EWIN = 192;
static constexpr int VSTAGES = 3, NVBOX = 6;  // up to 6 boxes of 64 channels = 384
static constexpr int KSTAGES = 4;
static constexpr float OVF_LOG2 = 96.0f;  // P <= 2^96: 1219 keys x |x| <= 2^10 still sums far below fp32 max
static constexpr float OVF_BOUND = 79228162514264337593543950336.0f;  // 2^96
static constexpr int COMPUTE_THREADS = 512, THREADS = COMPUTE_THREADS + 64;
static constexpr int MMA_WARP = COMPUTE_THREADS / 32, LOAD_WARP = MMA_WARP + 1;

struct alignas(16) Smem {  // the key mask that follows is read with 16-byte loads
  bf16 v[VSTAGES][NVBOX][BKV][64];  // MN-major SWIZZLE_128B boxes: [key][64 channels] = 8 KB each
  bf16 p[2][BQ][BKV];               // K-major SWIZZLE_128B: [query][64 keys]
  __half k[KSTAGES][BKV][DQ];       // SWIZZLE_64B boxes, 4 KB each
  __half q[BQ][DQ];                 // SWIZZLE_64B, two 64-row boxes
  float4 e[KSTAGES][EWIN];
  float red[2][BQ];                 // cross-warp (key half) reductions of max / sum
  uint64_t vfull[VSTAGES], vfree[VSTAGES], pfull[2], pfree[2], obar;
  uint64_t kfull[KSTAGES], kempty[KSTAGES], qbar;
  uint32_t tmem_base;
  uint32_t redo;  // a score exceeded the reference maximum by more than OVF_LOG2: repeat with exact maxima
  // followed by: uint8_t mask[n_tiles * BKV] (0 valid, 1 padded key, 2 beyond the sequence), uint8_t flag[n_tiles]
};

ZV_DEVINL void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
// byte offset of 16-byte chunk c of row r in a SWIZZLE_64B tile with 64-byte rows
ZV_DEVINL uint32_t sw64(int r, int c) { return (uint32_t)(r * 64 + ((c ^ ((r >> 1) & 3)) << 4)); }
// 1-D bulk copy global -> shared, completion counted on an mbarrier (bytes multiple of 16, 16-byte aligned)
ZV_DEVINL void bulk_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
ZV_DEVINL void mma_f16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// MN-major SWIZZLE_128B operand: 64 MN-elements (128 B) per K row, 8 K rows per 1024-byte swizzle atom
// (SBO), 64-element MN blocks LBO bytes apart  (cute mma_traits_sm100.hpp, UmmaDescriptor Major-MN, B128).
ZV_DEVINL uint64_t make_sdesc_mn_sw128(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(2) << 61;
  return d;
}
// kind::f16 instruction descriptor, A = bf16 K-major, B = bf16 MN-major (bit 16), D = fp32
ZV_DEVINL uint32_t make_idesc_bf16_bmn(uint32_t M, uint32_t N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// Scores of this warp's 16 rows x 32 keys of the staged tile: s[nt4][0..1] row g, [2..3] row g + 8.
// Returns the two row maxima through mx0 / mx1.
// MASKED is a template parameter so that the caller takes ONE warp-uniform branch per tile instead of
// ~45 predicated mask instructions that occupy issue slots on mask-free tiles.
// EXP: the value sweep -- the MMA accumulators start at minus the reference (scores are in log2 units), so a
// finished score is the exponent of P = exp2(S - ms); the exponentials are taken right there and interleave with the
// position FMAs of the next block; mx0 / mx1 then hold the largest P of the two rows.
template <bool MASKED, bool EXP>
ZV_DEVINL void score_tile(const Smem& sm, int buf, const uint8_t* mrow, int kh, int lane, int rl0,
                          const uint32_t (&qa)[2][4], const float (&pq0)[4], const float (&pq1)[4], float (&s)[4][4],
                          float& mx0, float& mx1, float ms0 = 0.f, float ms1 = 0.f) {
  const int qd = lane & 3;
  const uint32_t kbase = smem_u32(&sm.k[buf][0][0]);
#pragma unroll
  for (int n4 = 0; n4 < 4; ++n4) {
    const int nt = kh * 4 + n4;
    s[n4][0] = s[n4][1] = EXP ? -ms0 : 0.f, s[n4][2] = s[n4][3] = EXP ? -ms1 : 0.f;
    uint32_t kb[4];
    ldsm_x4(kb, kbase + sw64(nt * 8 + (lane & 7), lane >> 3));
    mma_f16(s[n4], qa[0], kb[0], kb[1]);
    mma_f16(s[n4], qa[1], kb[2], kb[3]);
  }
  const float4* ew = &sm.e[buf][kh * 32 + 2 * qd - rl0 + (BQ - 1)];
  float4 ep0 = ew[-8], ep1 = ew[-7];
  mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
  for (int n4 = 0; n4 < 4; ++n4) {
    const float4 ec0 = ew[n4 * 8], ec1 = ew[n4 * 8 + 1];
    float a0 = s[n4][0] + pq0[0] * ec0.x + pq0[1] * ec0.y + pq0[2] * ec0.z + pq0[3] * ec0.w;
    float a1 = s[n4][1] + pq0[0] * ec1.x + pq0[1] * ec1.y + pq0[2] * ec1.z + pq0[3] * ec1.w;
    float b0 = s[n4][2] + pq1[0] * ep0.x + pq1[1] * ep0.y + pq1[2] * ep0.z + pq1[3] * ep0.w;
    float b1 = s[n4][3] + pq1[0] * ep1.x + pq1[1] * ep1.y + pq1[2] * ep1.z + pq1[3] * ep1.w;
    if (MASKED) {
      const uint8_t k0 = mrow[kh * 32 + n4 * 8 + 2 * qd], k1 = mrow[kh * 32 + n4 * 8 + 2 * qd + 1];
      if (k0) a0 = (k0 == 1 ? ATTN_MASKED - (EXP ? ms0 : 0.f) : -INFINITY), b0 = (k0 == 1 ? ATTN_MASKED - (EXP ? ms1 : 0.f) : -INFINITY);
      if (k1) a1 = (k1 == 1 ? ATTN_MASKED - (EXP ? ms0 : 0.f) : -INFINITY), b1 = (k1 == 1 ? ATTN_MASKED - (EXP ? ms1 : 0.f) : -INFINITY);
    }
    if (EXP) a0 = ex2_approx(a0), a1 = ex2_approx(a1), b0 = ex2_approx(b0), b1 = ex2_approx(b1);
    s[n4][0] = a0, s[n4][1] = a1, s[n4][2] = b0, s[n4][3] = b1;
    mx0 = fmaxf(mx0, fmaxf(a0, a1)), mx1 = fmaxf(mx1, fmaxf(b0, b1));
    ep0 = ec0, ep1 = ec1;
  }
}

}  // namespace natc

using namespace natc;

__global__ void __launch_bounds__(natc::THREADS, 1)
attn_na_tc_kernel(const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmQK,
                  const __grid_constant__ CUtensorMap tmG, AttnParams p, int nvbox, int n_mma0, int n_mma1, int dv_slice) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // align inside the shared window (pointer + integer keeps the shared address space: LDS/STS, not generic LD/ST)
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.y;
  const int i0 = blockIdx.x * BQ;
  // value-channel slice of this CTA (small grids: blockIdx.z slices of nvbox boxes each; nvbox, n_mma* are per slice)
  const int zc = blockIdx.z * dv_slice;
  const int L = p.L, H = p.H, R = 2 * L - 1;
  const int n_tiles = (L + BKV - 1) / BKV;
  const int kd = min((i0 + BQ / 2) / BKV, n_tiles - 1);  // key tile under the middle of this CTA's diagonal
  const __half* qkp = reinterpret_cast<const __half*>(p.qkp) + (size_t)b * L * p.ldq;  // head 0
  const float4* E = reinterpret_cast<const float4*>(p.E);
  const uint8_t* mask = p.mask ? p.mask + (size_t)b * L : nullptr;
  const uint32_t v_stage_bytes = (uint32_t)nvbox * BKV * 64 * 2;
  uint8_t* smask = reinterpret_cast<uint8_t*>(&sm) + sizeof(Smem);
  uint8_t* sflag = smask + n_tiles * BKV;

  if (tid == 0) {
    for (int i = 0; i < VSTAGES; ++i) mbar_init(&sm.vfull[i], 1), mbar_init(&sm.vfree[i], 1);
    for (int i = 0; i < 2; ++i) mbar_init(&sm.pfull[i], COMPUTE_THREADS / 32), mbar_init(&sm.pfree[i], 1);
    for (int i = 0; i < KSTAGES; ++i) mbar_init(&sm.kfull[i], 1), mbar_init(&sm.kempty[i], COMPUTE_THREADS / 32);
    mbar_init(&sm.obar, 1);
    mbar_init(&sm.qbar, 1);
    sm.redo = (p.dbg & 4) ? 1u : 0u;
    fence_barrier_init();
  }
  if (warp == MMA_WARP) tmem_alloc(&sm.tmem_base, 512);
  // window entries outside the table are never copied: keep them finite
  for (int i = tid; i < KSTAGES * EWIN; i += THREADS) sm.e[i / EWIN][i % EWIN] = make_float4(0.f, 0.f, 0.f, 0.f);
  // key-padding mask of the whole sequence (0 valid, 1 padded, 2 beyond the sequence) and per-tile flags, once
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

  if (warp == LOAD_WARP) {
    // ---------------- producer: Q once, then K tile + position window of every key tile of every sweep
    // (the whole warp walks the loops, one elected lane issues: operands stay in uniform registers, see elect_one)
    const bool issuer = elect_one();
    int g = 0;
    auto load_k = [&](int kt) {
      const int st = g % KSTAGES;
      if (g >= KSTAGES) mbar_wait_backoff(&sm.kempty[st], ((g / KSTAGES) - 1) & 1);
      const int j0 = kt * BKV;
      const int rr0 = (j0 - i0 - (BQ - 1)) + (L - 1);  // table row of window entry 0
      const int lo = max(rr0, 0), hi = min(rr0 + EWIN, R);
      const uint32_t ebytes = hi > lo ? (uint32_t)(hi - lo) * 16u : 0u;
      if (issuer) {
        mbar_arrive_expect_tx(&sm.kfull[st], BKV * DQ * 2 + ebytes);
        tma_load_2d(&sm.k[st][0][0], &tmQK, &sm.kfull[st], H * DQ, b * L + j0);
        if (ebytes) bulk_load_1d(&sm.e[st][lo - rr0], E + lo, ebytes, &sm.kfull[st]);
      }
      __syncwarp();
      ++g;
    };
    {
      if (issuer) {
        tma_prefetch_desc(&tmQK);
        mbar_arrive_expect_tx(&sm.qbar, BQ * DQ * 2);
        tma_load_2d(&sm.q[0][0], &tmQK, &sm.qbar, 0, b * L + i0);
        tma_load_2d(&sm.q[64][0], &tmQK, &sm.qbar, 0, b * L + i0 + 64);
      }
      load_k(0);
      if (kd != 0) load_k(kd);
      for (int kt = 0; kt < n_tiles; ++kt) load_k(kt);
    }
    __syncwarp();
    named_bar_sync(2, THREADS);
    if (sm.redo != 0) {
      for (int kt = 0; kt < 2 * n_tiles; ++kt) load_k(kt < n_tiles ? kt : kt - n_tiles);
    }
    __syncwarp();
  } else if (warp == MMA_WARP) {
    // ---------------- producer: X tiles by TMA, P.X by tcgen05.mma.  c counts tiles over both value sweeps.
    const bool issuer = elect_one();  // uniform loops, one elected lane issues
    // v counts the uses of the X ring (value tiles, then the two halves of the gate tile of every sweep)
    auto load_v = [&](int v, int kt) {
      const int c = v;
      const int st = c % VSTAGES;
      if (c >= VSTAGES) mbar_wait(&sm.vfree[st], ((c / VSTAGES) - 1) & 1);  // tile c - VSTAGES consumed
      if (p.dbg & 1) {
        if (issuer) mbar_arrive(&sm.vfull[st]);
        return;
      }
      if (issuer) {
        mbar_arrive_expect_tx(&sm.vfull[st], v_stage_bytes);
        for (int cc = 0; cc < nvbox; ++cc)
          tma_load_2d(&sm.v[st][cc][0][0], &tmV, &sm.vfull[st], zc + cc * 64, b * L + kt * BKV);
      }
      __syncwarp();
    };
    // The gate tile y[128 queries][<= 384] of the epilogue rides behind the value tiles in the same ring: its two
    // halves (3 boxes of 128 rows x 64 channels = one 48 KB stage each) are uses n_tiles and n_tiles + 1 of the
    // sweep, so they arrive while the last two key tiles are still being worked on.
    auto load_gate = [&](int v, int part) {
      const int st = v % VSTAGES;
      if (v >= VSTAGES) mbar_wait(&sm.vfree[st], ((v / VSTAGES) - 1) & 1);
      const int nbox = min(3, nvbox - 3 * part);
      if (!p.gate || nbox <= 0) {
        if (issuer) mbar_arrive(&sm.vfull[st]);
        return;
      }
      if (issuer) {
        mbar_arrive_expect_tx(&sm.vfull[st], (uint32_t)nbox * BQ * 128);
        for (int cc = 0; cc < nbox; ++cc)
          tma_load_2d(reinterpret_cast<uint8_t*>(&sm.v[st][0][0][0]) + cc * (BQ * 128), &tmG, &sm.vfull[st],
                      zc + (3 * part + cc) * 64, b * L + i0);
      }
      __syncwarp();
    };
    const uint32_t id0 = make_idesc_bf16_bmn(BQ, n_mma0), id1 = make_idesc_bf16_bmn(BQ, max(n_mma1, 16));
    // c0 / v0: P-buffer and X-ring use counters at entry
    auto value_sweep = [&](int c0, int v0) {
      auto issue = [&](int idx) {
        if (idx < n_tiles) load_v(v0 + idx, idx);
        else if (idx < n_tiles + 2) load_gate(v0 + idx, idx - n_tiles);
      };
      issue(0);
      issue(1);
      for (int kt = 0; kt < n_tiles; ++kt) {
        const int c = c0 + kt, v = v0 + kt;
        issue(kt + 2);
        const int st = v % VSTAGES, pb = c & 1;
        mbar_wait(&sm.vfull[st], (v / VSTAGES) & 1);
        mbar_wait(&sm.pfull[pb], (c >> 1) & 1);
        tc_fence_after();
        const uint64_t da = make_sdesc_sw128(smem_u32(&sm.p[pb][0][0]));
        const uint32_t vaddr = smem_u32(&sm.v[st][0][0][0]);
        if (issuer) {
#pragma unroll
          for (int k = 0; k < ((p.dbg & 2) ? 0 : BKV / 16); ++k) {
            const uint64_t db0 = make_sdesc_mn_sw128(vaddr + k * 2048, BKV * 128, 1024);
            umma_bf16(tmem_base, da + 2 * k, db0, id0, (kt | k) != 0 ? 1u : 0u);
            if (n_mma1 > 0) {
              const uint64_t db1 =
                  make_sdesc_mn_sw128(vaddr + (n_mma0 / 64) * (BKV * 128) + k * 2048, BKV * 128, 1024);
              umma_bf16(tmem_base + n_mma0, da + 2 * k, db1, id1, (kt | k) != 0 ? 1u : 0u);
            }
          }
          umma_commit(&sm.vfree[st]);
          umma_commit(&sm.pfree[pb]);
          if (kt == n_tiles - 1) umma_commit(&sm.obar);
        }
        __syncwarp();
      }
    };
    {
      if (issuer) {
        tma_prefetch_desc(&tmV);
        if (p.gate) tma_prefetch_desc(&tmG);
      }
      value_sweep(0, 0);
    }
    __syncwarp();
    named_bar_sync(2, THREADS);
    if (sm.redo != 0) {
      // the gate halves of the abandoned sweep are not consumed by any MMA: retire them by hand
      for (int v = n_tiles; v < n_tiles + 2; ++v) {
        mbar_wait(&sm.vfull[v % VSTAGES], (v / VSTAGES) & 1);
        if (issuer) mbar_arrive(&sm.vfree[v % VSTAGES]);
        __syncwarp();
      }
      value_sweep(n_tiles, n_tiles + 2);
    }
    __syncwarp();
  } else {
    // ---------------- compute warps
    const int rg = warp & 7, kh = warp >> 3;
    const int g = lane >> 2, qd = lane & 3;
    const int rl0 = rg * 16 + g, rl1 = rl0 + 8;
    auto cbar = [] { named_bar_sync(1, COMPUTE_THREADS); };
    float pq0[4], pq1[4];
    {
      const __half* pa = qkp + (size_t)min(i0 + rl0, L - 1) * p.ldq + 2 * H * DQ;
      const __half* pb = qkp + (size_t)min(i0 + rl1, L - 1) * p.ldq + 2 * H * DQ;
#pragma unroll
      for (int c = 0; c < 4; ++c) pq0[c] = __half2float(pa[c]), pq1[c] = __half2float(pb[c]);
    }
    uint32_t qa[2][4];
    float s[4][4];
    float m0, m1;
    mbar_wait(&sm.qbar, 0);
    {
      const uint32_t qbase = smem_u32(&sm.q[0][0]);
#pragma unroll
      for (int ks = 0; ks < 2; ++ks) ldsm_x4(qa[ks], qbase + sw64(rg * 16 + (lane & 15), ks * 2 + (lane >> 4)));
    }
    // the K / E ring is consumed tile by tile: wait full, read into registers, release
    int gt = 0;  // running tile counter over both sweeps
    auto ring_release = [&](int st) {
      __syncwarp();
      if (lane == 0) mbar_arrive(&sm.kempty[st]);
    };
    // maxima of the listed key tiles -> m0 / m1 of this thread's two rows, agreed between the two key halves
    auto maxima = [&](bool full) {
      m0 = -INFINITY, m1 = -INFINITY;
      const int count = full ? n_tiles : (kd != 0 ? 2 : 1);
      for (int it = 0; it < count; ++it, ++gt) {
        const int kt = full ? it : (it == 0 ? 0 : kd);
        const int buf = gt % KSTAGES;
        mbar_wait(&sm.kfull[buf], (gt / KSTAGES) & 1);
        float mx0, mx1;
        if (sflag[kt] != 0) score_tile<true, false>(sm, buf, smask + kt * BKV, kh, lane, rl0, qa, pq0, pq1, s, mx0, mx1);
        else score_tile<false, false>(sm, buf, smask + kt * BKV, kh, lane, rl0, qa, pq0, pq1, s, mx0, mx1);
        ring_release(buf);
        m0 = fmaxf(m0, mx0), m1 = fmaxf(m1, mx1);
      }
      m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 1)), m0 = fmaxf(m0, __shfl_xor_sync(0xffffffffu, m0, 2));
      m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 1)), m1 = fmaxf(m1, __shfl_xor_sync(0xffffffffu, m1, 2));
      if (qd == 0) sm.red[kh][rl0] = m0, sm.red[kh][rl1] = m1;
      cbar();
      m0 = fmaxf(sm.red[0][rl0], sm.red[1][rl0]), m1 = fmaxf(sm.red[0][rl1], sm.red[1][rl1]);
      cbar();
    };
    // P = exp(S - m) of every tile -> shared memory -> tcgen05.mma; c0 = running value-tile counter at entry.
    // Returns true if some score of this thread exceeded its reference by more than OVF_LOG2.
    float l0 = 0.f, l1 = 0.f;
    auto value_sweep = [&](int c0) -> bool {
      const float ms0 = m0, ms1 = m1;  // log2 units
      bool ovf = false;
      l0 = 0.f, l1 = 0.f;
      for (int kt = 0; kt < n_tiles; ++kt, ++gt) {
        const int c = c0 + kt;
        const int buf = gt % KSTAGES, pb = c & 1;
        mbar_wait(&sm.kfull[buf], (gt / KSTAGES) & 1);
        float mx0, mx1;
        if (sflag[kt] != 0)
          score_tile<true, true>(sm, buf, smask + kt * BKV, kh, lane, rl0, qa, pq0, pq1, s, mx0, mx1, ms0, ms1);
        else
          score_tile<false, true>(sm, buf, smask + kt * BKV, kh, lane, rl0, qa, pq0, pq1, s, mx0, mx1, ms0, ms1);
        ring_release(buf);
        ovf |= !(fmaxf(mx0, mx1) <= OVF_BOUND);  // largest P of the tile against 2^OVF_LOG2 (also catches NaN)
        if (c >= 2) {  // the MMAs of tile c-2 have read p[pb]
          if (lane == 0) mbar_wait(&sm.pfree[pb], ((c >> 1) - 1) & 1);
          __syncwarp();
        }
        uint8_t* prow0 = reinterpret_cast<uint8_t*>(&sm.p[pb][rl0][0]);
        uint8_t* prow1 = reinterpret_cast<uint8_t*>(&sm.p[pb][rl1][0]);
#pragma unroll
        for (int n4 = 0; n4 < 4; ++n4) {
          const float a0 = s[n4][0], a1 = s[n4][1], b0 = s[n4][2], b1 = s[n4][3];
          l0 += a0 + a1, l1 += b0 + b1;
          const int chunk = kh * 4 + n4;  // 16-byte chunk = 8 keys
          *reinterpret_cast<uint32_t*>(prow0 + ((chunk ^ (rl0 & 7)) << 4) + qd * 4) = pack_bf16x2(a0, a1);
          *reinterpret_cast<uint32_t*>(prow1 + ((chunk ^ (rl1 & 7)) << 4) + qd * 4) = pack_bf16x2(b0, b1);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.pfull[pb]);  // one arrival per warp
      }
      return ovf;
    };
    // ---- reference maxima from key tile 0 (always holds valid keys) and the tile under the diagonal
    maxima(false);
    // ---- value sweep against the reference
    if (value_sweep(0)) sm.redo = 1u;
    named_bar_sync(2, THREADS);
    const bool redo = sm.redo != 0;
    if (redo) {  // exact two-sweep repeat, once the MMAs of the abandoned sweep have drained
      mbar_wait(&sm.obar, 0);
      maxima(true);
      value_sweep(n_tiles);
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1), l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1), l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    if (qd == 0) sm.red[kh][rl0] = l0, sm.red[kh][rl1] = l1;
    cbar();
    // ---- epilogue: O / l * gate -> bf16.  Warp w reads TMEM lane quarter w % 4, column slice w / 4.
    const int q4 = warp & 3, cs = warp >> 2;
    const int row = q4 * 32 + lane;
    const int i = i0 + row;
    const int ncols = nvbox * 64;
    const size_t grow = (size_t)b * L + min(i, L - 1);
    // the gate tile sits in the two X-ring stages behind the last value tile (SWIZZLE_128B boxes of 128 rows)
    const int vg = (redo ? n_tiles + 2 : 0) + n_tiles;
    const uint8_t* gst[2];
#pragma unroll
    for (int part = 0; part < 2; ++part) {
      const int v = vg + part;
      mbar_wait(&sm.vfull[v % VSTAGES], (v / VSTAGES) & 1);
      gst[part] = reinterpret_cast<const uint8_t*>(&sm.v[v % VSTAGES][0][0][0]);
    }
    mbar_wait(&sm.obar, redo ? 1 : 0);
    tc_fence_after();
    const float inv = 1.0f / (sm.red[0][row] + sm.red[1][row]);
#pragma unroll
    for (int it = 0; it < 3; ++it) {
      const int c0 = cs * 32 + it * 128;
      if (c0 >= ncols) break;
      uint32_t raw[32];
      tmem_ld32(tmem_base + (static_cast<uint32_t>(q4 * 32) << 16) + c0, raw);
      tmem_ld_wait();
      if (i < L && zc + c0 < p.dv) {
        bf16* op = p.out + grow * p.ldo + zc + c0;
#pragma unroll
        for (int j4 = 0; j4 < 4; ++j4) {
          const int jj = j4 * 8;
          if (zc + c0 + jj >= p.dv) break;
          float v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) v[u] = __uint_as_float(raw[jj + u]) * inv;
          uint4 g4 = make_uint4(0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u);  // bf16 1.0 pairs
          if (p.gate) {
            const int col = c0 + jj, box = col >> 6;
            g4 = *reinterpret_cast<const uint4*>(gst[box >= 3] + (box % 3) * (BQ * 128) + row * 128 +
                                                 ((((col & 63) >> 3) ^ (row & 7)) << 4));
          }
          v[0] *= bf16_lo(g4.x), v[1] *= bf16_hi(g4.x), v[2] *= bf16_lo(g4.y), v[3] *= bf16_hi(g4.y);
          v[4] *= bf16_lo(g4.z), v[5] *= bf16_hi(g4.z), v[6] *= bf16_lo(g4.w), v[7] *= bf16_hi(g4.w);
          *reinterpret_cast<uint4*>(op + jj) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                                          pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == MMA_WARP) tmem_dealloc(tmem_base, 512);
}

// Host launcher; returns -1 if the shape is not supported by this kernel (caller falls back to attn_kernel).
int attention_na_tc(const AttnParams& p, cudaStream_t s) {
  if (p.heads != 1 || p.dv % 8 || p.dv > 384 || p.ldv % 8 || p.ldo % 8 || (p.gate && p.ldg % 8)) return -1;
  int nvbox = (p.dv + 63) / 64;
  // Single utterances: a handful of (query tile, sequence) CTAs each pull every 48 KB value tile through one SM.
  // Slices of 128 value channels (blockIdx.z) recompute the scores but spread the A.X product, the value traffic and
  // the epilogue over 2-3 x as many SMs.  ZV_NA_DVSLICE=0 switches it off.
  static const int slice_on = getenv("ZV_NA_DVSLICE") ? atoi(getenv("ZV_NA_DVSLICE")) : 1;
  int slices = 1;
  if (slice_on && nvbox > 2 && (long long)((p.L + BQ - 1) / BQ) * p.nb * ((nvbox + 1) / 2) <= 148) {
    slices = (nvbox + 1) / 2;
    nvbox = 2;
  }
  const int ncols = nvbox * 64;
  const int n0 = ncols > 256 ? 256 : ncols, n1 = ncols - n0;
  CUtensorMap tmV;
  int rc = make_tmap(&tmV, p.V, 0, (uint64_t)p.nb * p.L, (uint64_t)p.dv, (uint64_t)p.ldv, BKV, 64, 128);
  if (rc) return rc;
  if (p.ldq % 8) return -1;
  CUtensorMap tmQK;  // fp16 q | k | p rows: 64-row x 32-column boxes (q of head 0 at column 0, k at column H * 32)
  rc = make_tmap(&tmQK, p.qkp, 0, (uint64_t)p.nb * p.L, (uint64_t)(2 * p.H * DQ + 4 * p.H), (uint64_t)p.ldq, BKV, DQ, 64);
  if (rc) return rc;
  CUtensorMap tmG = tmV;  // gate rows as 128 x 64 boxes (not read by the kernel without a gate)
  if (p.gate) {
    rc = make_tmap(&tmG, p.gate, 0, (uint64_t)p.nb * p.L, (uint64_t)p.dv, (uint64_t)p.ldg, BQ, 64, 128);
    if (rc) return rc;
  }
  static unsigned long long attr = 0;
  const int n_tiles = (p.L + BKV - 1) / BKV;
  const int smem = (int)sizeof(Smem) + 1024 + n_tiles * BKV + ((n_tiles + 15) & ~15);
  if (smem > 227 * 1024) return -1;  // sequence too long for the in-kernel mask copy: mma.sync kernel instead
  if (first_use_on_device(attr)) {
    cudaFuncSetAttribute(attn_na_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  }
  dim3 grid((p.L + BQ - 1) / BQ, p.nb, slices);
  attn_na_tc_kernel<<<grid, THREADS, smem, s>>>(tmV, tmQK, tmG, p, nvbox, n0, n1, ncols);
  return (int)cudaGetLastError();
}

}  // namespace zv
