// Fused relative-position attention (zipformer.py:1149-1306 + 1359-1396 / 1499-1544):
//   S = q k^T + (p . E[L-1-i+j]),  key-padded -> -1000,  A = softmax_j(S),  O = A V  (* gate)
// flash-style: the L x L matrix is never materialised; the score tile is recomputed for each of
// the three consumers of A (NonlinAttention head 0, SelfAttention 1 and 2).
// Block = 128 queries of one (sequence, head, value chunk); 8 warps x 16 query rows; key tiles of
// 64 double-buffered with cp.async; QK^T and PV on bf16 tensor-core MMAs with fp32 accumulation,
// the position term on CUDA cores from a 191-entry window of the projected table in shared memory.
#include "zv_common.cuh"
#include "zv_kernels.h"

namespace zv {

static constexpr int BQ = 128, BKV = 64, DQ = 32, QP = DQ + 8, EWIN = 192;

ZV_DEVINL void cp_async16(void* dst, const void* src, bool valid) {
  const int sz = valid ? 16 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(sz) : "memory");
}
ZV_DEVINL void cp_async8(void* dst, const void* src, bool valid) {
  const int sz = valid ? 8 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(sz) : "memory");
}
ZV_DEVINL void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
ZV_DEVINL void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
ZV_DEVINL void ldsm_x4(uint32_t (&r)[4], const void* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_u32(p)));
}
ZV_DEVINL void ldsm_x4_t(uint32_t (&r)[4], const void* p) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(smem_u32(p)));
}
ZV_DEVINL void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

ZV_DEVINL void mma_f16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int DV>
struct AttnSmem {
  bf16 q[BQ][QP];
  bf16 k[3][BKV][QP];
  bf16 v[3][BKV][DV + 8];
  float4 e[3][EWIN];
  // followed by: uint8_t mask[n_tiles * BKV] (0 valid, 1 padded key, 2 beyond the sequence) and
  // uint8_t flag[n_tiles] (!= 0 if the tile holds masked keys -> slow masking path)
};

template <int DV>
__global__ void __launch_bounds__(256, (DV <= 16) ? 2 : 1) attn_kernel(AttnParams p, int chunks, int dv_valid) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  AttnSmem<DV>& sm = *reinterpret_cast<AttnSmem<DV>*>(smem_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.z;
  const int head = blockIdx.y / chunks, chunk = blockIdx.y % chunks;
  const int i0 = blockIdx.x * BQ;
  const int L = p.L, H = p.H;
  const int R = 2 * L - 1;
  const int vcol0 = head * p.dv + chunk * DV;  // first value / output column of this block
  const bf16* qkp = p.qkp + (size_t)b * L * p.ldq;
  const bf16* V = p.V + (size_t)b * L * p.ldv + head * (p.vhs ? p.vhs : p.dv) + chunk * DV;
  const float4* E = reinterpret_cast<const float4*>(p.E) + (size_t)head * R;
  const uint8_t* mask = p.mask ? p.mask + (size_t)b * L : nullptr;
  const int n_tiles = (L + BKV - 1) / BKV;
  uint8_t* smask = smem_raw + sizeof(AttnSmem<DV>);
  uint8_t* sflag = smask + n_tiles * BKV;
  // key-padding mask of the whole sequence and per-tile flags, once
  for (int j = tid; j < n_tiles * BKV; j += 256) smask[j] = (j < L) ? (mask ? mask[j] : 0) : 2;
  __syncthreads();
  for (int t = tid; t < n_tiles; t += 256) {
    const uint4* mp = reinterpret_cast<const uint4*>(smask + t * BKV);
    const uint4 a = mp[0], b4 = mp[1], c = mp[2], d = mp[3];
    sflag[t] = ((a.x | a.y | a.z | a.w | b4.x | b4.y | b4.z | b4.w | c.x | c.y | c.z | c.w | d.x | d.y | d.z | d.w) != 0);
  }

  // zero the value tiles once (columns >= dv_valid are never written by the loads)
  if (dv_valid < DV) {
    uint32_t* z = reinterpret_cast<uint32_t*>(&sm.v[0][0][0]);
    for (int i = tid; i < (int)(sizeof(sm.v) / 4); i += 256) z[i] = 0;
    __syncthreads();
  }
  // ---- Q tile (once)
  for (int c = tid; c < BQ * (DQ / 8); c += 256) {
    const int r = c / (DQ / 8), cc = c % (DQ / 8);
    const int i = i0 + r;
    cp_async16(&sm.q[r][cc * 8], qkp + (size_t)min(i, L - 1) * p.ldq + head * DQ + cc * 8, i < L);
  }
  auto load_tile = [&](int kt, int buf) {
    const int j0 = kt * BKV;
    {  // K: 64 rows x 4 chunks of 16 B
      const int r = tid >> 2, cc = tid & 3;
      const int j = j0 + r;
      cp_async16(&sm.k[buf][r][cc * 8], qkp + (size_t)min(j, L - 1) * p.ldq + H * DQ + head * DQ + cc * 8, j < L);
    }
    if (DV >= 32) {
      constexpr int CPR = DV / 8;  // 16-byte chunks per row
      for (int c = tid; c < BKV * CPR; c += 256) {
        const int r = c / CPR, cc = c % CPR;
        const int j = j0 + r;
        cp_async16(&sm.v[buf][r][cc * 8], V + (size_t)min(j, L - 1) * p.ldv + cc * 8, j < L);
      }
    } else {
      const int gpr = dv_valid / 4;  // 8-byte granules per row
      for (int c = tid; c < BKV * gpr; c += 256) {
        const int r = c / gpr, cc = c % gpr;
        const int j = j0 + r;
        cp_async8(&sm.v[buf][r][cc * 4], V + (size_t)min(j, L - 1) * p.ldv + cc * 4, j < L);
      }
    }
    if (tid < EWIN) {
      const int rr = (j0 - i0 - (BQ - 1)) + (L - 1) + tid;  // table row of window entry tid
      const bool ok = rr >= 0 && rr < R;
      cp_async16(&sm.e[buf][tid], E + (ok ? rr : 0), ok);
    }
  };
  load_tile(0, 0);
  cp_async_commit();
  if (n_tiles > 1) load_tile(1, 1);
  cp_async_commit();

  // per-thread rows and position queries
  const int g = lane >> 2, qd = lane & 3;
  const int rl0 = warp * 16 + g, rl1 = rl0 + 8;
  float pq0[4], pq1[4];
  {
    const int ia = min(i0 + rl0, L - 1), ib = min(i0 + rl1, L - 1);
    const __half* pa = reinterpret_cast<const __half*>(qkp + (size_t)ia * p.ldq + 2 * H * DQ + head * 4);
    const __half* pb = reinterpret_cast<const __half*>(qkp + (size_t)ib * p.ldq + 2 * H * DQ + head * 4);
#pragma unroll
    for (int c = 0; c < 4; ++c) pq0[c] = __half2float(pa[c]), pq1[c] = __half2float(pb[c]);
  }
  float o[DV / 8][4];
#pragma unroll
  for (int n = 0; n < DV / 8; ++n) o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  uint32_t qa[2][4];

  for (int kt = 0; kt < n_tiles; ++kt) {
    // three-stage ring, one barrier per tile: after it every warp has finished tile kt - 1, whose
    // buffer is the one tile kt + 2 is loaded into
    const int buf = kt % 3;
    cp_async_wait<1>();
    __syncthreads();
    if (kt + 2 < n_tiles) load_tile(kt + 2, (kt + 2) % 3);
    cp_async_commit();
    const uint8_t* mrow = smask + kt * BKV;
    if (kt == 0) {
#pragma unroll
      for (int ks = 0; ks < 2; ++ks)
        ldsm_x4(qa[ks], &sm.q[warp * 16 + (lane & 15)][ks * 16 + (lane >> 4) * 8]);
    }
    // ---- S = Q K^T
    float s[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
      uint32_t kb[4];
      ldsm_x4(kb, &sm.k[buf][nt * 8 + (lane & 7)][(lane >> 3) * 8]);
      mma_f16(s[nt], qa[0], kb[0], kb[1]);  // q, k, p are IEEE fp16 (attention in_proj epilogue)
      mma_f16(s[nt], qa[1], kb[2], kb[3]);
    }
    // ---- position term, masks, running max.  Rows rl0 and rl1 = rl0 + 8 see the same relative offset
    // at columns 8 apart, so the table entry loaded for row rl0 in n-tile nt serves row rl1 in nt + 1.
    float mx0 = -INFINITY, mx1 = -INFINITY;
    {
      const float4* ew = &sm.e[buf][2 * qd - rl0 + (BQ - 1)];
      float4 ep0 = ew[-8], ep1 = ew[-7];
      if (sflag[kt] == 0) {
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          const float4 ec0 = ew[nt * 8], ec1 = ew[nt * 8 + 1];
          const float a0 = s[nt][0] + pq0[0] * ec0.x + pq0[1] * ec0.y + pq0[2] * ec0.z + pq0[3] * ec0.w;
          const float a1 = s[nt][1] + pq0[0] * ec1.x + pq0[1] * ec1.y + pq0[2] * ec1.z + pq0[3] * ec1.w;
          const float b0 = s[nt][2] + pq1[0] * ep0.x + pq1[1] * ep0.y + pq1[2] * ep0.z + pq1[3] * ep0.w;
          const float b1 = s[nt][3] + pq1[0] * ep1.x + pq1[1] * ep1.y + pq1[2] * ep1.z + pq1[3] * ep1.w;
          s[nt][0] = a0, s[nt][1] = a1, s[nt][2] = b0, s[nt][3] = b1;
          mx0 = fmaxf(mx0, fmaxf(a0, a1)), mx1 = fmaxf(mx1, fmaxf(b0, b1));
          ep0 = ec0, ep1 = ec1;
        }
      } else {
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          const float4 ec0 = ew[nt * 8], ec1 = ew[nt * 8 + 1];
          float a0 = s[nt][0] + pq0[0] * ec0.x + pq0[1] * ec0.y + pq0[2] * ec0.z + pq0[3] * ec0.w;
          float a1 = s[nt][1] + pq0[0] * ec1.x + pq0[1] * ec1.y + pq0[2] * ec1.z + pq0[3] * ec1.w;
          float b0 = s[nt][2] + pq1[0] * ep0.x + pq1[1] * ep0.y + pq1[2] * ep0.z + pq1[3] * ep0.w;
          float b1 = s[nt][3] + pq1[0] * ep1.x + pq1[1] * ep1.y + pq1[2] * ep1.z + pq1[3] * ep1.w;
          const uint8_t k0 = mrow[nt * 8 + 2 * qd], k1 = mrow[nt * 8 + 2 * qd + 1];
          if (k0) a0 = b0 = (k0 == 1 ? ATTN_MASKED : -INFINITY);
          if (k1) a1 = b1 = (k1 == 1 ? ATTN_MASKED : -INFINITY);
          s[nt][0] = a0, s[nt][1] = a1, s[nt][2] = b0, s[nt][3] = b1;
          mx0 = fmaxf(mx0, fmaxf(a0, a1)), mx1 = fmaxf(mx1, fmaxf(b0, b1));
          ep0 = ec0, ep1 = ec1;
        }
      }
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float mn0 = fmaxf(m0, mx0), mn1 = fmaxf(m1, mx1);
    const bool grow = (mn0 != m0) || (mn1 != m1);
    const float ms0 = mn0, ms1 = mn1;  // scores are in log2 units
    float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const float pa = ex2_approx(s[nt][e] - ms0), pb = ex2_approx(s[nt][2 + e] - ms1);
        s[nt][e] = pa, s[nt][2 + e] = pb;
        rs0 += pa, rs1 += pb;
      }
    }
    if (__any_sync(0xffffffffu, grow)) {  // warp-uniform: rescale only when some row's running max moved
      const float sc0 = ex2_approx(m0 - mn0), sc1 = ex2_approx(m1 - mn1);
      l0 *= sc0, l1 *= sc1;
#pragma unroll
      for (int n = 0; n < DV / 8; ++n) o[n][0] *= sc0, o[n][1] *= sc0, o[n][2] *= sc1, o[n][3] *= sc1;
      m0 = mn0, m1 = mn1;
    }
    l0 += rs0, l1 += rs1;
    // ---- O += P V
#pragma unroll
    for (int kk = 0; kk < BKV / 16; ++kk) {
      uint32_t pa[4];
      pa[0] = pack_bf16x2(s[2 * kk][0], s[2 * kk][1]);
      pa[1] = pack_bf16x2(s[2 * kk][2], s[2 * kk][3]);
      pa[2] = pack_bf16x2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
      pa[3] = pack_bf16x2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
      for (int n2 = 0; n2 < DV / 16; ++n2) {
        uint32_t vb[4];
        ldsm_x4_t(vb, &sm.v[buf][kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8][n2 * 16 + (lane >> 4) * 8]);
        mma_bf16(o[2 * n2], pa, vb[0], vb[1]);
        mma_bf16(o[2 * n2 + 1], pa, vb[2], vb[3]);
      }
    }
  }
  // ---- normalise, gate, store
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1), l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1), l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float inv0 = 1.f / l0, inv1 = 1.f / l1;
  const int ia = i0 + rl0, ib = i0 + rl1;
#pragma unroll
  for (int n = 0; n < DV / 8; ++n) {
    const int col = n * 8 + 2 * qd;
    if (col >= dv_valid) continue;
    float a0 = o[n][0] * inv0, a1 = o[n][1] * inv0, b0 = o[n][2] * inv1, b1 = o[n][3] * inv1;
    if (ia < L) {
      const size_t row = (size_t)b * L + ia;
      if (p.gate) {
        const uint32_t gv = *reinterpret_cast<const uint32_t*>(p.gate + row * p.ldg + vcol0 + col);
        a0 *= bf16_lo(gv), a1 *= bf16_hi(gv);
      }
      *reinterpret_cast<uint32_t*>(p.out + row * p.ldo + vcol0 + col) = pack_bf16x2(a0, a1);
    }
    if (ib < L) {
      const size_t row = (size_t)b * L + ib;
      if (p.gate) {
        const uint32_t gv = *reinterpret_cast<const uint32_t*>(p.gate + row * p.ldg + vcol0 + col);
        b0 *= bf16_lo(gv), b1 *= bf16_hi(gv);
      }
      *reinterpret_cast<uint32_t*>(p.out + row * p.ldo + vcol0 + col) = pack_bf16x2(b0, b1);
    }
  }
}

template <int DV>
static int launch_attn(const AttnParams& p, int chunks, int dv_valid, cudaStream_t s) {
  static unsigned long long attr = 0;
  const int n_tiles = (p.L + BKV - 1) / BKV;
  const int smem = (int)sizeof(AttnSmem<DV>) + n_tiles * BKV + ((n_tiles + 15) & ~15);
  if (smem > 227 * 1024) return 931;  // sequence too long for the in-kernel mask copy
  if (first_use_on_device(attr)) {
    cudaFuncSetAttribute(attn_kernel<DV>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  }
  dim3 grid((p.L + BQ - 1) / BQ, p.heads * chunks, p.nb);
  attn_kernel<DV><<<grid, 256, smem, s>>>(p, chunks, dv_valid);
  return (int)cudaGetLastError();
}

int attention_na_tc(const AttnParams& p, cudaStream_t s);
int attention_sa_tma(const AttnParams& p, cudaStream_t s);
int g_attn_impl = 0;  // 0: TMA-fed SelfAttention (zv_attn_sa.cu) + tcgen05 NonlinAttention (zv_attn_tc.cu);
                      // 1: the cp.async / mma.sync flash kernel below for everything
                      // 2: as 0, with the exact-maxima repeat of the NonlinAttention kernel forced (tests)

int attention(const AttnParams& p, cudaStream_t s) {
  if (p.heads == 1 && g_attn_impl != 1) {
    AttnParams q = p;
    if (g_attn_impl == 2) q.dbg |= 4;
    const int rc = attention_na_tc(q, s);
    if (rc >= 0) return rc;
  }
  if (p.heads == p.H && p.dv == 12 && g_attn_impl != 1) {
    const int rc = attention_sa_tma(p, s);
    if (rc >= 0) return rc;
  }
  if (p.dv == 12) return launch_attn<16>(p, 1, 12, s);
  if (p.dv % 192 == 0) return launch_attn<192>(p, p.dv / 192, 192, s);
  if (p.dv % 128 == 0) return launch_attn<128>(p, p.dv / 128, 128, s);
  if (p.dv % 48 == 0) return launch_attn<48>(p, p.dv / 48, 48, s);
  return 930;
}

}  // namespace zv
