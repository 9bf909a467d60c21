// Bandwidth-bound kernels of the Zipformer layer: coalesced, 128-bit vectorised where the
// layout allows, shared-memory staged where there is reuse (depthwise conv), warp-shuffle
// reductions (BiasNorm).  Grids are sized from the data; all launches are asynchronous.
#include <cuda.h>

#include "zv_common.cuh"
#include "zv_kernels.h"

namespace zv {

static inline unsigned blocks_for(size_t n, int threads) { return (unsigned)((n + threads - 1) / threads); }

// ------------------------------------------------------------------ input packing (+ CFG doubling)
// One CTA per output row (no per-element index division); a thread packs one pair of columns.  Fx and Ft are even
// (checked by the launcher), so a pair never straddles two segments and is fetched with one 8-byte load.
__global__ void pack_input_kernel(const float* __restrict__ x, const float* __restrict__ text,
                                  const float* __restrict__ speech, const uint8_t* __restrict__ mask, int B, int T,
                                  int Fx, int Ft, int cfg, int zero_speech_uncond, bf16* __restrict__ out, int ldo,
                                  uint8_t* __restrict__ mask_out) {
  const int row = blockIdx.x;  // bo * T + t
  const int bo = row / T, t = row - bo * T;
  const int b = (cfg && bo >= B) ? bo - B : bo;
  const bool uncond = cfg && bo < B;
  // row = [x (Fx) | text (Ft) | speech (Fx)]: Fx = Ft for the mono models, Fx = 2 Ft for the two-channel dialog model
  const size_t rowi = (size_t)b * T + t;
  for (int c = threadIdx.x * 2; c < ldo; c += blockDim.x * 2) {
    float2 v = make_float2(0.f, 0.f);
    if (c < Fx) v = *reinterpret_cast<const float2*>(x + rowi * Fx + c);
    else if (c < Fx + Ft) {
      if (!uncond) v = *reinterpret_cast<const float2*>(text + rowi * Ft + c - Fx);
    } else if (c < 2 * Fx + Ft) {
      if (!(uncond && zero_speech_uncond)) v = *reinterpret_cast<const float2*>(speech + rowi * Fx + c - Fx - Ft);
    }
    *reinterpret_cast<uint32_t*>(out + (size_t)row * ldo + c) = pack_bf16x2(v.x, v.y);
  }
  if (threadIdx.x == 0 && mask_out) mask_out[row] = mask ? mask[rowi] : 0;
}

int pack_input(const float* x, const float* text, const float* speech, const uint8_t* mask, int B, int T, int Fx,
               int Ft, int cfg, int zero_speech_uncond, bf16* out, int ldo, uint8_t* mask_out, cudaStream_t s) {
  if ((Fx & 1) || (Ft & 1) || (ldo & 1)) return 913;
  const int rows = (cfg ? 2 * B : B) * T;
  const int threads = ((ldo / 2 + 31) / 32) * 32;
  pack_input_kernel<<<rows, threads < 256 ? threads : 256, 0, s>>>(x, text, speech, mask, B, T, Fx, Ft, cfg,
                                                                    zero_speech_uncond, out, ldo, mask_out);
  return (int)cudaGetLastError();
}

// ------------------------------------------------------------------ CFG combine + Euler update
__global__ void cfg_euler_kernel(float* __restrict__ x, const float* __restrict__ v,
                                 const float* __restrict__ guidance, int guidance_len, float g_mult, int B,
                                 size_t per_b, int cfg, float dt, float* __restrict__ v_out) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= (size_t)B * per_b) return;
  float vel;
  if (cfg) {
    const int b = (int)(gid / per_b);
    const float g = guidance[guidance_len == 1 ? 0 : b] * g_mult;
    const float u = v[gid], c = v[gid + (size_t)B * per_b];
    vel = (1.f + g) * c - g * u;
  } else {
    vel = v[gid];
  }
  if (v_out) v_out[gid] = vel;
  if (x) x[gid] = x[gid] + vel * dt;
}

int cfg_euler_update(float* x, const float* v, const float* guidance, int guidance_len, float g_mult, int B, int T,
                     int F, int cfg, float dt, float* v_out, cudaStream_t s) {
  const size_t per_b = (size_t)T * F;
  cfg_euler_kernel<<<blocks_for((size_t)B * per_b, 256), 256, 0, s>>>(x, v, guidance, guidance_len, g_mult, B, per_b,
                                                                      cfg, dt, v_out);
  return (int)cudaGetLastError();
}

__global__ void f32_to_bf16_kernel(const float* __restrict__ in, bf16* __restrict__ out, size_t n) {
  const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i + 3 < n) {
    const float4 v = *reinterpret_cast<const float4*>(in + i);
    *reinterpret_cast<uint2*>(out + i) = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
  } else {
    for (size_t j = i; j < n; ++j) out[j] = __float2bfloat16(in[j]);
  }
}
int f32_to_bf16(const float* in, bf16* out, size_t n, cudaStream_t s) {
  f32_to_bf16_kernel<<<blocks_for((n + 3) / 4, 256), 256, 0, s>>>(in, out, n);
  return (int)cudaGetLastError();
}

// ------------------------------------------------------------------ depthwise conv + SwooshR
// Persistent CTAs (two per SM) walk (sequence, 256-channel chunk, TT-frame tile) units.  The input tile plus its
// (K-1)-frame halo arrives by ONE 3-D TMA box [1][TT+K-1][256 channels] per unit -- frames before / after the
// sequence and channels beyond C are zero-filled by the tensor map, which is exactly the conv's "same" padding
// -- into a double-buffered shared-memory tile, so the load of unit i+1 overlaps the FMAs of unit i.  Taps and
// accumulators are packed fp32 pairs (channels c, c+1): one FFMA2 per tap and frame; every thread produces
// FG = 8 consecutive frames x 2 channels per step so each shared-memory word is read once per 8 outputs.
template <int K, int TT, int FG>
__global__ void __launch_bounds__(256, 2)
dwconv_kernel(const __grid_constant__ CUtensorMap tm_in, const float* __restrict__ w, const float* __restrict__ bias,
              int nb, int L, int C, bf16* __restrict__ out, int ldo) {
  constexpr int PAD = K / 2, ROWS = TT + K - 1;
  constexpr uint32_t TILE_BYTES = ROWS * 128 * 4;
  extern __shared__ __align__(128) uint8_t dw_smem[];
  uint8_t* base = dw_smem + ((128u - (smem_u32(dw_smem) & 127u)) & 127u);
  uint64_t* full = reinterpret_cast<uint64_t*>(base + 2 * TILE_BYTES);
  const int tid = threadIdx.y * 128 + threadIdx.x;
  const int cp = threadIdx.x;  // channel pair within the chunk
  const int n_t = (L + TT - 1) / TT, n_c = (C + 255) / 256;
  const int units = nb * n_c * n_t;
  if (tid == 0) {
    mbar_init(&full[0], 1), mbar_init(&full[1], 1);
    fence_barrier_init();
  }
  __syncthreads();
  auto issue = [&](int u, int buf) {  // thread 0 only
    const int tt = u % n_t, b = (u / n_t) % nb, cb = u / (n_t * nb);  // chunk index slowest: taps reload once
    mbar_arrive_expect_tx(&full[buf], TILE_BYTES);
    tma_load_3d(base + buf * TILE_BYTES, &tm_in, &full[buf], cb * 256, tt * TT - PAD, b);
  };
  if (tid == 0 && (int)blockIdx.x < units) issue(blockIdx.x, 0);
  int it = 0, cb_prev = -1;
  f32x2 wp[K];
  f32x2 bp = f2_pack(0.f, 0.f);
  for (int u = blockIdx.x; u < units; u += gridDim.x, ++it) {
    const int buf = it & 1;
    const int tt0 = (u % n_t) * TT, b = (u / n_t) % nb, cb = u / (n_t * nb);
    // every thread finished reading buffer buf^1 at the barrier that closed the previous unit
    if (tid == 0 && u + (int)gridDim.x < units) issue(u + gridDim.x, buf ^ 1);
    const int c = cb * 256 + cp * 2;
    const bool active = c < C;
    if (cb != cb_prev && active) {  // taps of this thread's channel pair
#pragma unroll
      for (int j = 0; j < K; ++j) wp[j] = f2_pack(__ldg(w + (size_t)c * K + j), __ldg(w + (size_t)(c + 1) * K + j));
      bp = f2_pack(__ldg(bias + c), __ldg(bias + c + 1));
      cb_prev = cb;
    }
    mbar_wait(&full[buf], (it >> 1) & 1);
    if (active) {
      const uint32_t(*tile)[128] = reinterpret_cast<const uint32_t(*)[128]>(base + buf * TILE_BYTES);
      bf16* dst = out + (size_t)b * L * ldo;
      constexpr int per_y = TT / 2;
      for (int tt = threadIdx.y * per_y; tt < (threadIdx.y + 1) * per_y; tt += FG) {
        if (tt0 + tt >= L) break;
        f32x2 acc[FG];
#pragma unroll
        for (int o = 0; o < FG; ++o) acc[o] = bp;
#pragma unroll
        for (int j = 0; j < K + FG - 1; ++j) {
          const uint32_t xv = tile[tt + j][cp];
          const f32x2 xp = f2_pack(bf16_lo(xv), bf16_hi(xv));
#pragma unroll
          for (int o = 0; o < FG; ++o) {
            const int tap = j - o;
            if (tap >= 0 && tap < K) acc[o] = f2_fma(wp[tap], xp, acc[o]);
          }
        }
        bf16* op = dst + (size_t)(tt0 + tt) * ldo + c;
        if (tt0 + tt + FG <= L) {  // whole group inside the sequence: no per-frame predicate
#pragma unroll
          for (int o = 0; o < FG; ++o, op += ldo) {
            float a0, a1;
            f2_unpack(swoosh_r_fast2(acc[o]), a0, a1);
            *reinterpret_cast<uint32_t*>(op) = pack_bf16x2(a0, a1);
          }
        } else {
#pragma unroll
          for (int o = 0; o < FG; ++o, op += ldo) {
            float a0, a1;
            f2_unpack(swoosh_r_fast2(acc[o]), a0, a1);
            if (tt0 + tt + o < L) *reinterpret_cast<uint32_t*>(op) = pack_bf16x2(a0, a1);
          }
        }
      }
    }
    __syncthreads();
  }
}

int make_tmap3d_16(void* tm, const void* ptr, uint64_t n0, uint64_t n1, uint64_t n2, uint64_t ld1, uint64_t ld2,
                   uint32_t box0, uint32_t box1);  // zv_gemm.cu

template <int K, int TT, int FG>
static int launch_dwconv(const bf16* in, const float* w, const float* bias, int nb, int L, int C, bf16* out, int ldo,
                         cudaStream_t s) {
  constexpr int smem = 2 * (TT + K - 1) * 128 * 4 + 128 + 64;
  static unsigned long long attr = 0;
  static int sms = 0;
  if (first_use_on_device(attr)) {
    cudaFuncSetAttribute(dwconv_kernel<K, TT, FG>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  CUtensorMap tm;
  int rc = make_tmap3d_16(&tm, in, (uint64_t)C, (uint64_t)L, (uint64_t)nb, (uint64_t)C, (uint64_t)L * C, 256,
                          TT + K - 1);
  if (rc) return rc;
  const int units = nb * ((C + 255) / 256) * ((L + TT - 1) / TT);
  const int grid = units < 2 * sms ? units : 2 * sms;
  dwconv_kernel<K, TT, FG><<<grid, dim3(128, 2), smem, s>>>(tm, w, bias, nb, L, C, out, ldo);
  return (int)cudaGetLastError();
}

int dwconv_swooshr(const bf16* in, const float* w, const float* bias, int nb, int L, int C, int k, bf16* out, int ldo,
                   cudaStream_t s) {
  if (ldo < C || (ldo & 1) || (C & 7)) return 911;
  static const int fg = [] {  // frames per thread step: 16 halves the shared-memory loads per output (ZV_DWCONV_FG=8|16)
    const char* e = getenv("ZV_DWCONV_FG");
    return e ? atoi(e) : 16;
  }();
  switch (k) {
    case 7: return launch_dwconv<7, 64, 8>(in, w, bias, nb, L, C, out, ldo, s);
    case 9: return launch_dwconv<9, 64, 8>(in, w, bias, nb, L, C, out, ldo, s);
    case 15:
      return fg == 16 ? launch_dwconv<15, 64, 16>(in, w, bias, nb, L, C, out, ldo, s)
                      : launch_dwconv<15, 64, 8>(in, w, bias, nb, L, C, out, ldo, s);
    case 31:
      return fg == 16 ? launch_dwconv<31, 64, 16>(in, w, bias, nb, L, C, out, ldo, s)
                      : launch_dwconv<31, 64, 8>(in, w, bias, nb, L, C, out, ldo, s);
    default: return 910;
  }
}

// ------------------------------------------------------------------ BiasNorm + bypass (one warp per row)
// All three kernels below read and write the residual stream in its s24 storage format (zv_common.cuh): a bf16 `hi`
// plane and a byte `lo` plane per tensor, 8 elements (16 + 8 bytes) per thread and access.
struct S24Ld {
  const bf16* hi;
  const uint8_t* lo;
  ZV_DEVINL void load8(size_t off, float (&v)[8]) const {
    s24_dec8(*reinterpret_cast<const uint4*>(hi + off), *reinterpret_cast<const uint2*>(lo + off), v);
  }
};
ZV_DEVINL void s24_store8(bf16* hi, uint8_t* lo, size_t off, const float (&v)[8]) {
  uint4 h;
  uint2 l;
  s24_enc8(v, h, l);
  *reinterpret_cast<uint4*>(hi + off) = h;
  *reinterpret_cast<uint2*>(lo + off) = l;
}

__global__ void biasnorm_bypass_kernel(S24Ld src, S24Ld orig, const float* __restrict__ nbias, float scale_exp,
                                       const float* __restrict__ bscale, int M, int D, bf16* __restrict__ out_hi,
                                       uint8_t* __restrict__ out_lo) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= M) return;
  const size_t base = (size_t)row * D;
  float v[2][8];  // up to D = 512: two groups of 8 per lane
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int col = (i * 32 + lane) * 8;
    if (col < D) {
      src.load8(base + col, v[i]);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const float4 nb = *reinterpret_cast<const float4*>(nbias + col + 4 * h);
        const float d0 = v[i][4 * h] - nb.x, d1 = v[i][4 * h + 1] - nb.y, d2 = v[i][4 * h + 2] - nb.z,
                    d3 = v[i][4 * h + 3] - nb.w;
        ss += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
      }
    }
  }
  ss = warp_sum(ss);
  const float sc = rsqrtf(ss / (float)D) * scale_exp;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int col = (i * 32 + lane) * 8;
    if (col < D) {
      float og[8], r[8];
      orig.load8(base + col, og);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const float4 bs = *reinterpret_cast<const float4*>(bscale + col + 4 * h);
        r[4 * h] = og[4 * h] + (v[i][4 * h] * sc - og[4 * h]) * bs.x;
        r[4 * h + 1] = og[4 * h + 1] + (v[i][4 * h + 1] * sc - og[4 * h + 1]) * bs.y;
        r[4 * h + 2] = og[4 * h + 2] + (v[i][4 * h + 2] * sc - og[4 * h + 2]) * bs.z;
        r[4 * h + 3] = og[4 * h + 3] + (v[i][4 * h + 3] * sc - og[4 * h + 3]) * bs.w;
      }
      s24_store8(out_hi, out_lo, base + col, r);
    }
  }
}

int biasnorm_bypass(const bf16* src_hi, const uint8_t* src_lo, const bf16* orig_hi, const uint8_t* orig_lo,
                    const float* nbias, float log_scale, const float* bscale, int M, int D, bf16* out_hi,
                    uint8_t* out_lo, cudaStream_t s) {
  if (D > 512 || (D & 7)) return 911;
  const int warps = 8;
  biasnorm_bypass_kernel<<<(M + warps - 1) / warps, warps * 32, 0, s>>>(
      S24Ld{src_hi, src_lo}, S24Ld{orig_hi, orig_lo}, nbias, expf(log_scale), bscale, M, D, out_hi, out_lo);
  return (int)cudaGetLastError();
}

// ------------------------------------------------------------------ SimpleDownsample / SimpleUpsample
__global__ void downsample_kernel(S24Ld src, float4 w, int L, int L2, int D, int ds, bf16* __restrict__ out_hi,
                                  uint8_t* __restrict__ out_lo, size_t total8) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total8) return;
  const int d8 = D >> 3;
  const int c = (int)(gid % d8) * 8;
  const size_t row = gid / d8;  // b * L2 + j
  const int j = (int)(row % L2);
  const int b = (int)(row / L2);
  const float ws[4] = {w.x, w.y, w.z, w.w};
  float acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = 0.f;
  for (int k = 0; k < ds; ++k) {
    const int t = min(j * ds + k, L - 1);
    float v[8];
    src.load8(((size_t)b * L + t) * D + c, v);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] += v[i] * ws[k];
  }
  s24_store8(out_hi, out_lo, row * D + c, acc);
}
__global__ void mask_subsample_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int nb, int L, int L2,
                                      int ds) {
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= nb * L2) return;
  const int b = gid / L2, j = gid % L2;
  out[gid] = in[(size_t)b * L + j * ds];  // padding_mask[..., ::ds] (zipformer.py:857-858)
}

int downsample(const bf16* src_hi, const uint8_t* src_lo, const float* w, int nb, int L, int D, int ds, bf16* out_hi,
               uint8_t* out_lo, uint8_t* mask_out, const uint8_t* mask_in, cudaStream_t s) {
  if (ds > 4 || (D & 7)) return 912;
  const int L2 = (L + ds - 1) / ds;
  float4 wv = make_float4(w[0], ds > 1 ? w[1] : 0.f, ds > 2 ? w[2] : 0.f, ds > 3 ? w[3] : 0.f);
  const size_t total8 = (size_t)nb * L2 * (D >> 3);
  downsample_kernel<<<blocks_for(total8, 256), 256, 0, s>>>(S24Ld{src_hi, src_lo}, wv, L, L2, D, ds, out_hi, out_lo,
                                                            total8);
  if (mask_out) mask_subsample_kernel<<<blocks_for((size_t)nb * L2, 256), 256, 0, s>>>(mask_in, mask_out, nb, L, L2, ds);
  return (int)cudaGetLastError();
}

__global__ void upsample_combine_kernel(bf16* __restrict__ orig_hi, uint8_t* __restrict__ orig_lo, S24Ld y,
                                        const float* __restrict__ bscale, int L, int L2, int D, int ds, size_t total8) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total8) return;
  const int d8 = D >> 3;
  const int c = (int)(gid % d8) * 8;
  const size_t row = gid / d8;  // b * L + t
  const int t = (int)(row % L);
  const int b = (int)(row / L);
  float o[8], v[8], r[8];
  S24Ld{orig_hi, orig_lo}.load8(row * D + c, o);
  y.load8(((size_t)b * L2 + t / ds) * D + c, v);
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const float4 sc = *reinterpret_cast<const float4*>(bscale + c + 4 * h);
    r[4 * h] = o[4 * h] + (v[4 * h] - o[4 * h]) * sc.x;
    r[4 * h + 1] = o[4 * h + 1] + (v[4 * h + 1] - o[4 * h + 1]) * sc.y;
    r[4 * h + 2] = o[4 * h + 2] + (v[4 * h + 2] - o[4 * h + 2]) * sc.z;
    r[4 * h + 3] = o[4 * h + 3] + (v[4 * h + 3] - o[4 * h + 3]) * sc.w;
  }
  s24_store8(orig_hi, orig_lo, row * D + c, r);
}

int upsample_combine(bf16* orig_hi, uint8_t* orig_lo, const bf16* y_hi, const uint8_t* y_lo, const float* bscale,
                     int nb, int L, int L2, int D, int ds, cudaStream_t s) {
  if (D & 7) return 912;
  const size_t total8 = (size_t)nb * L * (D >> 3);
  upsample_combine_kernel<<<blocks_for(total8, 256), 256, 0, s>>>(orig_hi, orig_lo, S24Ld{y_hi, y_lo}, bscale, L, L2, D,
                                                                  ds, total8);
  return (int)cudaGetLastError();
}

}  // namespace zv
