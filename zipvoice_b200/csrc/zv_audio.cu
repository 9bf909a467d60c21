// VocosFbank mel front-end (zipvoice/utils/feature.py:36-120) and the Vocos vocoder head
// (third-party `vocos`, call site zipvoice/bin/infer_zipvoice.py:249-260,378) as CUDA kernels.
//   mel      : frame (reflect pad 512) x periodic Hann -> real FFT-1024 (packed 512-point radix-8 Stockham
//              in shared memory, zv_fft.cuh) -> |.| -> sparse HTK mel filterbank (<= 2 filters per bin)
//              -> log(max(., 1e-7)).  One launch; 5.56 algorithmic bytes per input sample.
//   vocoder  : embed Conv1d(100->512, k7) as an implicit-im2col tcgen05 GEMM (the TMA tensor map of
//              the A operand has a 104-element row pitch and 728 columns, so consecutive "rows" are
//              overlapping 7-frame windows of the zero-padded channel-last mel), LayerNorm kernels,
//              fused depthwise-conv7 + LayerNorm, pwconv GEMMs with GELU / residual epilogues
//              (layer scale gamma folded into pwconv2), head GEMM, and the iSTFT: spectrum ->
//              inverse real FFT-1024 -> Hann -> overlap-add / window-envelope normalisation.
// Activations of sequence b live at rows [b * Tp, b * Tp + T) with Tp = T + 6 (the 6 slack rows per
// sequence come from the implicit im2col; every kernel treats frames >= len_b as zero input).
#include <math.h>

#include "zv_common.cuh"
#include "zv_engine.h"
#include "zv_fft.cuh"

namespace zv {

static constexpr int N_FFT = 1024, HOP = 256, N_BINS = 513, N_MELS = 100;
static constexpr int VC_DIM = 512, VC_INTER = 1536, VC_LAYERS = 8, VC_MELP = 104, VC_K = 7;

struct Audio {
  // constant tables
  float2* tw512 = nullptr;    // exp(-2 pi i k / 512)
  float2* tw1024 = nullptr;   // exp(-2 pi i k / 1024)
  int* fb_start = nullptr;    // first FFT bin of mel filter m
  int* fb_count = nullptr;
  int* fb_off = nullptr;      // offset of its weights in fb_w
  float* fb_w = nullptr;
  // the same four arrays for the BigVGAN features: librosa Slaney mel basis with area normalisation
  int *sl_start = nullptr, *sl_count = nullptr, *sl_off = nullptr;
  float* sl_w = nullptr;
  // vocoder weights
  bool has_vocoder = false;
  Lin embed, pw1[VC_LAYERS], pw2[VC_LAYERS], head;
  float *norm_w = nullptr, *norm_b = nullptr, *fnorm_w = nullptr, *fnorm_b = nullptr;
  float *dw_w[VC_LAYERS], *dw_b[VC_LAYERS], *ln_w[VC_LAYERS], *ln_b[VC_LAYERS];
  // vocoder workspace
  size_t cap_rows = 0;
  int cap_B = 0;
  __nv_bfloat16 *melp = nullptr, *h = nullptr, *g = nullptr;
  float *x = nullptr, *emb = nullptr, *spec = nullptr, *frames = nullptr;
  int* lens_dev = nullptr;
  std::vector<void*> ws_allocs;
  // resampler filter banks by reduced (orig, new) ratio
  struct ResampleBank {
    int orig = 0, nw = 0, width = 0, K = 0;
    float* w = nullptr;  // [new][K]
  };
  std::vector<ResampleBank> banks;
};

static int audio_tables(Engine& e, Audio& a) {
  if (a.tw512) return 0;
  std::vector<float2> t512(512), t1024(1024);
  for (int i = 0; i < 512; ++i) t512[i] = make_float2((float)cos(-2.0 * M_PI * i / 512), (float)sin(-2.0 * M_PI * i / 512));
  for (int i = 0; i < 1024; ++i)
    t1024[i] = make_float2((float)cos(-2.0 * M_PI * i / 1024), (float)sin(-2.0 * M_PI * i / 1024));
  // torchaudio.functional.melscale_fbanks(513, 0, 12000, 100, 24000, norm=None, mel_scale="htk") in double
  const double f_max = 12000.0;
  auto hz2mel = [](double f) { return 2595.0 * log10(1.0 + f / 700.0); };
  std::vector<double> fpts(N_MELS + 2);
  const double mmin = hz2mel(0.0), mmax = hz2mel(f_max);
  for (int i = 0; i < N_MELS + 2; ++i) {
    const double m = mmin + (mmax - mmin) * i / (N_MELS + 1);
    fpts[i] = 700.0 * (pow(10.0, m / 2595.0) - 1.0);
  }
  std::vector<int> start(N_MELS), count(N_MELS), off(N_MELS);
  std::vector<float> w;
  for (int m = 0; m < N_MELS; ++m) {
    int first = -1, last = -2;
    std::vector<float> row(N_BINS, 0.f);
    for (int k = 0; k < N_BINS; ++k) {
      const double f = f_max * k / (N_BINS - 1);
      const double down = (f - fpts[m]) / (fpts[m + 1] - fpts[m]);
      const double up = (fpts[m + 2] - f) / (fpts[m + 2] - fpts[m + 1]);
      const double v = fmin(down, up);
      if (v > 0.0) {
        row[k] = (float)v;
        if (first < 0) first = k;
        last = k;
      }
    }
    start[m] = first < 0 ? 0 : first, count[m] = first < 0 ? 0 : last - first + 1, off[m] = (int)w.size();
    for (int k = 0; k < count[m]; ++k) w.push_back(row[start[m] + k]);
  }
  // librosa.filters.mel(sr=24000, n_fft=1024, n_mels=100, fmin=0, fmax=None): Slaney scale (200/3 Hz per mel below
  // 1 kHz, logarithmic above with step ln(6.4)/27), triangles on the rfft bin frequencies, each filter scaled by
  // 2 / (f[m+2] - f[m]); double arithmetic, fp32 weights (zipvoice/utils/_bigvgan_mel_feature.py:76-79)
  std::vector<int> sl_start(N_MELS), sl_count(N_MELS), sl_off(N_MELS);
  std::vector<float> sl_w;
  {
    const double f_sp = 200.0 / 3.0, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp, logstep = log(6.4) / 27.0;
    auto hz2sl = [&](double f) { return f >= min_log_hz ? min_log_mel + log(f / min_log_hz) / logstep : f / f_sp; };
    auto sl2hz = [&](double m) { return m >= min_log_mel ? min_log_hz * exp(logstep * (m - min_log_mel)) : f_sp * m; };
    std::vector<double> mf(N_MELS + 2);
    const double lo = hz2sl(0.0), hi = hz2sl(f_max);
    for (int i = 0; i < N_MELS + 2; ++i) mf[i] = sl2hz(lo + (hi - lo) * i / (N_MELS + 1));
    for (int m = 0; m < N_MELS; ++m) {
      int first = -1, last = -2;
      std::vector<float> row(N_BINS, 0.f);
      const double enorm = 2.0 / (mf[m + 2] - mf[m]);
      for (int k = 0; k < N_BINS; ++k) {
        const double f = 24000.0 * k / N_FFT;
        const double lower = (f - mf[m]) / (mf[m + 1] - mf[m]);
        const double upper = (mf[m + 2] - f) / (mf[m + 2] - mf[m + 1]);
        const double v = fmin(lower, upper);
        if (v > 0.0) {
          row[k] = (float)(v * enorm);
          if (first < 0) first = k;
          last = k;
        }
      }
      sl_start[m] = first < 0 ? 0 : first, sl_count[m] = first < 0 ? 0 : last - first + 1, sl_off[m] = (int)sl_w.size();
      for (int k = 0; k < sl_count[m]; ++k) sl_w.push_back(row[sl_start[m] + k]);
    }
  }
  auto up = [&](const void* src, size_t bytes, void** dst) {
    int rc = dev_alloc(e, e.weight_allocs, dst, bytes);
    if (rc) return rc;
    return (int)cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice);
  };
  int rc = up(t512.data(), 512 * sizeof(float2), (void**)&a.tw512);
  if (!rc) rc = up(t1024.data(), 1024 * sizeof(float2), (void**)&a.tw1024);
  if (!rc) rc = up(start.data(), N_MELS * 4, (void**)&a.fb_start);
  if (!rc) rc = up(count.data(), N_MELS * 4, (void**)&a.fb_count);
  if (!rc) rc = up(off.data(), N_MELS * 4, (void**)&a.fb_off);
  if (!rc) rc = up(w.data(), w.size() * 4, (void**)&a.fb_w);
  if (!rc) rc = up(sl_start.data(), N_MELS * 4, (void**)&a.sl_start);
  if (!rc) rc = up(sl_count.data(), N_MELS * 4, (void**)&a.sl_count);
  if (!rc) rc = up(sl_off.data(), N_MELS * 4, (void**)&a.sl_off);
  if (!rc) rc = up(sl_w.data(), sl_w.size() * 4, (void**)&a.sl_w);
  return rc;
}

// ------------------------------------------------------------------ mel front-end
// CTA = 256 threads = 4 frame groups of 64 threads; each group transforms one frame per iteration.
// The two feature types of zipvoice/utils/feature.py share the kernel:
//   VocosFbank   (:36-120)  frames centred (reflect pad 512), N / 256 + 1 STFT frames, |X|, HTK basis, log clamp 1e-7
//   BigVGANFbank (:133-204) reflect pad (1024 - 256) / 2 = 384 and no centring, N / 256 frames,
//                           sqrt(|X|^2 + 1e-9), Slaney basis, log clamp 1e-5
struct FbankVariant {
  int pad;           // samples of reflect padding before frame 0
  int extra_frames;  // STFT frames = len / HOP + extra_frames
  float mag_eps, log_floor;
};
struct FftSmem {
  float2 tw512[512];
  float2 tw1024[1024];
  float2 buf[4][512];
  float mag[4][520];
};

__global__ void __launch_bounds__(256) fbank_kernel(const float* __restrict__ wav, const int64_t* __restrict__ lens,
                                                    int N, int T, const float2* __restrict__ tw512g,
                                                    const float2* __restrict__ tw1024g,
                                                    const int* __restrict__ fb_start, const int* __restrict__ fb_count,
                                                    const int* __restrict__ fb_off, const float* __restrict__ fb_w,
                                                    int frames_per_cta, FbankVariant var, float* __restrict__ mel) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  FftSmem& sm = *reinterpret_cast<FftSmem*>(smem_raw);
  const int tid = threadIdx.x, g = tid >> 6, j = tid & 63;
  const int b = blockIdx.y;
  for (int i = tid; i < 512; i += 256) sm.tw512[i] = tw512g[i];
  for (int i = tid; i < 1024; i += 256) sm.tw1024[i] = tw1024g[i];
  __syncthreads();
  const long long len = lens ? lens[b] : (long long)N;
  const int Tb = (int)((len + HOP / 2) / HOP);  // lhotse compute_num_frames (round half up)
  const int n_spec = max((int)(len / HOP) + var.extra_frames, 1);  // frames the reference STFT computes
  const float* x = wav + (size_t)b * N;
  const int t_begin = blockIdx.x * frames_per_cta;
  for (int it = 0; it < frames_per_cta; it += 4) {
    const int t = t_begin + it + g;
    // frames beyond torchaudio's count replicate the last one (feature.py:105-110)
    const int ts = min(t, n_spec - 1);
    const bool live = t < T && t < Tb;
    float2 v[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int n = j + 64 * r;
      float2 z = make_float2(0.f, 0.f);
      if (live) {
        long long s0 = (long long)ts * HOP - var.pad + 2 * n, s1 = s0 + 1;
        s0 = s0 < 0 ? -s0 : (s0 >= len ? 2 * (len - 1) - s0 : s0);
        s1 = s1 < 0 ? -s1 : (s1 >= len ? 2 * (len - 1) - s1 : s1);
        s0 = max(0ll, min(s0, len - 1)), s1 = max(0ll, min(s1, len - 1));
        const float w0 = 0.5f - 0.5f * sm.tw1024[2 * n].x, w1 = 0.5f - 0.5f * sm.tw1024[2 * n + 1].x;
        z = make_float2(x[s0] * w0, x[s1] * w1);
      }
      v[r] = z;
    }
    fft512_block<false>(v, j, sm.buf[g], sm.tw512);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int k = j + 64 * r;
      const float2 X = rfft_unpack(sm.buf[g], k, sm.tw1024[k]);
      sm.mag[g][k] = sqrtf(X.x * X.x + X.y * X.y + var.mag_eps);
    }
    if (j == 0) {
      const float2 X = rfft_unpack(sm.buf[g], 512, make_float2(-1.f, 0.f));
      sm.mag[g][512] = sqrtf(X.x * X.x + X.y * X.y + var.mag_eps);
    }
    __syncthreads();
    for (int m = j; m < N_MELS; m += 64) {
      const int st = fb_start[m], cnt = fb_count[m];
      const float* w = fb_w + fb_off[m];
      float acc = 0.f;
      for (int k = 0; k < cnt; ++k) acc = fmaf(sm.mag[g][st + k], __ldg(w + k), acc);
      if (t < T) mel[((size_t)b * T + t) * N_MELS + m] = live ? logf(fmaxf(acc, var.log_floor)) : 0.f;
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------ vocoder kernels
// mel (B, 100, T) fp32 channel-major -> zero-padded channel-last bf16 rows [(B * Tp + 6)][104]:
// row b*Tp + 3 + t holds frame t; frames >= len_b and the 3 + 3 border rows are zero (Conv1d padding).
__global__ void vc_pack_mel_kernel(const float* __restrict__ mel, const int* __restrict__ lens, int B, int T, int Tp,
                                   __nv_bfloat16* __restrict__ out, size_t total) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int c = (int)(gid % VC_MELP);
  const size_t row = gid / VC_MELP;
  const int b = (int)(row / Tp), tp = (int)(row % Tp);
  const int t = tp - 3;
  float v = 0.f;
  if (b < B && c < N_MELS && t >= 0 && t < T && t < lens[b]) v = mel[((size_t)b * N_MELS + c) * T + t];
  out[gid] = __float2bfloat16(v);
}

// LayerNorm over 512 channels, one warp per row (two-pass statistics in registers, eps 1e-6).
__global__ void vc_layernorm_kernel(const float* __restrict__ in, const float* __restrict__ w,
                                    const float* __restrict__ bias, int rows, float* __restrict__ out_f32,
                                    __nv_bfloat16* __restrict__ out_bf16) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float* x = in + (size_t)row * VC_DIM;
  float4 v[4];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[i] = *reinterpret_cast<const float4*>(x + (i * 32 + lane) * 4);
    s += v[i].x + v[i].y + v[i].z + v[i].w;
  }
  const float mean = warp_sum(s) * (1.0f / VC_DIM);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
    q += a * a + b * b + c * c + d * d;
  }
  const float rstd = rsqrtf(warp_sum(q) * (1.0f / VC_DIM) + 1e-6f);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int col = (i * 32 + lane) * 4;
    const float4 ww = *reinterpret_cast<const float4*>(w + col), bb = *reinterpret_cast<const float4*>(bias + col);
    float4 r;
    r.x = (v[i].x - mean) * rstd * ww.x + bb.x, r.y = (v[i].y - mean) * rstd * ww.y + bb.y;
    r.z = (v[i].z - mean) * rstd * ww.z + bb.z, r.w = (v[i].w - mean) * rstd * ww.w + bb.w;
    if (out_f32) *reinterpret_cast<float4*>(out_f32 + (size_t)row * VC_DIM + col) = r;
    if (out_bf16)
      *reinterpret_cast<uint2*>(out_bf16 + (size_t)row * VC_DIM + col) =
          make_uint2(pack_bf16x2(r.x, r.y), pack_bf16x2(r.z, r.w));
  }
}

// ConvNeXt block head: depthwise Conv1d(k=7, pad 3) over time + LayerNorm -> bf16 (GEMM A operand).
// HBM-bound (4 B in + 2 B out per element). CTA = 128 threads walking DW_SUB frames at a time along one
// sequence; a thread owns 4 channels: its 7 x 4 taps and a sliding window of 6 + DW_SUB input rows stay
// in registers, every row is read from global exactly once per CTA (plus a 6-row halo per run of
// frames_per_cta) as one coalesced 2 KB line, and the next DW_SUB rows are in flight while the current
// ones are computed. LayerNorm statistics (two passes, as the reference's F.layer_norm) go through a
// warp shuffle + 4-warp shared-memory reduction per group of frames.
static constexpr int DW_SUB = 4;
__global__ void __launch_bounds__(128, 4) vc_dwconv_ln_kernel(const float* __restrict__ in,
                                                              const int* __restrict__ lens, int T, int Tp,
                                                              int frames_per_cta, const float* __restrict__ dw_w,
                                                              const float* __restrict__ dw_b,
                                                              const float* __restrict__ ln_w,
                                                              const float* __restrict__ ln_b,
                                                              __nv_bfloat16* __restrict__ out) {
  __shared__ float red_s[4][DW_SUB], red_q[4][DW_SUB];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int b = blockIdx.y, t0 = blockIdx.x * frames_per_cta;
  const int t_end = min(t0 + frames_per_cta, T);
  const int len = min(lens[b], T);
  const int col = tid * 4;
  const float* src = in + (size_t)b * Tp * VC_DIM + col;
  auto load_row = [&](int t) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (t >= 0 && t < len) v = __ldg(reinterpret_cast<const float4*>(src + (size_t)t * VC_DIM));
    return v;
  };
  float4 w[VC_K];  // weight (512, 1, 7): channel-major
  {
    float wf[4 * VC_K];
#pragma unroll
    for (int i = 0; i < VC_K; ++i)
      *reinterpret_cast<float4*>(&wf[4 * i]) = __ldg(reinterpret_cast<const float4*>(dw_w + (size_t)col * VC_K) + i);
#pragma unroll
    for (int k = 0; k < VC_K; ++k) w[k] = make_float4(wf[k], wf[VC_K + k], wf[2 * VC_K + k], wf[3 * VC_K + k]);
  }
  const float4 cb = __ldg(reinterpret_cast<const float4*>(dw_b + col));
  const float4 lw = __ldg(reinterpret_cast<const float4*>(ln_w + col));
  const float4 lb = __ldg(reinterpret_cast<const float4*>(ln_b + col));
  float4 r[6 + DW_SUB], nxt[DW_SUB];  // r[j] = row tc - 3 + j
#pragma unroll
  for (int j = 0; j < 6; ++j) r[j] = load_row(t0 - 3 + j);
#pragma unroll
  for (int i = 0; i < DW_SUB; ++i) nxt[i] = load_row(t0 + 3 + i);
  for (int tc = t0; tc < t_end; tc += DW_SUB) {
#pragma unroll
    for (int i = 0; i < DW_SUB; ++i) r[6 + i] = nxt[i];
    if (tc + DW_SUB < t_end) {
#pragma unroll
      for (int i = 0; i < DW_SUB; ++i) nxt[i] = load_row(tc + DW_SUB + 3 + i);
    }
    float4 acc[DW_SUB];
    float s[DW_SUB];
#pragma unroll
    for (int f = 0; f < DW_SUB; ++f) {
      float4 a = cb;
#pragma unroll
      for (int k = 0; k < VC_K; ++k) {
        a.x = fmaf(r[f + k].x, w[k].x, a.x), a.y = fmaf(r[f + k].y, w[k].y, a.y);
        a.z = fmaf(r[f + k].z, w[k].z, a.z), a.w = fmaf(r[f + k].w, w[k].w, a.w);
      }
      acc[f] = a;
      s[f] = warp_sum((a.x + a.y) + (a.z + a.w));
    }
    if (lane == 0) {
#pragma unroll
      for (int f = 0; f < DW_SUB; ++f) red_s[warp][f] = s[f];
    }
    __syncthreads();
    float mean[DW_SUB], q[DW_SUB];
#pragma unroll
    for (int f = 0; f < DW_SUB; ++f) {
      mean[f] = ((red_s[0][f] + red_s[1][f]) + (red_s[2][f] + red_s[3][f])) * (1.0f / VC_DIM);
      const float a0 = acc[f].x - mean[f], a1 = acc[f].y - mean[f], a2 = acc[f].z - mean[f], a3 = acc[f].w - mean[f];
      q[f] = warp_sum((a0 * a0 + a1 * a1) + (a2 * a2 + a3 * a3));
    }
    if (lane == 0) {
#pragma unroll
      for (int f = 0; f < DW_SUB; ++f) red_q[warp][f] = q[f];
    }
    __syncthreads();
#pragma unroll
    for (int f = 0; f < DW_SUB; ++f) {
      const int t = tc + f;
      if (t < t_end) {
        const float var = ((red_q[0][f] + red_q[1][f]) + (red_q[2][f] + red_q[3][f])) * (1.0f / VC_DIM);
        const float rstd = rsqrtf(var + 1e-6f);
        const float r0 = (acc[f].x - mean[f]) * rstd * lw.x + lb.x, r1 = (acc[f].y - mean[f]) * rstd * lw.y + lb.y;
        const float r2 = (acc[f].z - mean[f]) * rstd * lw.z + lb.z, r3 = (acc[f].w - mean[f]) * rstd * lw.w + lb.w;
        *reinterpret_cast<uint2*>(out + ((size_t)b * Tp + t) * VC_DIM + col) =
            make_uint2(pack_bf16x2(r0, r1), pack_bf16x2(r2, r3));
      }
    }
#pragma unroll
    for (int j = 0; j < 6; ++j) r[j] = r[j + DW_SUB];
  }
}

// ISTFTHead, per frame: spec row = [513 log-magnitudes | 513 phases] (ld = ldspec) ->
// S = min(exp(m), 100) (cos p + i sin p) -> inverse real FFT-1024 -> x Hann -> frames[row][1024].
__global__ void __launch_bounds__(256) vc_istft_frames_kernel(const float* __restrict__ spec, int ldspec,
                                                              const int* __restrict__ lens, int T, int Tp,
                                                              const float2* __restrict__ tw512g,
                                                              const float2* __restrict__ tw1024g,
                                                              int frames_per_cta, float* __restrict__ frames) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  FftSmem& sm = *reinterpret_cast<FftSmem*>(smem_raw);
  const int tid = threadIdx.x, g = tid >> 6, j = tid & 63;
  const int b = blockIdx.y;
  for (int i = tid; i < 512; i += 256) sm.tw512[i] = tw512g[i];
  for (int i = tid; i < 1024; i += 256) sm.tw1024[i] = tw1024g[i];
  __syncthreads();
  const int len = min(lens[b], T);
  const int t_begin = blockIdx.x * frames_per_cta;
  for (int it = 0; it < frames_per_cta; it += 4) {
    const int t = t_begin + it + g;
    const bool live = t < len;
    const float* row = spec + ((size_t)b * Tp + min(t, T - 1)) * ldspec;
    __shared__ float2 nyq[4];  // bin 512 (bins 0..511 go to buf)
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int k = j + 64 * r;
      float2 s = make_float2(0.f, 0.f);
      if (live) {
        const float mag = fminf(expf(row[k]), 100.0f);
        float sn, cs;
        sincosf(row[N_BINS + k], &sn, &cs);
        s = make_float2(mag * cs, mag * sn);
      }
      sm.buf[g][k] = s;
    }
    if (j == 0) {
      float2 s = make_float2(0.f, 0.f);
      if (live) {
        const float mag = fminf(expf(row[512]), 100.0f);
        float sn, cs;
        sincosf(row[N_BINS + 512], &sn, &cs);
        s = make_float2(mag * cs, mag * sn);
      }
      nyq[g] = s;
    }
    __syncthreads();
    float2 v[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int k = j + 64 * r;
      float2 xk = sm.buf[g][k];
      float2 xn = k == 0 ? nyq[g] : sm.buf[g][512 - k];
      if (k == 0) xk.y = 0.f, xn.y = 0.f;  // irfft ignores the imaginary parts of DC and Nyquist
      v[r] = irfft_pack(xk, xn, sm.tw1024[k]);
    }
    __syncthreads();  // all reads of buf done before the transform overwrites it
    fft512_block<true>(v, j, sm.buf[g], sm.tw512);
    if (t < T) {
      float* o = frames + ((size_t)b * Tp + t) * N_FFT;
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const int n = j + 64 * r;
        const float2 z = sm.buf[g][n];
        const float w0 = 0.5f - 0.5f * sm.tw1024[2 * n].x, w1 = 0.5f - 0.5f * sm.tw1024[2 * n + 1].x;
        *reinterpret_cast<float2*>(o + 2 * n) =
            make_float2(z.x * (1.0f / N_FFT) * w0, z.y * (1.0f / N_FFT) * w1);
      }
    }
    __syncthreads();
  }
}

// Overlap-add with window-envelope normalisation (torch.istft, center=True): output sample n of
// sequence b sits at padded coordinate c = n + 512 and sums frames t with 0 <= c - 256 t < 1024.
__global__ void vc_ola_kernel(const float* __restrict__ frames, const int* __restrict__ lens, int T, int Tp,
                              const float2* __restrict__ tw1024g, int n_out, float* __restrict__ wav) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  const int b = blockIdx.y;
  if (n >= n_out) return;
  const int len = min(lens[b], T);
  float out = 0.f;
  if (n < HOP * (len - 1)) {
    const int c = n + N_FFT / 2;
    const int t_hi = min(len - 1, c / HOP);
    float acc = 0.f, env = 0.f;
    for (int t = t_hi; t >= 0 && c - HOP * t < N_FFT; --t) {
      const int i = c - HOP * t;
      const float w = 0.5f - 0.5f * __ldg(&tw1024g[i]).x;
      acc += frames[((size_t)b * Tp + t) * N_FFT + i];
      env += w * w;
    }
    out = acc / env;
  }
  wav[(size_t)b * n_out + n] = out;
}

__global__ void vc_lens_kernel(const int64_t* __restrict__ lens, int B, int T, int* __restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) out[b] = lens ? (int)max(0ll, min((long long)lens[b], (long long)T)) : T;
}

#define AU_TRY(expr)                                                                   \
  do {                                                                                 \
    int _rc = (expr);                                                                  \
    if (_rc) {                                                                         \
      if (e.err.empty()) {                                                             \
        char _b[512];                                                                  \
        snprintf(_b, sizeof(_b), "%s failed with %d at %s:%d", #expr, _rc, __FILE__, __LINE__); \
        e.err = _b;                                                                    \
      }                                                                                \
      return _rc;                                                                      \
    }                                                                                  \
  } while (0)
#define AU_K(cls, work, s, ...)                  \
  do {                                           \
    ++g_launch_count;                            \
    if (e.prof.on) e.prof.begin(cls, work, s);   \
    __VA_ARGS__;                                 \
    if (e.prof.on) e.prof.end(s);                \
    AU_TRY((int)cudaGetLastError());             \
  } while (0)

static Audio& audio_of(Engine& e) {
  if (!e.audio) e.audio = new Audio();
  return *static_cast<Audio*>(e.audio);
}

void audio_destroy(Engine& e) {
  if (!e.audio) return;
  Audio* a = static_cast<Audio*>(e.audio);
  for (void* p : a->ws_allocs) cudaFree(p);
  delete a;
  e.audio = nullptr;
}

static const HostTensor* vfind(Engine& e, const std::string& name) {
  auto it = e.raw.find("vocoder." + name);
  if (it == e.raw.end()) {
    e.err = "missing vocoder weight: vocoder." + name;
    return nullptr;
  }
  return &it->second;
}
static int vupload(Engine& e, const std::string& name, size_t numel, float** out) {
  const HostTensor* t = vfind(e, name);
  if (!t) return 940;
  if (t->numel() != numel) {
    e.err = "unexpected shape for vocoder." + name;
    return 944;
  }
  AU_TRY(dev_alloc(e, e.weight_allocs, (void**)out, numel * 4));
  AU_TRY((int)cudaMemcpy(*out, t->data.data(), numel * 4, cudaMemcpyHostToDevice));
  return 0;
}

static int vocoder_build(Engine& e, Audio& a) {
  if (a.has_vocoder) return 0;
  const HostTensor* ew = vfind(e, "backbone.embed.weight");
  const HostTensor* eb = vfind(e, "backbone.embed.bias");
  if (!ew || !eb) return 940;
  if (ew->shape.size() != 3 || ew->shape[0] != VC_DIM || ew->shape[1] != N_MELS || ew->shape[2] != VC_K) {
    e.err = "vocoder.backbone.embed.weight must be (512, 100, 7)";
    return 944;
  }
  {  // (out, in, k) -> [out][k * 104 + in]: the implicit-im2col column order
    std::vector<float> w((size_t)VC_DIM * VC_K * VC_MELP, 0.f);
    for (int n = 0; n < VC_DIM; ++n)
      for (int c = 0; c < N_MELS; ++c)
        for (int k = 0; k < VC_K; ++k)
          w[(size_t)n * VC_K * VC_MELP + k * VC_MELP + c] = ew->data[((size_t)n * N_MELS + c) * VC_K + k];
    AU_TRY(pack_linear_host(e, w.data(), eb->data.data(), VC_DIM, VC_K * VC_MELP, EPI_F32, &a.embed, 0, nullptr));
  }
  AU_TRY(vupload(e, "backbone.norm.weight", VC_DIM, &a.norm_w));
  AU_TRY(vupload(e, "backbone.norm.bias", VC_DIM, &a.norm_b));
  AU_TRY(vupload(e, "backbone.final_layer_norm.weight", VC_DIM, &a.fnorm_w));
  AU_TRY(vupload(e, "backbone.final_layer_norm.bias", VC_DIM, &a.fnorm_b));
  for (int i = 0; i < VC_LAYERS; ++i) {
    const std::string p = "backbone.convnext." + std::to_string(i);
    AU_TRY(vupload(e, p + ".dwconv.weight", (size_t)VC_DIM * VC_K, &a.dw_w[i]));
    AU_TRY(vupload(e, p + ".dwconv.bias", VC_DIM, &a.dw_b[i]));
    AU_TRY(vupload(e, p + ".norm.weight", VC_DIM, &a.ln_w[i]));
    AU_TRY(vupload(e, p + ".norm.bias", VC_DIM, &a.ln_b[i]));
    const HostTensor *w1 = vfind(e, p + ".pwconv1.weight"), *b1 = vfind(e, p + ".pwconv1.bias");
    const HostTensor *w2 = vfind(e, p + ".pwconv2.weight"), *b2 = vfind(e, p + ".pwconv2.bias");
    const HostTensor* gm = vfind(e, p + ".gamma");
    if (!w1 || !b1 || !w2 || !b2 || !gm) return 940;
    AU_TRY(pack_linear_host(e, w1->data.data(), b1->data.data(), VC_INTER, VC_DIM, EPI_GELU_BF16, &a.pw1[i], 0, nullptr));
    // gamma * (W2 g + b2): the layer scale is folded into the rows of pwconv2
    AU_TRY(pack_linear_host(e, w2->data.data(), b2->data.data(), VC_DIM, VC_INTER, EPI_RESIDUAL, &a.pw2[i], 0,
                            gm->data.data()));
  }
  const HostTensor *hw = vfind(e, "head.out.weight"), *hb = vfind(e, "head.out.bias");
  if (!hw || !hb) return 940;
  AU_TRY(pack_linear_host(e, hw->data.data(), hb->data.data(), N_FFT + 2, VC_DIM, EPI_F32, &a.head, 0, nullptr));
  a.has_vocoder = true;
  return 0;
}

static int vocoder_reserve(Engine& e, Audio& a, int B, int T) {
  const int Tp = T + 6;
  size_t rows = (size_t)B * Tp;
  if (rows <= a.cap_rows && B <= a.cap_B) return 0;
  // capacity only grows, in both dimensions: a sweep that alternates many short and few long clips must not
  // shrink one of them and reallocate (and clear) gigabytes on every call
  if (rows < a.cap_rows) rows = a.cap_rows;
  if (B < a.cap_B) B = a.cap_B;
  cudaDeviceSynchronize();
  for (void* p : a.ws_allocs) cudaFree(p);
  a.ws_allocs.clear();
  a.cap_rows = rows, a.cap_B = B;
  auto A = [&](void** p, size_t bytes) {
    int rc = dev_alloc(e, a.ws_allocs, p, bytes);
    if (!rc) rc = (int)cudaMemset(*p, 0, bytes);
    return rc;
  };
  AU_TRY(A((void**)&a.melp, (rows + 8) * VC_MELP * 2));
  AU_TRY(A((void**)&a.x, rows * VC_DIM * 4));
  AU_TRY(A((void**)&a.emb, rows * VC_DIM * 4));
  AU_TRY(A((void**)&a.h, rows * VC_DIM * 2));
  AU_TRY(A((void**)&a.g, rows * VC_INTER * 2));
  AU_TRY(A((void**)&a.spec, rows * 1028 * 4));
  AU_TRY(A((void**)&a.frames, rows * N_FFT * 4));
  AU_TRY(A((void**)&a.lens_dev, (size_t)B * 4));
  return 0;
}

}  // namespace zv

namespace zv {

// torchaudio sinc_interp_hann resampling (_apply_sinc_resample_kernel): the waveform is zero-padded by `width` on
// the left, and output sample k * new + p is the dot product of phase filter p with K input samples starting at
// k * orig - width.  One thread per output sample; the filter bank ([new][K] fp32, a few KB) stays in L1.
__global__ void resample_kernel(const float* __restrict__ wav, const int64_t* __restrict__ lens, int N,
                                const float* __restrict__ bank, int orig, int nw, int width, int K, int n_out,
                                float* __restrict__ out) {
  const int b = blockIdx.y;
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_out) return;
  const int64_t len = lens ? min((int64_t)N, lens[b]) : (int64_t)N;
  const int64_t valid_out = (len * nw + orig - 1) / orig;  // ceil(new * len / orig)
  float acc = 0.f;
  if (n < valid_out) {
    const int p = n % nw, k = n / nw;
    const float* w = bank + (size_t)p * K;
    const float* x = wav + (size_t)b * N;
    const int64_t base = (int64_t)k * orig - width;
    const int j0 = base < 0 ? (int)-base : 0;
    const int j1 = (int)min((int64_t)K, len - base);
    for (int j = j0; j < j1; ++j) acc = fmaf(__ldg(w + j), __ldg(x + base + j), acc);
  }
  out[(size_t)b * n_out + n] = acc;
}

// _get_sinc_resample_kernel(orig, new, gcd, lowpass_filter_width=6, rolloff=0.99, "sinc_interp_hann") in double,
// stored as fp32 like torchaudio does.
static int resample_bank(Engine& e, Audio& a, int orig_freq, int new_freq, const Audio::ResampleBank** out) {
  int g = orig_freq, h = new_freq;
  while (h) {
    const int t = g % h;
    g = h, h = t;
  }
  const int orig = orig_freq / g, nw = new_freq / g;
  for (const auto& bk : a.banks)
    if (bk.orig == orig && bk.nw == nw) {
      *out = &bk;
      return 0;
    }
  const double lpw = 6.0, rolloff = 0.99;
  const double base_freq = (double)(orig < nw ? orig : nw) * rolloff;
  Audio::ResampleBank bk;
  bk.orig = orig, bk.nw = nw;
  bk.width = (int)ceil(lpw * orig / base_freq);
  bk.K = 2 * bk.width + orig;
  std::vector<float> w((size_t)nw * bk.K);
  const double scale = base_freq / orig;
  for (int p = 0; p < nw; ++p)
    for (int j = 0; j < bk.K; ++j) {
      const double idx = (double)(j - bk.width) / orig;
      double t = ((double)(-p) / nw + idx) * base_freq;
      t = t < -lpw ? -lpw : (t > lpw ? lpw : t);
      const double c = cos(t * M_PI / lpw / 2.0);
      const double window = c * c;
      t *= M_PI;
      const double sinc = t == 0.0 ? 1.0 : sin(t) / t;
      w[(size_t)p * bk.K + j] = (float)(sinc * (window * scale));
    }
  AU_TRY(dev_alloc(e, e.weight_allocs, (void**)&bk.w, w.size() * 4));
  AU_TRY((int)cudaMemcpy(bk.w, w.data(), w.size() * 4, cudaMemcpyHostToDevice));
  a.banks.push_back(bk);
  *out = &a.banks.back();
  return 0;
}

}  // namespace zv

using namespace zv;

static int fbank_launch(void* handle, const float* wav, const int64_t* lens, int B, int N, float* mel_out,
                        void* stream, bool bigvgan) {
  Engine* ep = (Engine*)handle;
  if (!ep) return 960;
  Engine& e = *ep;
  e.err.clear();
  if (B <= 0 || N <= N_FFT / 2) {
    e.err = "zv_fbank: need B > 0 and more than 512 samples (reflect padding)";
    return 964;
  }
  cudaStream_t s = (cudaStream_t)stream;
  Audio& a = audio_of(e);
  AU_TRY(audio_tables(e, a));
  const int T = (N + HOP / 2) / HOP;
  const int fpc = 16;
  static unsigned long long attr = 0;
  if (first_use_on_device(attr)) {
    cudaFuncSetAttribute(fbank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FftSmem));
  }
  dim3 grid((T + fpc - 1) / fpc, B);
  // algorithmic bytes: 4 B per input sample + 400 B per output frame
  const FbankVariant var = bigvgan ? FbankVariant{(N_FFT - HOP) / 2, 0, 1e-9f, 1e-5f} : FbankVariant{N_FFT / 2, 1, 0.f, 1e-7f};
  AU_K(KC_FBANK, (double)B * N * 4.0 + (double)B * T * N_MELS * 4.0, s,
       (fbank_kernel<<<grid, 256, sizeof(FftSmem), s>>>(
           wav, lens, N, T, a.tw512, a.tw1024, bigvgan ? a.sl_start : a.fb_start, bigvgan ? a.sl_count : a.fb_count,
           bigvgan ? a.sl_off : a.fb_off, bigvgan ? a.sl_w : a.fb_w, fpc, var, mel_out)));
  return 0;
}

extern "C" int zv_fbank(void* handle, const float* wav, const int64_t* lens, int B, int N, float* mel_out,
                        void* stream) {
  return fbank_launch(handle, wav, lens, B, N, mel_out, stream, false);
}

extern "C" int zv_fbank_bigvgan(void* handle, const float* wav, const int64_t* lens, int B, int N, float* mel_out,
                                void* stream) {
  return fbank_launch(handle, wav, lens, B, N, mel_out, stream, true);
}

extern "C" int zv_vocos_decode(void* handle, const float* mel, const int64_t* lens, int B, int T, float* wav_out,
                               void* stream) {
  Engine* ep = (Engine*)handle;
  if (!ep) return 960;
  Engine& e = *ep;
  e.err.clear();
  if (B <= 0 || T < 2) {
    e.err = "zv_vocos_decode: need B > 0 and T >= 2 frames";
    return 964;
  }
  cudaStream_t s = (cudaStream_t)stream;
  Audio& a = audio_of(e);
  AU_TRY(audio_tables(e, a));
  AU_TRY(vocoder_build(e, a));
  AU_TRY(vocoder_reserve(e, a, B, T));
  const int Tp = T + 6, M = B * Tp;
  static unsigned long long attr = 0;
  if (first_use_on_device(attr)) {
    cudaFuncSetAttribute(vc_istft_frames_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FftSmem));
  }
  AU_K(KC_VOCODER_ELEM, B * 12.0, s, (vc_lens_kernel<<<(B + 127) / 128, 128, 0, s>>>(lens, B, T, a.lens_dev)));
  const size_t tot = (size_t)(M + 6) * VC_MELP;
  AU_K(KC_VOCODER_ELEM, (double)B * T * N_MELS * 4.0 + tot * 2.0, s,
       (vc_pack_mel_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, s>>>(mel, a.lens_dev, B, T, Tp, a.melp, tot)));
  EpiParams p0;
  p0.L = Tp, p0.out_f32 = a.emb, p0.ld_f32 = VC_DIM;
  AU_TRY(run_gemm(e, a.embed, a.melp, VC_MELP, M, p0, s));
  AU_K(KC_VOCODER_ELEM, (double)M * VC_DIM * 8.0, s,
       (vc_layernorm_kernel<<<(M + 7) / 8, 256, 0, s>>>(a.emb, a.norm_w, a.norm_b, M, a.x, nullptr)));
  // longest run of frames per CTA that still leaves >= 8 CTAs for each of the 148 SMs (6 halo rows per run)
  int dw_run = 32;
  while (dw_run > 8 && (long long)B * ((T + dw_run - 1) / dw_run) < 8ll * 148) dw_run >>= 1;
  if (const char* ev = getenv("ZV_DW_RUN")) dw_run = atoi(ev);
  dim3 dgrid((T + dw_run - 1) / dw_run, B);
  for (int i = 0; i < VC_LAYERS; ++i) {
    AU_K(KC_VOCODER_ELEM, (double)B * T * VC_DIM * 6.0, s,
         (vc_dwconv_ln_kernel<<<dgrid, 128, 0, s>>>(a.x, a.lens_dev, T, Tp, dw_run, a.dw_w[i], a.dw_b[i],
                                                    a.ln_w[i], a.ln_b[i], a.h)));
    EpiParams p1;
    p1.L = Tp, p1.out_bf16 = a.g, p1.ld_bf16 = VC_INTER;
    AU_TRY(run_gemm(e, a.pw1[i], a.h, VC_DIM, M, p1, s));
    EpiParams p2;
    p2.L = Tp, p2.res = a.x, p2.out_f32 = a.x, p2.ld_f32 = VC_DIM;
    AU_TRY(run_gemm(e, a.pw2[i], a.g, VC_INTER, M, p2, s));
  }
  AU_K(KC_VOCODER_ELEM, (double)M * VC_DIM * 6.0, s,
       (vc_layernorm_kernel<<<(M + 7) / 8, 256, 0, s>>>(a.x, a.fnorm_w, a.fnorm_b, M, nullptr, a.h)));
  EpiParams ph;
  ph.L = Tp, ph.out_f32 = a.spec, ph.ld_f32 = 1028;
  AU_TRY(run_gemm(e, a.head, a.h, VC_DIM, M, ph, s));
  const int fpc = 16;
  dim3 igrid((T + fpc - 1) / fpc, B);
  AU_K(KC_VOCODER_ELEM, (double)B * T * (1026.0 + 1024.0) * 4.0, s,
       (vc_istft_frames_kernel<<<igrid, 256, sizeof(FftSmem), s>>>(a.spec, 1028, a.lens_dev, T, Tp, a.tw512,
                                                                   a.tw1024, fpc, a.frames)));
  const int n_out = HOP * (T - 1);
  dim3 ogrid((n_out + 255) / 256, B);
  AU_K(KC_VOCODER_ELEM, (double)B * T * N_FFT * 4.0 + (double)B * n_out * 4.0, s,
       (vc_ola_kernel<<<ogrid, 256, 0, s>>>(a.frames, a.lens_dev, T, Tp, a.tw1024, n_out, wav_out)));
  return 0;
}

extern "C" int zv_resample(void* handle, const float* wav, const int64_t* lens, int B, int N, int orig_freq,
                           int new_freq, float* out, int n_out, void* stream) {
  Engine* ep = (Engine*)handle;
  if (!ep) return 960;
  Engine& e = *ep;
  e.err.clear();
  if (B <= 0 || N <= 0 || orig_freq <= 0 || new_freq <= 0) {
    e.err = "zv_resample: need B, N, orig_freq, new_freq > 0";
    return 964;
  }
  const long long want = ((long long)N * new_freq + orig_freq - 1) / orig_freq;
  if (n_out != want) {
    e.err = "zv_resample: n_out must be ceil(new_freq * N / orig_freq)";
    return 964;
  }
  cudaStream_t s = (cudaStream_t)stream;
  Audio& a = audio_of(e);
  const Audio::ResampleBank* bk = nullptr;
  AU_TRY(resample_bank(e, a, orig_freq, new_freq, &bk));
  dim3 grid((n_out + 255) / 256, B);
  AU_K(KC_FBANK, (double)B * N * 4.0 + (double)B * n_out * 4.0, s,
       (resample_kernel<<<grid, 256, 0, s>>>(wav, lens, N, bk->w, bk->orig, bk->nw, bk->width, bk->K, n_out, out)));
  return 0;
}

// ------------------------------------------------------------------ zv_resample_static (no allocation)
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops).  zv_resample above cudaMallocs + cudaMemcpys its filter bank
// on the first use of a rate pair; this variant treats the bank as scratch that every call rebuilds on the device
// into the file-scope array below, so the call allocates nothing and copies nothing.
//
// g_resample_static_scratch is a HOST-SIDE STAND-IN in host code, not a device address: handing it to a kernel
// would fault (or silently use the wrong memory where the GPU can reach host memory).  Host code gets the real
// address once per device from cudaGetSymbolAddress (resample_static_setup) and passes THAT to the kernels.
// Every GPU that loads the library holds these 256 KB until the process exits.  All calls on a GPU share the array
// and it keeps the last call's contents; nothing has to start at zero because resample_bank_static_kernel writes
// every one of the nw * K entries resample_kernel reads.  Calls that overlap on the GPU race on it.
namespace zv {

static __device__ __align__(256) float g_resample_static_scratch[ZV_RESAMPLE_STATIC_MAX_FLOATS];

static constexpr int RS_MAX_DEVICES = 64;
static float* g_resample_static_ptr[RS_MAX_DEVICES];  // device address of the array on device i; null = not set up

// _get_sinc_resample_kernel in double, the expression order of resample_bank() above; one thread per bank entry.
__global__ void resample_bank_static_kernel(float* __restrict__ bank, int orig, int nw, int width, int K) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nw * K) return;
  const int p = i / K, j = i - p * K;
  const double lpw = 6.0, rolloff = 0.99;
  const double base_freq = (double)(orig < nw ? orig : nw) * rolloff;
  const double scale = base_freq / orig;
  const double idx = (double)(j - width) / orig;
  double t = ((double)(-p) / nw + idx) * base_freq;
  t = t < -lpw ? -lpw : (t > lpw ? lpw : t);
  const double c = cos(t * M_PI / lpw / 2.0);
  const double window = c * c;
  t *= M_PI;
  const double sinc = t == 0.0 ? 1.0 : sin(t) / t;
  bank[i] = (float)(sinc * (window * scale));
}

// Once per device, with that device current: the device address of the scratch array.
static int resample_static_setup(Engine& e, float** out) {
  int dev = -1;
  AU_TRY((int)cudaGetDevice(&dev));
  if (dev < 0 || dev >= RS_MAX_DEVICES) {
    e.err = "zv_resample_static: device ordinal out of range";
    return 966;
  }
  if (!g_resample_static_ptr[dev]) {
    float* p = nullptr;
    AU_TRY((int)cudaGetSymbolAddress((void**)&p, g_resample_static_scratch));
    if (!p) {
      e.err = "zv_resample_static: cudaGetSymbolAddress returned a null address";
      return 966;
    }
    g_resample_static_ptr[dev] = p;
  }
  *out = g_resample_static_ptr[dev];
  return 0;
}

}  // namespace zv

extern "C" int zv_resample_static(void* handle, const float* wav, const int64_t* lens, int B, int N, int orig_freq,
                                  int new_freq, float* out, int n_out, void* stream) {
  Engine* ep = (Engine*)handle;
  if (!ep) return 960;
  Engine& e = *ep;
  e.err.clear();
  if (B <= 0 || N <= 0 || orig_freq <= 0 || new_freq <= 0) {
    e.err = "zv_resample_static: need B, N, orig_freq, new_freq > 0";
    return 964;
  }
  const long long want = ((long long)N * new_freq + orig_freq - 1) / orig_freq;
  if (n_out != want) {
    e.err = "zv_resample_static: n_out must be ceil(new_freq * N / orig_freq)";
    return 964;
  }
  int g = orig_freq, h = new_freq;
  while (h) {
    const int t = g % h;
    g = h, h = t;
  }
  const int orig = orig_freq / g, nw = new_freq / g;
  const double base_freq = (double)(orig < nw ? orig : nw) * 0.99;
  const int width = (int)ceil(6.0 * orig / base_freq);
  const long long K = 2LL * width + orig;
  if ((long long)nw * K > (long long)ZV_RESAMPLE_STATIC_MAX_FLOATS || B > 65535) {
    char msg[320];
    snprintf(msg, sizeof(msg),
             "zv_resample_static: %d -> %d Hz needs a filter bank of %d x %lld = %lld floats (B = %d); the static scratch "
             "holds ZV_RESAMPLE_STATIC_MAX_FLOATS = %d and B <= 65535 -- use zv_resample", orig_freq, new_freq, nw, K,
             (long long)nw * K, B, ZV_RESAMPLE_STATIC_MAX_FLOATS);
    e.err = msg;
    return 967;
  }
  float* bank = nullptr;
  AU_TRY(resample_static_setup(e, &bank));
  cudaStream_t s = (cudaStream_t)stream;
  const int n_bank = nw * (int)K;
  g_launch_count += 2;  // no zv_prof events here: the profiler's event pool allocates
  resample_bank_static_kernel<<<(n_bank + 255) / 256, 256, 0, s>>>(bank, orig, nw, width, (int)K);
  AU_TRY((int)cudaGetLastError());
  dim3 grid((n_out + 255) / 256, B);
  resample_kernel<<<grid, 256, 0, s>>>(wav, lens, N, bank, orig, nw, width, (int)K, n_out, out);
  AU_TRY((int)cudaGetLastError());
  return 0;
}
