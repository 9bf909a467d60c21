// Weight packing (finalize) and the per-NFE launch schedule of the TTSZipformer
// (reference: zipvoice/models/modules/zipformer.py:242-293, 489-642, 702-744, 850-870).
#include "zv_engine.h"

#include <math.h>

#include <algorithm>
#include <stdio.h>
#include <string.h>

namespace zv {

long long g_launch_count = 0;

#define ZV_TRY(expr)                                                                  \
  do {                                                                                \
    int _rc = (expr);                                                                 \
    if (_rc) {                                                                        \
      char _b[640];                                                                   \
      snprintf(_b, sizeof(_b), "%s failed with %d at %s:%d", #expr, _rc, __FILE__, __LINE__); \
      e.err = _b;                                                                     \
      return _rc;                                                                     \
    }                                                                                 \
  } while (0)
// One kernel launch of class `cls` doing `work` algorithmic FLOPs / bytes.
// ZV_SYNC_EACH=1 (debugging): synchronise after every launch outside stream capture and name the class that failed
static const bool g_sync_each = getenv("ZV_SYNC_EACH") && atoi(getenv("ZV_SYNC_EACH")) != 0;
static int sync_each_check(Engine& e, int cls, cudaStream_t s);
const char* kclass_name(int c);
#define ZV_K(cls, work, s, expr)              \
  do {                                        \
    ++g_launch_count;                         \
    if (e.prof.on) e.prof.begin(cls, work, s); \
    int _krc = (expr);                        \
    if (e.prof.on) e.prof.end(s);             \
    if (_krc) fprintf(stderr, "[zv] launch %lld (class %s) returned %d\n", (long long)g_launch_count, kclass_name(cls), _krc); \
    ZV_TRY(_krc);                             \
    if (g_sync_each) ZV_TRY(sync_each_check(e, cls, s)); \
  } while (0)

static const char* kNames[KC_COUNT] = {"gemm_bias", "gemm_swooshl", "gemm_glu", "gemm_nonlin", "gemm_residual",
                                       "gemm_f32", "gemm_gelu", "attn_self", "attn_nonlin", "dwconv_swooshr", "biasnorm_bypass",
                                       "resample", "pack_input", "cfg_euler", "time_embed", "text_glue", "fbank",
                                       "vocoder_elem"};
const char* kclass_name(int c) { return c >= 0 && c < KC_COUNT ? kNames[c] : "?"; }
static int sync_each_check(Engine& e, int cls, cudaStream_t s) {
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(s, &cap);
  if (cap != cudaStreamCaptureStatusNone) return 0;
  const cudaError_t rc = cudaStreamSynchronize(s);
  if (rc != cudaSuccess) {
    fprintf(stderr, "[zv] ZV_SYNC_EACH: launch %lld of class %s failed: %s\n", (long long)g_launch_count,
            kclass_name(cls), cudaGetErrorString(rc));
    e.err = std::string("kernel of class ") + kclass_name(cls) + " failed: " + cudaGetErrorString(rc);
    return (int)rc;
  }
  return 0;
}
bool kclass_is_tensor(int c) { return c <= KC_ATTN_NONLIN; }

cudaEvent_t Prof::get() {
  if (used == pool.size()) {
    cudaEvent_t ev;
    cudaEventCreate(&ev);
    pool.push_back(ev);
  }
  return pool[used++];
}
void Prof::begin(int cls, double work, cudaStream_t s) {
  Rec r{cls, work, get(), get()};
  cudaEventRecord(r.a, s);
  recs.push_back(r);
}
void Prof::end(cudaStream_t s) { cudaEventRecord(recs.back().b, s); }
int Prof::read(double* ms, double* work, long long* launches) {
  int rc = (int)cudaDeviceSynchronize();
  for (int i = 0; i < KC_COUNT; ++i) ms[i] = 0.0, work[i] = 0.0, launches[i] = 0;
  for (const Rec& r : recs) {
    float t = 0.f;
    if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess) ms[r.cls] += t;
    work[r.cls] += r.work, launches[r.cls] += 1;
  }
  recs.clear();
  used = 0;
  return rc;
}
void Prof::release() {
  for (cudaEvent_t ev : pool) cudaEventDestroy(ev);
  pool.clear(), recs.clear(), used = 0;
}

int dev_alloc(Engine& e, std::vector<void*>& pool, void** p, size_t bytes) {
  if (bytes == 0) bytes = 16;
  cudaError_t rc = cudaMalloc(p, bytes);
  if (rc != cudaSuccess) {
    e.err = std::string("cudaMalloc failed: ") + cudaGetErrorString(rc);
    return (int)rc;
  }
  pool.push_back(*p);
  return 0;
}

static const HostTensor* find(Engine& e, const std::string& name) {
  auto it = e.raw.find(name);
  if (it == e.raw.end()) {
    e.err = "missing weight: " + name;
    return nullptr;
  }
  return &it->second;
}

static int upload_f32(Engine& e, const std::vector<float>& v, float** out) {
  ZV_TRY(dev_alloc(e, e.weight_allocs, (void**)out, v.size() * sizeof(float)));
  ZV_TRY((int)cudaMemcpy(*out, v.data(), v.size() * sizeof(float), cudaMemcpyHostToDevice));
  return 0;
}
static int upload_named(Engine& e, const std::string& name, float** out) {
  const HostTensor* t = find(e, name);
  if (!t) return 940;
  return upload_f32(e, t->data, out);
}

// Row permutation applied to the (out_features) axis for the fused epilogues.
//   GLU     (conv in_proj, rows = [x (D) | s (D)])           -> groups of 32 = [16 x | 16 s]
//   NONLIN  (nonlin in_proj, rows = [s (h) | x (h) | y (h)])  -> groups of 32 = [16 s | 16 x], then y
static std::vector<int> row_perm(int mode, int N) {
  std::vector<int> p(N);
  if (mode == EPI_GLU_BF16) {
    const int D = N / 2;
    for (int g = 0; g < D / 16; ++g)
      for (int i = 0; i < 16; ++i) p[g * 32 + i] = g * 16 + i, p[g * 32 + 16 + i] = D + g * 16 + i;
  } else if (mode == EPI_NONLIN_BF16) {
    const int h = N / 3;
    for (int g = 0; g < h / 16; ++g)
      for (int i = 0; i < 16; ++i) p[g * 32 + i] = g * 16 + i, p[g * 32 + 16 + i] = h + g * 16 + i;
    for (int i = 0; i < h; ++i) p[2 * h + i] = 2 * h + i;
  } else {
    for (int i = 0; i < N; ++i) p[i] = i;
  }
  return p;
}

static inline uint16_t f2bf(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;
  u += 0x7fffu + ((u >> 16) & 1u);  // round to nearest even
  return (uint16_t)(u >> 16);
}

int pack_linear_host(Engine& e, const float* w, const float* bias_in, int N, int K, int mode, Lin* lin, int k_pad_to,
                     const float* row_scale) {
  lin->N = N, lin->K = k_pad_to ? k_pad_to : K, lin->K_valid = K, lin->mode = mode;
  lin->gate_cols = mode == EPI_NONLIN_BF16 ? 2 * (N / 3) : 0;
  if ((mode == EPI_GLU_BF16 && (N / 2) % 16) || (mode == EPI_NONLIN_BF16 && (N / 3) % 16)) {
    e.err = "fused epilogue needs channel counts that are multiples of 16";
    return 941;
  }
  lin->block_n = gemm_pick_block_n(N, mode, lin->gate_cols);
  lin->N_pad = (N + lin->block_n - 1) / lin->block_n * lin->block_n;
  lin->ldw = (lin->K + 63) / 64 * 64;
  const std::vector<int> perm = row_perm(mode, N);
  std::vector<uint16_t> wb((size_t)lin->N_pad * lin->ldw, 0);
  for (int n = 0; n < N; ++n) {
    const float* src = w + (size_t)perm[n] * K;
    const float sc = row_scale ? row_scale[perm[n]] : 1.0f;
    uint16_t* dst = wb.data() + (size_t)n * lin->ldw;
    for (int k = 0; k < K; ++k) dst[k] = f2bf(src[k] * sc);
  }
  ZV_TRY(dev_alloc(e, e.weight_allocs, (void**)&lin->w, wb.size() * 2));
  ZV_TRY((int)cudaMemcpy(lin->w, wb.data(), wb.size() * 2, cudaMemcpyHostToDevice));
  std::vector<float> bias(lin->N_pad, 0.f);
  if (bias_in)
    for (int n = 0; n < N; ++n) bias[n] = bias_in[perm[n]] * (row_scale ? row_scale[perm[n]] : 1.0f);
  ZV_TRY(upload_f32(e, bias, &lin->bias));
  return 0;
}

static int pack_linear(Engine& e, const std::string& name, int mode, Lin* lin, int k_pad_to = 0) {
  const HostTensor* w = find(e, name + ".weight");
  if (!w) return 940;
  auto it = e.raw.find(name + ".bias");
  int rc = pack_linear_host(e, w->data.data(), it != e.raw.end() ? it->second.data.data() : nullptr, (int)w->shape[0],
                            (int)w->shape[1], mode, lin, k_pad_to, nullptr);
  if (rc && e.err.find(name) == std::string::npos) e.err += ": " + name;
  return rc;
}

// SelfAttention.out_proj is folded into the ConvolutionModule that follows it (zipformer.py:564-587,600-624):
//   x3 = x2 + Wso o + bso (+ te),   x4 = x3 + Wco conv(GLU(Wci x3 + bci)) + bco
// becomes, with o = the 48-wide attention output kept next to the conv hidden state,
//   GLU input = [x2 | o] [Wci | Wci Wso]^T + (bci + Wci bso) (+ Wci te as a per-batch row bias)
//   x4        = x2 + [h | o] [Wco | Wso]^T + (bco + bso) (+ te)
// so the stream makes ONE fp32 round trip through HBM for the two modules instead of two.
static int fold_self_attn_into_conv(Engine& e, const std::string& so, const std::string& c, Lin* conv_in,
                                    Lin* conv_out) {
  const HostTensor *wso = find(e, so + ".weight"), *bso = find(e, so + ".bias");
  const HostTensor *wci = find(e, c + ".in_proj.weight"), *bci = find(e, c + ".in_proj.bias");
  const HostTensor *wco = find(e, c + ".out_proj.weight"), *bco = find(e, c + ".out_proj.bias");
  if (!wso || !bso || !wci || !bci || !wco || !bco) return 940;
  const int D = (int)wso->shape[0], V = (int)wso->shape[1], N2 = (int)wci->shape[0];
  if ((int)wci->shape[1] != D || (int)wco->shape[0] != D || (int)wco->shape[1] != D || V > 64 || D % 64) {
    e.err = "unexpected conv / self-attention projection shapes at " + c;
    return 944;
  }
  const int Ke = D + V;
  std::vector<float> wi((size_t)N2 * Ke), bi(N2), wo((size_t)D * Ke), bo(D);
  for (int n = 0; n < N2; ++n) {
    const float* w = &wci->data[(size_t)n * D];
    float* dst = &wi[(size_t)n * Ke];
    memcpy(dst, w, (size_t)D * 4);
    double accb = bci->data[n];
    for (int d = 0; d < D; ++d) accb += (double)w[d] * bso->data[d];
    bi[n] = (float)accb;
    for (int j = 0; j < V; ++j) {
      double acc = 0.0;
      for (int d = 0; d < D; ++d) acc += (double)w[d] * wso->data[(size_t)d * V + j];
      dst[D + j] = (float)acc;
    }
  }
  for (int n = 0; n < D; ++n) {
    memcpy(&wo[(size_t)n * Ke], &wco->data[(size_t)n * D], (size_t)D * 4);
    memcpy(&wo[(size_t)n * Ke + D], &wso->data[(size_t)n * V], (size_t)V * 4);
    bo[n] = bco->data[n] + bso->data[n];
  }
  ZV_TRY(pack_linear_host(e, wi.data(), bi.data(), N2, Ke, EPI_GLU_BF16, conv_in, D + 64, nullptr));
  conv_in->K_valid = D;  // algorithmic FLOPs of the reference in_proj; the folded 48 columns are counted below
  ZV_TRY(pack_linear_host(e, wo.data(), bo.data(), D, Ke, EPI_RES24, conv_out, D + 64, nullptr));
  return 0;
}

static int build_layer(Engine& e, Net& net, const std::string& p, int kernel, Layer* L) {
  {
    // The attention kernels take their scores in log2 units (ATTN_LOG2E, zv_kernels.h): the q rows of the in_proj
    // (rows [0, H * 32) of [q | k | p], zipformer.py:1220-1233) and linear_pos are scaled by log2(e) once, here.
    const HostTensor* w = find(e, p + ".self_attn_weights.in_proj.weight");
    const HostTensor* b = find(e, p + ".self_attn_weights.in_proj.bias");
    const HostTensor* wp = find(e, p + ".self_attn_weights.linear_pos.weight");
    if (!w || !b || !wp) return 940;
    const int N = (int)w->shape[0], K = (int)w->shape[1], nq = net.H * 32;
    if (N < 2 * nq) {
      e.err = "unexpected attention in_proj shape at " + p;
      return 944;
    }
    std::vector<float> ws(w->data), bs(b->data), wps(wp->data);
    for (size_t i = 0; i < (size_t)nq * K; ++i) ws[i] *= ATTN_LOG2E;
    for (int i = 0; i < nq; ++i) bs[i] *= ATTN_LOG2E;
    for (float& v : wps) v *= ATTN_LOG2E;
    ZV_TRY(pack_linear_host(e, ws.data(), bs.data(), N, K, EPI_BIAS_BF16, &L->attn_in, 0, nullptr));
    ZV_TRY(upload_f32(e, wps, &L->w_pos));
    // [feed_forward1.in_proj | attention in_proj] as one weight matrix (Layer::ff1_qkp)
    const HostTensor* w1 = find(e, p + ".feed_forward1.in_proj.weight");
    const HostTensor* b1 = find(e, p + ".feed_forward1.in_proj.bias");
    if (!w1 || !b1) return 940;
    const int N1 = (int)w1->shape[0];
    if ((int)w1->shape[1] == K && N1 % 32 == 0) {
      std::vector<float> wc(w1->data), bc(b1->data);
      wc.insert(wc.end(), ws.begin(), ws.end());
      bc.insert(bc.end(), bs.begin(), bs.end());
      ZV_TRY(pack_linear_host(e, wc.data(), bc.data(), N1 + N, K, EPI_SWOOSHL_BF16, &L->ff1_qkp, 0, nullptr));
      L->ff1_qkp.gate_cols = N1;
    }
  }
  for (int i = 0; i < 3; ++i) {
    const std::string f = p + ".feed_forward" + std::to_string(i + 1);
    ZV_TRY(pack_linear(e, f + ".in_proj", EPI_SWOOSHL_BF16, &L->ff_in[i]));
    ZV_TRY(pack_linear(e, f + ".out_proj", EPI_RES24, &L->ff_out[i]));
  }
  ZV_TRY(pack_linear(e, p + ".nonlin_attention.in_proj", EPI_NONLIN_BF16, &L->na_in));
  ZV_TRY(pack_linear(e, p + ".nonlin_attention.out_proj", EPI_RES24, &L->na_out));
  for (int i = 0; i < 2; ++i) {
    const std::string a = p + ".self_attn" + std::to_string(i + 1);
    {  // value projection padded to 16 columns per head (12 + 4 zero rows): 16-byte aligned TMA boxes per head
      const HostTensor* w = find(e, a + ".in_proj.weight");
      const HostTensor* bv = find(e, a + ".in_proj.bias");
      if (!w || !bv) return 940;
      const int dv = e.cfg.value_head_dim, K = (int)w->shape[1], Hh = (int)w->shape[0] / dv;
      std::vector<float> wp((size_t)Hh * 16 * K, 0.f), bp((size_t)Hh * 16, 0.f);
      for (int h = 0; h < Hh; ++h)
        for (int c = 0; c < dv; ++c) {
          memcpy(&wp[((size_t)h * 16 + c) * K], &w->data[((size_t)h * dv + c) * K], (size_t)K * 4);
          bp[h * 16 + c] = bv->data[h * dv + c];
        }
      // column dv of every head is the constant 1 (zero weights, unit bias): P.V then carries the softmax
      // row sum in that column and the SelfAttention kernel needs no separate accumulation of it
      for (int h = 0; h < Hh; ++h) bp[h * 16 + dv] = 1.0f;
      ZV_TRY(pack_linear_host(e, wp.data(), bp.data(), Hh * 16, K, EPI_BIAS_BF16, &L->sa_in[i], 0, nullptr));
    }
    const std::string c = p + ".conv_module" + std::to_string(i + 1);
    ZV_TRY(fold_self_attn_into_conv(e, a + ".out_proj", c, &L->conv_in[i], &L->conv_out[i]));
    ZV_TRY(upload_named(e, c + ".depthwise_conv.weight", &L->dw_w[i]));
    ZV_TRY(upload_named(e, c + ".depthwise_conv.bias", &L->dw_b[i]));
    const HostTensor* dw = find(e, c + ".depthwise_conv.weight");
    if ((int)dw->shape[2] != kernel) {
      e.err = "depthwise kernel size mismatch at " + c;
      return 942;
    }
  }
  ZV_TRY(upload_named(e, p + ".norm.bias", &L->norm_bias));
  const HostTensor* ls = find(e, p + ".norm.log_scale");
  if (!ls) return 940;
  L->log_scale = ls->data[0];
  ZV_TRY(upload_named(e, p + ".bypass.bypass_scale", &L->bypass_scale));
  ZV_TRY(upload_named(e, p + ".bypass_mid.bypass_scale", &L->bypass_mid_scale));
  (void)net;
  return 0;
}

static int build_net(Engine& e, Net& net, const std::string& prefix, int in_dim, int out_dim, int D, int F, int H,
                     int n_stacks, const int* ds, const int* nl, const int* kern, bool use_time, bool use_guidance,
                     bool two_stream = false) {
  const zv_config& c = e.cfg;
  net.prefix = prefix, net.in_dim = in_dim, net.out_dim = out_dim, net.D = D, net.F = F, net.H = H;
  net.hidden_na = 3 * D / 4;
  net.in_pad = (in_dim + 63) / 64 * 64;
  net.use_time = use_time, net.use_guidance = use_guidance;
  if (c.query_head_dim != 32 || c.pos_head_dim != 4 || c.value_head_dim != 12 || H * c.pos_head_dim > 16) {
    e.err = "attention kernel is specialised for query_head_dim 32, pos_head_dim 4, value_head_dim 12, <=4 heads";
    return 943;
  }
  if (D % 64 || D > 512) {
    e.err = "encoder dim must be a multiple of 64 and <= 512";
    return 943;
  }
  // TTSZipformerTwoStream keeps two projection pairs in ModuleLists; inference of the stereo dialog model
  // goes through the first one (zipformer_two_stream.py:240-263, x.size(2) == in_dim[0])
  const std::string pj = two_stream ? ".0" : "";
  ZV_TRY(pack_linear(e, prefix + ".in_proj" + pj, EPI_RES24, &net.in_proj, net.in_pad));
  ZV_TRY(pack_linear(e, prefix + ".out_proj" + pj, EPI_F32, &net.out_proj));
  net.in_pad_max = net.in_pad, net.out_dim_max = net.out_dim;
  net.two_stream = two_stream;
  if (two_stream) {
    // the mono pair: in_proj.1 (3 feat_dim -> D), out_proj.1 (D -> feat_dim)
    net.alt_in_dim = 3 * c.feat_dim, net.alt_out_dim = c.feat_dim;
    net.alt_in_pad = (net.alt_in_dim + 63) / 64 * 64;
    ZV_TRY(pack_linear(e, prefix + ".in_proj.1", EPI_RES24, &net.alt_in_proj, net.alt_in_pad));
    ZV_TRY(pack_linear(e, prefix + ".out_proj.1", EPI_F32, &net.alt_out_proj));
    if (net.alt_in_pad > net.in_pad_max) net.in_pad_max = net.alt_in_pad;
    if (net.alt_out_dim > net.out_dim_max) net.out_dim_max = net.alt_out_dim;
  }
  const int td = c.time_embed_dim;
  if (use_time) {
    ZV_TRY(upload_named(e, prefix + ".time_embed.0.weight", &net.te_w0));
    ZV_TRY(upload_named(e, prefix + ".time_embed.0.bias", &net.te_b0));
    ZV_TRY(upload_named(e, prefix + ".time_embed.2.weight", &net.te_w2));
    ZV_TRY(upload_named(e, prefix + ".time_embed.2.bias", &net.te_b2));
    if (use_guidance) ZV_TRY(upload_named(e, prefix + ".guidance_scale_embed.weight", &net.te_wg));
  }
  net.stacks.resize(n_stacks);
  size_t te_cols = 0;
  for (int i = 0; i < n_stacks; ++i) {
    Stack& st = net.stacks[i];
    st.ds = ds[i], st.kernel = kern[i];
    const std::string sp = prefix + ".encoders." + std::to_string(i);
    const std::string ep = st.ds == 1 ? sp : sp + ".encoder";
    if (st.ds != 1) {
      if (st.ds > 4) {
        e.err = "downsampling factor > 4 unsupported";
        return 943;
      }
      const HostTensor* b = find(e, sp + ".downsample.bias");
      if (!b) return 940;
      float mx = -1e30f, sum = 0.f;
      for (int k = 0; k < st.ds; ++k) mx = fmaxf(mx, b->data[k]);
      for (int k = 0; k < st.ds; ++k) sum += expf(b->data[k] - mx);
      for (int k = 0; k < st.ds; ++k) st.ds_w[k] = expf(b->data[k] - mx) / sum;
      ZV_TRY(upload_named(e, sp + ".out_combiner.bypass_scale", &st.out_comb_scale));
    }
    st.layers.resize(nl[i]);
    if (use_time) st.te_off = te_cols, te_cols += D;
    for (int j = 0; j < nl[i]; ++j) {
      ZV_TRY(build_layer(e, net, ep + ".layers." + std::to_string(j), st.kernel, &st.layers[j]));
      if (use_time) {
        st.layers[j].te_off_ff1 = te_cols, te_cols += st.layers[j].ff_in[0].N;
        for (int i2 = 0; i2 < 2; ++i2) st.layers[j].te_off_conv[i2] = te_cols, te_cols += st.layers[j].conv_in[i2].N;
      }
    }
  }
  net.N_te = (int)te_cols;
  if (use_time) {
    // Fold: te_stack = Ws a + bs (zipformer.py:677-680), FF1 row-bias = Wff1 te_stack (in_proj is linear,
    // so in_proj(src + te) = in_proj(src) + Wff1 te).  Rows are stored in projection-column order.
    ZV_TRY(dev_alloc(e, e.weight_allocs, (void**)&net.te_proj_w, te_cols * td * sizeof(float)));
    ZV_TRY(dev_alloc(e, e.weight_allocs, (void**)&net.te_proj_b, te_cols * sizeof(float)));
    for (int i = 0; i < n_stacks; ++i) {
      Stack& st = net.stacks[i];
      const std::string sp = prefix + ".encoders." + std::to_string(i);
      const std::string ep = st.ds == 1 ? sp : sp + ".encoder";
      const HostTensor* ws = find(e, ep + ".time_emb.1.weight");
      const HostTensor* bs = find(e, ep + ".time_emb.1.bias");
      if (!ws || !bs) return 940;
      ZV_TRY((int)cudaMemcpy(net.te_proj_w + st.te_off * td, ws->data.data(), (size_t)D * td * 4,
                             cudaMemcpyHostToDevice));
      ZV_TRY((int)cudaMemcpy(net.te_proj_b + st.te_off, bs->data.data(), (size_t)D * 4, cudaMemcpyHostToDevice));
      for (int j = 0; j < nl[i]; ++j) {
        const std::string f = ep + ".layers." + std::to_string(j) + ".feed_forward1.in_proj.weight";
        const HostTensor* wf = find(e, f);
        if (!wf) return 940;
        const int F1 = (int)wf->shape[0];
        float* wf_dev = nullptr;
        ZV_TRY((int)cudaMalloc(&wf_dev, (size_t)F1 * D * 4));
        cudaMemcpy(wf_dev, wf->data.data(), (size_t)F1 * D * 4, cudaMemcpyHostToDevice);
        const size_t off = st.layers[j].te_off_ff1;
        int rc = sgemm_nn(wf_dev, net.te_proj_w + st.te_off * td, net.te_proj_w + off * td, F1, td, D, 0);
        if (!rc) rc = sgemm_nn(wf_dev, net.te_proj_b + st.te_off, net.te_proj_b + off, F1, 1, D, 0);
        cudaDeviceSynchronize();
        cudaFree(wf_dev);
        ZV_TRY(rc);
        // conv in_proj sees x + te as well (the time embedding added after each SelfAttention rides in the
        // folded GEMMs): row bias = Wci te_stack, rows in accumulator (GLU-permuted) order
        for (int i2 = 0; i2 < 2; ++i2) {
          const HostTensor* wc = find(e, ep + ".layers." + std::to_string(j) + ".conv_module" + std::to_string(i2 + 1) +
                                             ".in_proj.weight");
          if (!wc) return 940;
          const int N2 = (int)wc->shape[0];
          const std::vector<int> perm = row_perm(EPI_GLU_BF16, N2);
          std::vector<float> wperm((size_t)N2 * D);
          for (int n = 0; n < N2; ++n) memcpy(&wperm[(size_t)n * D], &wc->data[(size_t)perm[n] * D], (size_t)D * 4);
          float* wc_dev = nullptr;
          ZV_TRY((int)cudaMalloc(&wc_dev, (size_t)N2 * D * 4));
          cudaMemcpy(wc_dev, wperm.data(), (size_t)N2 * D * 4, cudaMemcpyHostToDevice);
          const size_t offc = st.layers[j].te_off_conv[i2];
          rc = sgemm_nn(wc_dev, net.te_proj_w + st.te_off * td, net.te_proj_w + offc * td, N2, td, D, 0);
          if (!rc) rc = sgemm_nn(wc_dev, net.te_proj_b + st.te_off, net.te_proj_b + offc, N2, 1, D, 0);
          cudaDeviceSynchronize();
          cudaFree(wc_dev);
          ZV_TRY(rc);
        }
      }
    }
  }
  return 0;
}

int engine_finalize(Engine& e) {
  const zv_config& c = e.cfg;
  if (e.raw.find("fm_decoder.in_proj.weight") == e.raw.end()) {
    // audio-only handle (mel front-end / vocoder): no acoustic model to pack
    bool any_model = false;
    for (auto& kv : e.raw) any_model = any_model || kv.first.rfind("vocoder.", 0) != 0;
    if (!any_model) {
      e.finalized = false;
      return 0;
    }
  }
  const int fx = c.stereo ? 2 * c.feat_dim : c.feat_dim;
  ZV_TRY(build_net(e, e.fm, "fm_decoder", 2 * fx + c.feat_dim, fx, c.fm_dim, c.fm_feedforward_dim, c.fm_num_heads,
                   c.fm_num_stacks, c.fm_downsampling_factor, c.fm_num_layers, c.fm_cnn_module_kernel, true,
                   c.use_guidance_scale_embed != 0, c.stereo != 0));
  const int one = 1;
  ZV_TRY(build_net(e, e.text, "text_encoder", c.text_embed_dim, c.feat_dim, c.text_dim, c.text_feedforward_dim,
                   c.text_num_heads, 1, &one, &c.text_num_layers, &c.text_cnn_module_kernel, false, false));
  ZV_TRY(upload_named(e, "embed.weight", &e.embed));
  if (c.dialog) ZV_TRY(upload_named(e, "spk_embed.weight", &e.spk_embed));
  ZV_TRY((int)cudaDeviceSynchronize());
  e.finalized = true;
  // keep only what zv_audio needs; model tensors are now on the device
  for (auto it = e.raw.begin(); it != e.raw.end();)
    it = (it->first.rfind("vocoder.", 0) == 0) ? ++it : e.raw.erase(it);
  return 0;
}

// ------------------------------------------------------------------ workspace
int ws_reserve(Engine& e, Net& net, Workspace& ws, int nb, int T) {
  size_t rows = (size_t)nb * T;
  if (rows <= ws.cap_rows && nb <= ws.cap_nb) return 0;
  // capacity only ever grows in BOTH dimensions: alternating shapes (many short utterances, then few long ones)
  // must not shrink one of them and reallocate ~3 GB (and drop every captured graph) on each call
  if (rows < ws.cap_rows) rows = ws.cap_rows;
  if (nb < ws.cap_nb) nb = ws.cap_nb;
  cudaDeviceSynchronize();
  for (void* p : ws.allocs) cudaFree(p);
  ws.allocs.clear();
  ws.cap_rows = rows, ws.cap_nb = nb, ws.D = net.D;
  ++ws.gen;
  const size_t D = net.D;
  const size_t Fmax = (size_t)net.F * 5 / 4;
  const size_t td = e.cfg.time_embed_dim;
  auto A = [&](void** p, size_t bytes) { return dev_alloc(e, ws.allocs, p, bytes); };
  ZV_TRY(A((void**)&ws.X_hi, rows * D * 2));
  ZV_TRY(A((void**)&ws.Xs_hi, rows * D * 2));
  ZV_TRY(A((void**)&ws.Y_hi, rows * D * 2));
  ZV_TRY(A((void**)&ws.X_lo, rows * D));
  ZV_TRY(A((void**)&ws.Xs_lo, rows * D));
  ZV_TRY(A((void**)&ws.Y_lo, rows * D));
  ZV_TRY(A((void**)&ws.vel, rows * net.out_dim_max * 4));
  ZV_TRY(A((void**)&ws.xcur, rows * net.out_dim_max * 4));
  ZV_TRY(A((void**)&ws.x_in, rows * net.in_pad_max * 2));
  ZV_TRY(A((void**)&ws.qkp, rows * 288 * 2));
  ZV_TRY(A((void**)&ws.hidden, rows * Fmax * 2));
  ZV_TRY(A((void**)&ws.na_v, rows * net.hidden_na * 2));
  ZV_TRY(A((void**)&ws.na_y, rows * net.hidden_na * 2));
  ZV_TRY(A((void**)&ws.na_o, rows * net.hidden_na * 2));
  ZV_TRY(A((void**)&ws.sa_v, rows * 64 * 2));
  ZV_TRY(A((void**)&ws.conv_a, rows * D * 2));
  // conv hidden state (D columns) + SelfAttention output (48 of the 64 extra columns; the rest stays zero)
  ZV_TRY(A((void**)&ws.conv_b, rows * (D + 64) * 2));
  ZV_TRY((int)cudaMemset(ws.conv_b, 0, rows * (D + 64) * 2));
  // key split of the SelfAttention kernel: only small grids use it, so the buffer is capped (8 splits of every row
  // of a small batch fit; the launcher falls back to fewer splits or none when they would not)
  ws.sa_part_cap = std::min<size_t>((size_t)64 << 20, rows * net.H * 8 * 17 * sizeof(float));
  ZV_TRY(A((void**)&ws.sa_part, ws.sa_part_cap));
  for (int i = 0; i < 3; ++i) ZV_TRY(A((void**)&ws.mask[i], rows));
  ZV_TRY(A((void**)&ws.t_dev, nb * 4));
  ZV_TRY(A((void**)&ws.ts_dev, 1025 * 4));
  ZV_TRY(A((void**)&ws.g_dev, nb * 4));
  ZV_TRY(A((void**)&ws.guid_dev, nb * 4));
  ZV_TRY(A((void**)&ws.te_a, nb * td * 4));
  ZV_TRY(A((void**)&ws.te_out, (size_t)nb * (net.N_te + 1) * 4));
  return 0;
}

int run_gemm(Engine& e, const Lin& lin, const __nv_bfloat16* A, int lda, int M, EpiParams ep, cudaStream_t s,
             const __nv_bfloat16* A2, int lda2, int k_split) {
  GemmPlan p;
  p.M = M, p.N_pad = lin.N_pad, p.K = lin.K, p.lda = lda, p.ldw = lin.ldw, p.block_n = lin.block_n;
  p.A2 = A2, p.lda2 = lda2, p.k_split = k_split;
  ep.mode = lin.mode, ep.M = M, ep.N = lin.N, ep.gate_cols = lin.gate_cols, ep.bias = lin.bias;
  // the s24 residual GEMMs are timed with the residual class; without a residual input (stack in_proj) with EPI_F32's
  const int cls = lin.mode == EPI_RES24 ? (ep.res_hi ? KC_GEMM_RESIDUAL : KC_GEMM_F32) : KC_GEMM_BIAS + lin.mode;
  ZV_K(cls, 2.0 * M * lin.N * (double)lin.K_valid, s, gemm_launch(A, lin.w, p, ep, e.gemm_impl, s));
  return 0;
}

// Algorithmic MACs per (query, key) pair and layer: the scores of the H heads are needed once
// (H * (dq + dp) = 144), each consumer adds its A.V product.  The three passes recompute the
// scores; only `score_heads` of them are counted as algorithmic work for this launch.
static int run_attn(Engine& e, int cls, const AttnParams& a, int score_heads, cudaStream_t s) {
  const double macs = (double)a.nb * a.L * a.L * (score_heads * (32.0 + 4.0) + (double)a.heads * a.dv);
  ZV_K(cls, 2.0 * macs, s, attention(a, s));
  return 0;
}
// algorithmic bytes: bf16 in + bf16 out
static int run_dwconv(Engine& e, const __nv_bfloat16* in, const float* w, const float* b, int nb, int L, int C, int k,
                      __nv_bfloat16* out, int ld_out, cudaStream_t s) {
  ZV_K(KC_DWCONV, 4.0 * nb * L * C, s, dwconv_swooshr(in, w, b, nb, L, C, k, out, ld_out, s));
  return 0;
}

static int get_pos_table(Engine& e, Net& net, Layer& ly, int L, cudaStream_t s, float** out, void** out16) {
  auto it = ly.pos_tab.find(L);
  if (it == ly.pos_tab.end()) {
    if (ly.pos_tab.size() >= kMaxPosTables) {
      // single-utterance serving sees a new T on almost every request: drop this layer's tables instead of growing
      // without bound.  Kernels in flight and captured graphs may still read them.
      cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
      cudaStreamIsCapturing(s, &cap);
      if (cap != cudaStreamCaptureStatusNone) {
        e.err = "position-table cache full during stream capture";
        return 951;  // the capture fails and zv_sample runs the plain loop, which evicts
      }
      ZV_TRY((int)cudaDeviceSynchronize());
      for (auto& kv : ly.pos_tab) cudaFree(kv.second);
      for (auto& kv : ly.pos_tab16) cudaFree(kv.second);
      ly.pos_tab.clear(), ly.pos_tab16.clear();
      ++e.ws_fm.gen;  // captured graphs hold the freed addresses
    }
    float* E = nullptr;
    const cudaError_t mrc = cudaMalloc((void**)&E, (size_t)net.H * (2 * L - 1) * 4 * sizeof(float));
    if (mrc != cudaSuccess) {
      e.err = std::string("cudaMalloc failed: ") + cudaGetErrorString(mrc);
      return (int)mrc;
    }
    ZV_K(KC_TIME_EMBED, (double)net.H * (2 * L - 1) * 16, s,
         pos_table(ly.w_pos, net.H, e.cfg.pos_head_dim, e.cfg.pos_dim, L, E, s));
    it = ly.pos_tab.emplace(L, E).first;
    void* E16 = nullptr;
    const cudaError_t hrc = cudaMalloc(&E16, (size_t)2 * net.H * pos_table_f16_stride(L) * 8);
    if (hrc != cudaSuccess) {
      e.err = std::string("cudaMalloc failed: ") + cudaGetErrorString(hrc);
      return (int)hrc;
    }
    ly.pos_tab16.emplace(L, E16);
    ZV_K(KC_TIME_EMBED, (double)net.H * (2 * L - 1) * 32, s, pos_table_f16(E, net.H, L, E16, s));
  }
  *out = it->second;
  *out16 = ly.pos_tab16[L];
  return 0;
}

void free_pos_tables(Net& net) {
  for (Stack& st : net.stacks)
    for (Layer& ly : st.layers) {
      for (auto& kv : ly.pos_tab) cudaFree(kv.second);
      for (auto& kv : ly.pos_tab16) cudaFree(kv.second);
      ly.pos_tab.clear(), ly.pos_tab16.clear();
    }
}

// Zipformer2EncoderLayer.forward in eval mode (zipformer.py:489-642).  X: layer input / output, ws.Y: working
// residual stream, both in the s24 storage format (hi plane = bf16 GEMM A operand, lo plane = 8 more mantissa bits).
static int layer_forward(Engine& e, Net& net, Workspace& ws, Layer& ly, int kernel, __nv_bfloat16* X_hi,
                         uint8_t* X_lo, int nb, int L,
                         const uint8_t* mask, const float* te_stack, const float* te_ff1, int te_ld, cudaStream_t s) {
  const int M = nb * L, D = net.D, hn = net.hidden_na;
  float* Etab = nullptr;
  void* Etab16 = nullptr;
  ZV_TRY(get_pos_table(e, net, ly, L, s, &Etab, &Etab16));
  EpiParams ep;
  // attention projections from the layer input (before the time embedding is added) -- :526-531
  // ZV_MERGE_QKP=0: separate launches (A/B measurements)
  static const bool merge_qkp = !(getenv("ZV_MERGE_QKP") && atoi(getenv("ZV_MERGE_QKP")) == 0);
  const bool merged = merge_qkp && ly.ff1_qkp.w != nullptr;
  if (!merged) {
    ep = EpiParams();
    ep.L = L, ep.out_bf16 = ws.qkp, ep.ld_bf16 = 288, ep.out_f16 = 1;  // q | k | p kept in fp16
    ZV_TRY(run_gemm(e, ly.attn_in, X_hi, D, M, ep, s));
  }

  // src: the stream state the module reads (X for the first feed-forward, ws.Y afterwards); every module writes ws.Y
  // from_x: the residual input is the layer input X (first module), else ws.Y updated in place
  auto residual = [&](const Lin& lin, const __nv_bfloat16* A, int lda, bool from_x, const float* resbias,
                      bool bypass_x, const float* bscale) {
    EpiParams r;
    r.L = L, r.res_hi = from_x ? X_hi : ws.Y_hi, r.res_lo = from_x ? X_lo : ws.Y_lo;
    r.out_bf16 = ws.Y_hi, r.out_lo = ws.Y_lo, r.ld_bf16 = D;
    r.resbias = resbias, r.bias_ld = te_ld, r.bypass_scale = bscale;
    if (bypass_x) r.orig_hi = X_hi, r.orig_lo = X_lo;
    return run_gemm(e, lin, A, lda, M, r, s);
  };
  auto ff = [&](int i, const float* rowbias, bool from_x, const float* resbias, bool bypass_x, const float* bscale) {
    EpiParams f;
    f.L = L, f.rowbias = rowbias, f.bias_ld = te_ld, f.out_bf16 = ws.hidden, f.ld_bf16 = ly.ff_in[i].N;
    int rc;
    if (i == 0 && merged) {  // + q | k | p (fp16, pitch 288) from the same rows, before the time embedding is added:
      f.out2_bf16 = ws.qkp, f.ld2_bf16 = 288;  // the row bias (W_ff1 . time_emb) applies to the first part only
      rc = run_gemm(e, ly.ff1_qkp, X_hi, D, M, f, s);
    } else {
      rc = run_gemm(e, ly.ff_in[i], from_x ? X_hi : ws.Y_hi, D, M, f, s);
    }
    if (rc) return rc;
    return residual(ly.ff_out[i], ws.hidden, ly.ff_in[i].N, from_x, resbias, bypass_x, bscale);
  };
  const int ldc = D + 64;  // conv_b row pitch: [conv hidden (D) | SelfAttention output (48) | zeros (16)]
  auto self_attn = [&](int i) {
    EpiParams v;
    v.L = L, v.out_bf16 = ws.sa_v, v.ld_bf16 = 16 * net.H;
    int rc = run_gemm(e, ly.sa_in[i], ws.Y_hi, D, M, v, s);
    if (rc) return rc;
    AttnParams a;
    a.qkp = ws.qkp, a.ldq = 288, a.E = Etab, a.mask = mask, a.V = ws.sa_v, a.ldv = 16 * net.H, a.vhs = 16, a.gate = nullptr, a.ldg = 0;
    a.out = ws.conv_b + D, a.ldo = ldc, a.nb = nb, a.L = L, a.H = net.H, a.heads = net.H, a.dv = 12;
    a.E16 = Etab16, a.e16_stride = pos_table_f16_stride(L);
    a.part = ws.sa_part, a.part_cap = ws.sa_part_cap;
    return run_attn(e, KC_ATTN_SELF, a, i == 0 ? net.H - 1 : 0, s);  // out_proj: folded into conv(i)
  };
  // SelfAttention(i).out_proj + time embedding + ConvolutionModule(i) with one pass over the fp32 stream
  auto conv = [&](int i, const float* te_conv, const float* resbias) {
    EpiParams c;
    c.L = L, c.mask = mask, c.out_bf16 = ws.conv_a, c.ld_bf16 = D, c.rowbias = te_conv, c.bias_ld = te_ld;
    int rc = run_gemm(e, ly.conv_in[i], ws.Y_hi, D, M, c, s, ws.conv_b + D, ldc, D);
    if (rc) return rc;
    rc = run_dwconv(e, ws.conv_a, ly.dw_w[i], ly.dw_b[i], nb, L, D, kernel, ws.conv_b, ldc, s);
    if (rc) return rc;
    return residual(ly.conv_out[i], ws.conv_b, ldc, false, resbias, false, nullptr);
  };

  // src += time_emb; src += FF1(src)  (:532-536): FF1 input bias folded, time_emb added to the stream here
  ZV_TRY(ff(0, te_ff1, true, te_stack, false, nullptr));
  // src += NonlinAttention(src, A[0:1])  (:542,558-562)
  ep = EpiParams();
  ep.L = L, ep.out_bf16 = ws.na_v, ep.ld_bf16 = hn, ep.out2_bf16 = ws.na_y, ep.ld2_bf16 = hn;
  ZV_TRY(run_gemm(e, ly.na_in, ws.Y_hi, D, M, ep, s));
  {
    AttnParams a;
    a.qkp = ws.qkp, a.ldq = 288, a.E = Etab, a.mask = mask, a.V = ws.na_v, a.ldv = hn, a.gate = ws.na_y, a.ldg = hn;
    a.out = ws.na_o, a.ldo = hn, a.nb = nb, a.L = L, a.H = net.H, a.heads = 1, a.dv = hn;
    ZV_TRY(run_attn(e, KC_ATTN_NONLIN, a, 1, s));
  }
  ZV_TRY(residual(ly.na_out, ws.na_o, hn, false, nullptr, false, nullptr));
  // src += SelfAttn1(src, A); src += time_emb  (:564-579)
  const float* te_c0 = te_stack ? ws.te_out + ly.te_off_conv[0] : nullptr;
  const float* te_c1 = te_stack ? ws.te_out + ly.te_off_conv[1] : nullptr;
  ZV_TRY(self_attn(0));
  ZV_TRY(conv(0, te_c0, te_stack));                                      // :581-587
  ZV_TRY(ff(1, nullptr, false, nullptr, true, ly.bypass_mid_scale));      // :593-598 (FF2 + bypass_mid)
  ZV_TRY(self_attn(1));                                                  // :600-616
  ZV_TRY(conv(1, te_c1, te_stack));                                      // :618-624
  ZV_TRY(ff(2, nullptr, false, nullptr, false, nullptr));                 // :630-632
  // src = bypass(src_orig, BiasNorm(src))  (:635-637), in place on X
  ZV_K(KC_BIASNORM, (double)M * D * 9, s,
       biasnorm_bypass(ws.Y_hi, ws.Y_lo, X_hi, X_lo, ly.norm_bias, ly.log_scale, ly.bypass_scale, M, D, X_hi, X_lo, s));
  return 0;
}

int net_forward(Engine& e, Net& net, Workspace& ws, int nb, int T, const uint8_t* mask_dev, const TimeArgs& ta,
                float* out, cudaStream_t s) {
  const int M = nb * T, D = net.D;
  const int td = e.cfg.time_embed_dim;
  // One time step (and guidance scale) for the whole batch -- the Euler solver's case -- needs ONE row of
  // time-embedding projections: the per-batch row biases are then read with pitch 0.
  const bool te_shared = ta.t_stride == 0 && (!ta.g || ta.g_stride == 0);
  const int nb_te = te_shared ? 1 : nb;
  const int te_ld = te_shared ? 0 : net.N_te;
  EpiParams ep;
  ep.L = T, ep.out_bf16 = ws.X_hi, ep.out_lo = ws.X_lo, ep.ld_bf16 = D;  // no residual input: X = in_proj(x)
  ZV_TRY(run_gemm(e, net.in_proj, ws.x_in, net.in_pad, M, ep, s));
  if (net.use_time) {
    if (!ta.t) {
      e.err = "this network needs a time step";
      return 950;
    }
    ZV_K(KC_TIME_EMBED, 4.0 * (4 * td * td + td), s,
         time_embed(ta.t, ta.t_stride, ta.g, ta.g_stride, nb_te, td, net.te_w0, net.te_b0, net.te_w2, net.te_b2,
                    net.te_wg, ws.te_a, s));
    ZV_K(KC_TIME_EMBED, 4.0 * ((double)net.N_te * td + (double)nb_te * (net.N_te + td)), s,
         small_linear(ws.te_a, net.te_proj_w, net.te_proj_b, nb_te, net.N_te, td, ws.te_out, s));
  }
  for (Stack& st : net.stacks) {
    const int ds = st.ds;
    const int L = (T + ds - 1) / ds;
    __nv_bfloat16* X_hi = ws.X_hi;
    uint8_t* X_lo = ws.X_lo;
    const uint8_t* mask = mask_dev;
    if (ds != 1) {
      uint8_t* msub = mask_dev ? ws.mask[ds == 2 ? 1 : 2] : nullptr;
      g_launch_count += msub ? 1 : 0;  // mask subsampling rides in the same launcher
      ZV_K(KC_RESAMPLE, (double)nb * D * (3.0 * T + 3.0 * L), s,
           downsample(ws.X_hi, ws.X_lo, st.ds_w, nb, T, D, ds, ws.Xs_hi, ws.Xs_lo, msub, mask_dev, s));
      X_hi = ws.Xs_hi, X_lo = ws.Xs_lo, mask = msub;
    }
    const float* te_stack = net.use_time ? ws.te_out + st.te_off : nullptr;
    for (Layer& ly : st.layers) {
      const float* te_ff1 = net.use_time ? ws.te_out + ly.te_off_ff1 : nullptr;
      // row pitch of the time-embedding projections is N_te for every layer of this net
      ZV_TRY(layer_forward(e, net, ws, ly, st.kernel, X_hi, X_lo, nb, L, mask, te_stack, te_ff1, te_ld, s));
    }
    if (ds != 1)
      ZV_K(KC_RESAMPLE, (double)nb * D * (6.0 * T + 3.0 * L), s,
           upsample_combine(ws.X_hi, ws.X_lo, ws.Xs_hi, ws.Xs_lo, st.out_comb_scale, nb, T, L, D, ds, s));
  }
  ep = EpiParams();
  ep.L = T, ep.out_f32 = out, ep.ld_f32 = net.out_dim;
  ZV_TRY(run_gemm(e, net.out_proj, ws.X_hi, D, M, ep, s));
  return 0;
}

}  // namespace zv
