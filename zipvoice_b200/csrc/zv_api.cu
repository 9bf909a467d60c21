// extern "C" surface of libzipvoice_b200.so (declared in include/zipvoice_b200.h).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <utility>

#include "zv_common.cuh"
#include "zv_engine.h"

using namespace zv;
namespace zv {
extern int g_attn_impl;
}

static thread_local std::string g_err;
static const char* kNullHandle = "null handle";

#define API_K(cls, work, s, expr)                 \
  do {                                             \
    ++g_launch_count;                              \
    if (e->prof.on) e->prof.begin(cls, work, s);   \
    expr;                                          \
    if (e->prof.on) e->prof.end(s);                \
    API_TRY((int)cudaGetLastError());              \
  } while (0)

#define API_TRY(expr)                                                                     \
  do {                                                                                    \
    int _rc = (expr);                                                                     \
    if (_rc) {                                                                            \
      if (e->err.empty()) {                                                               \
        char _b[512];                                                                     \
        snprintf(_b, sizeof(_b), "%s failed with %d (%s)", #expr, _rc,                    \
                 _rc < 900 ? cudaGetErrorString((cudaError_t)_rc) : "zipvoice_b200 error"); \
        e->err = _b;                                                                      \
      }                                                                                   \
      return _rc;                                                                         \
    }                                                                                     \
  } while (0)

// ------------------------------------------------------------------ small device kernels of the text path
// Embedding lookup -> bf16 rows [B*S][in_pad] (zipvoice.py:203); also the key padding mask.
__global__ void embed_kernel(const int64_t* __restrict__ tokens, const int64_t* __restrict__ lens,
                             const float* __restrict__ table, int S, int dim, int in_pad, int vocab,
                             __nv_bfloat16* __restrict__ out, uint8_t* __restrict__ mask, size_t total) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int c = (int)(gid % in_pad);
  const size_t row = gid / in_pad;
  long long tok = tokens[row];
  tok = tok < 0 ? 0 : (tok >= vocab ? vocab - 1 : tok);
  out[gid] = __float2bfloat16(c < dim ? table[(size_t)tok * dim + c] : 0.f);
  if (c == 0) mask[row] = (long long)(row % S) >= lens[row / S] ? 1 : 0;
}
// Dialog speaker rows (zipvoice_dialog.py:118-125,149-158): one thread per sequence scans the turn count.
__global__ void spk_add_kernel(const int64_t* __restrict__ tokens, int B, int S, int F, int pad_id, int spk_a,
                               int spk_b, const float* __restrict__ spk, float* __restrict__ embed) {
  const int b = blockIdx.x;
  int count = 0;
  for (int s0 = 0; s0 < S; ++s0) {
    const long long tok = tokens[(size_t)b * S + s0];
    if (tok == spk_a || tok == spk_b) ++count;
    if (tok == pad_id) continue;
    const float* sv = spk + (size_t)(count & 1) * F;
    for (int c = threadIdx.x; c < F; c += blockDim.x) embed[((size_t)b * S + s0) * F + c] += sv[c];
  }
}
// text_condition[b][t] = embed[b][idx], idx = t / floor(T_b / ntok_b) for the first ntok_b*d frames, else ntok_b
// (utils/common.py:246-252,271-295); padding_mask[b][t] = t >= T_b.
__global__ void text_condition_kernel(const float* __restrict__ embed, const int64_t* __restrict__ tokens_lens,
                                      const int64_t* __restrict__ features_lens, int S, int T, int F,
                                      float* __restrict__ out, uint8_t* __restrict__ mask, size_t total) {
  const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int c = (int)(gid % F);
  const size_t row = gid / F;
  const int t = (int)(row % T), b = (int)(row / T);
  const long long n = tokens_lens[b], fl = features_lens[b];
  const long long d = n > 0 ? fl / n : 0;
  long long idx = n;
  if (d > 0 && t < n * d) idx = t / d;
  if (idx > S - 1) idx = S - 1;
  out[gid] = embed[((size_t)b * S + idx) * F + c];
  if (c == 0 && mask) mask[row] = t >= fl ? 1 : 0;
}

extern "C" {

int zv_create(const zv_config* cfg, void** handle) {
  if (!cfg || !handle) return 960;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    g_err = "zipvoice_b200 needs a CUDA device (sm_100a); there is no CPU fallback";
    return 961;
  }
  Engine* e = new Engine();
  e->cfg = *cfg;
  *handle = e;
  return 0;
}

void zv_destroy(void* handle) {
  Engine* e = (Engine*)handle;
  if (!e) return;
  cudaDeviceSynchronize();
  e->prof.release();
  for (auto& g : e->graphs)
    if (g.exec) cudaGraphExecDestroy(g.exec);
  for (void* p : e->gallocs) cudaFree(p);
  if (e->gstream) cudaStreamDestroy(e->gstream);
  if (e->gev_in) cudaEventDestroy(e->gev_in);
  if (e->gev_out) cudaEventDestroy(e->gev_out);
  audio_destroy(*e);
  if (e->pin_host) {
    cudaFreeHost(e->pin_host);
    for (int i = 0; i < Engine::kPinSlots; ++i) cudaEventDestroy(e->pin_ev[i]);
  }
  free_pos_tables(e->fm);
  free_pos_tables(e->text);
  for (void* p : e->weight_allocs) cudaFree(p);
  for (void* p : e->ws_fm.allocs) cudaFree(p);
  for (void* p : e->ws_text.allocs) cudaFree(p);
  delete e;
}

const char* zv_last_error(void* handle) {
  Engine* e = (Engine*)handle;
  return e ? e->err.c_str() : g_err.c_str();
}

int zv_set_weight(void* handle, const char* name, const float* data, const int64_t* shape, int ndim, int on_device) {
  Engine* e = (Engine*)handle;
  if (!e) return 960;
  HostTensor t;
  size_t n = 1;
  for (int i = 0; i < ndim; ++i) t.shape.push_back(shape[i]), n *= (size_t)shape[i];
  t.data.resize(n);
  if (on_device) API_TRY((int)cudaMemcpy(t.data.data(), data, n * 4, cudaMemcpyDeviceToHost));
  else memcpy(t.data.data(), data, n * 4);
  e->raw[name] = std::move(t);
  return 0;
}

int zv_finalize_weights(void* handle) {
  Engine* e = (Engine*)handle;
  if (!e) return 960;
  e->err.clear();
  return engine_finalize(*e);
}

int zv_set_option(void* handle, const char* key, int value) {
  Engine* e = (Engine*)handle;
  if (!e) return 960;
  if (!strcmp(key, "gemm_impl")) e->gemm_impl = value;
  else if (!strcmp(key, "attn_impl")) zv::g_attn_impl = value;  // process-wide
  else if (!strcmp(key, "graph_max_rows")) e->graph_max_rows = value;
  else if (!strcmp(key, "stream")) {
    // TTSZipformerTwoStream: 0 = the two-channel projection pair, 1 = the mono pair (zipformer_two_stream.py:238-244)
    Net& net = e->fm;
    if (!net.two_stream || value < 0 || value > 1) {
      e->err = "option stream needs a two-stream (stereo) model and a value of 0 or 1";
      return 962;
    }
    if (value != net.stream) {
      std::swap(net.in_proj, net.alt_in_proj), std::swap(net.out_proj, net.alt_out_proj);
      std::swap(net.in_dim, net.alt_in_dim), std::swap(net.in_pad, net.alt_in_pad), std::swap(net.out_dim, net.alt_out_dim);
      net.stream = value;
    }
  }
  else {
    e->err = std::string("unknown option ") + key;
    return 962;
  }
  return 0;
}

int zv_time_steps(double t_start, double t_end, int num_step, double t_shift, float* out) {
  // torch.linspace on fp32 (ATen RangeFactories: step in fp32, filled from both ends with fused
  // multiply-add), then the tensor expression t_shift*t / (1 + (t_shift-1)*t) op by op in fp32 with
  // python-double scalars rounded to fp32 (solver.py:277-279).
  const int steps = num_step + 1;
  const float start = (float)t_start, end = (float)t_end;
  const float step = (end - start) / (float)(steps - 1);
  const int halfway = steps / 2;
  const float sh = (float)t_shift, sm1 = (float)(t_shift - 1.0);
  for (int i = 0; i < steps; ++i) {
    const float t = i < halfway ? fmaf(step, (float)i, start) : fmaf(-step, (float)(steps - i - 1), end);
    volatile float num = sh * t;
    volatile float prod = sm1 * t;
    volatile float den = 1.0f + prod;
    out[i] = num / den;
  }
  return 0;
}

int zv_prof_enable(void* handle, int on) {
  Engine* e = (Engine*)handle;
  if (!e) return 960;
  cudaDeviceSynchronize();
  e->prof.recs.clear(), e->prof.used = 0;
  e->prof.on = on != 0;
  return 0;
}
int zv_prof_num_classes(void) { return KC_COUNT; }
const char* zv_prof_class_name(int c) { return kclass_name(c); }
int zv_prof_class_is_tensor(int c) { return kclass_is_tensor(c) ? 1 : 0; }
int zv_prof_read(void* handle, double* ms, double* work, long long* launches) {
  Engine* e = (Engine*)handle;
  if (!e) return 960;
  return e->prof.read(ms, work, launches);
}

long long zv_launch_count(void) { return g_launch_count; }
void zv_launch_count_reset(void) { g_launch_count = 0; }

}  // extern "C"

static int upload_small(Engine* e, float* dst, const float* host, int n, cudaStream_t s) {
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(s, &cap);
  if (n > Engine::kPinSlotFloats || cap != cudaStreamCaptureStatusNone)  // oversized / capturing: plain staged copy
    return (int)cudaMemcpyAsync(dst, host, (size_t)n * 4, cudaMemcpyHostToDevice, s);
  if (!e->pin_host) {
    API_TRY((int)cudaHostAlloc((void**)&e->pin_host, (size_t)Engine::kPinSlots * Engine::kPinSlotFloats * 4,
                               cudaHostAllocDefault));
    for (int i = 0; i < Engine::kPinSlots; ++i)
      API_TRY((int)cudaEventCreateWithFlags(&e->pin_ev[i], cudaEventDisableTiming));
  }
  const int slot = e->pin_next;
  e->pin_next = (slot + 1) % Engine::kPinSlots;
  API_TRY((int)cudaEventSynchronize(e->pin_ev[slot]));  // never recorded: returns at once
  float* stage = e->pin_host + (size_t)slot * Engine::kPinSlotFloats;
  memcpy(stage, host, (size_t)n * 4);
  API_TRY((int)cudaMemcpyAsync(dst, stage, (size_t)n * 4, cudaMemcpyHostToDevice, s));
  return (int)cudaEventRecord(e->pin_ev[slot], s);
}

// One DiffusionModel.forward / DistillDiffusionModel.forward.  t_dev points at the device scalar of
// this step; guidance already lives in ws.guid_dev when needed.  If x_update != null the Euler step
// x += dt * v is applied instead of writing v (fused CFG combine + update).
static int velocity_step(Engine* e, int variant, float t_host, const float* t_dev, const float* x, const float* text,
                         const float* speech, const uint8_t* mask, int B, int T, bool all_zero_g, int guidance_len,
                         float* v_out, float* x_update, float dt, cudaStream_t s) {
  Net& net = e->fm;
  Workspace& ws = e->ws_fm;
  const int Ft = e->cfg.feat_dim, F = e->fm.out_dim;  // F: width of x / speech / velocity (active stream)
  const bool cfg = (variant == 0) && !all_zero_g;
  const int nb = cfg ? 2 * B : B;
  const bool hi = t_host > 0.5f;  // solver.py:90 decided on the host from the precomputed time table
  // bytes: x read once, text / speech once, bf16 rows written
  API_K(KC_PACK, (double)B * T * (2 * F + Ft) * 4.0 + (double)nb * T * net.in_pad * 2.0, s,
        API_TRY(pack_input(x, text, speech, mask, B, T, F, Ft, cfg ? 1 : 0, hi ? 1 : 0, ws.x_in, net.in_pad,
                           mask ? ws.mask[0] : nullptr, s)));
  TimeArgs ta;
  ta.t = t_dev, ta.t_stride = 0;
  if (variant == 1) ta.g = ws.guid_dev, ta.g_stride = guidance_len == 1 ? 0 : 1;
  API_TRY(net_forward(*e, net, ws, nb, T, mask ? ws.mask[0] : nullptr, ta, ws.vel, s));
  API_K(KC_CFG_EULER, (double)B * T * F * 4.0 * (cfg ? 4 : 3), s,
        API_TRY(cfg_euler_update(x_update, ws.vel, ws.guid_dev, guidance_len, hi ? 1.0f : 2.0f, B, T, F, cfg ? 1 : 0,
                                 dt, v_out, s)));
  return 0;
}

static int check_ready(Engine* e, int variant, int B, int T, const float* guidance, int guidance_len) {
  if (!e->finalized) {
    e->err = "zv_finalize_weights has not been called";
    return 963;
  }
  if (B <= 0 || T <= 0 || !(guidance_len == 1 || guidance_len == B) || !guidance) {
    e->err = "bad arguments (B, T > 0; guidance_len must be 1 or B)";
    return 964;
  }
  if (variant == 1 && !e->cfg.use_guidance_scale_embed) {
    e->err = "variant 1 (distill) needs a model built with use_guidance_scale_embed";
    return 965;
  }
  return 0;
}

// The NFE loop of zv_sample on x1 in place; the time table and the guidance values are already on the device.
static int sample_loop(Engine* e, int variant, const std::vector<float>& ts, float* x1, const float* text,
                       const float* speech, const uint8_t* mask, int B, int T, int num_step, bool all_zero,
                       int guidance_len, cudaStream_t s) {
  const float* ts_dev = e->ws_fm.ts_dev;
  for (int i = 0; i < num_step; ++i)
    API_TRY(velocity_step(e, variant, ts[i], ts_dev + i, x1, text, speech, mask, B, T, all_zero, guidance_len, nullptr,
                          x1, ts[i + 1] - ts[i], s));
  return 0;
}

// Launch-bound batches (a single utterance is ~350 launches of a few microseconds per NFE): the loop is
// captured once per (shape, schedule, guidance mode) as a CUDA graph working on engine-owned staging buffers
// and replayed on later calls.  First call of a key: eager (fills the lazily built position tables, tensor-map
// cache and function attributes); second call: capture + launch; afterwards: launch.  The graph holds workspace
// addresses, so every graph is dropped when the workspace or the staging buffers are reallocated.
static int sample_graphed(Engine* e, int variant, const float* x0, const float* text, const float* speech,
                          const uint8_t* mask, int B, int T, int num_step, bool all_zero, int guidance_len,
                          double t_start, double t_end, double t_shift, const std::vector<float>& ts, float* x1,
                          cudaStream_t s) {
  const int Ft = e->cfg.feat_dim, F = e->fm.out_dim;
  const size_t bt = (size_t)B * T;
  if (!e->gstream) {
    API_TRY((int)cudaStreamCreateWithFlags(&e->gstream, cudaStreamNonBlocking));
    API_TRY((int)cudaEventCreateWithFlags(&e->gev_in, cudaEventDisableTiming));
    API_TRY((int)cudaEventCreateWithFlags(&e->gev_out, cudaEventDisableTiming));
  }
  auto drop_graphs = [&] {
    for (auto& g : e->graphs)
      if (g.exec) cudaGraphExecDestroy(g.exec);
    e->graphs.clear();
  };
  if (bt > e->g_cap) {
    cudaDeviceSynchronize();
    drop_graphs();
    for (void* p : e->gallocs) cudaFree(p);
    e->gallocs.clear();
    API_TRY(dev_alloc(*e, e->gallocs, (void**)&e->gx, bt * F * 4));
    API_TRY(dev_alloc(*e, e->gallocs, (void**)&e->gtext, bt * Ft * 4));
    API_TRY(dev_alloc(*e, e->gallocs, (void**)&e->gspeech, bt * F * 4));
    API_TRY(dev_alloc(*e, e->gallocs, (void**)&e->gmask, bt));
    e->g_cap = bt;
  }
  if (e->graphs_ws_gen != e->ws_fm.gen) {
    drop_graphs();
    e->graphs_ws_gen = e->ws_fm.gen;
  }
  cudaStream_t gs = e->gstream;
  API_TRY((int)cudaEventRecord(e->gev_in, s));
  API_TRY((int)cudaStreamWaitEvent(gs, e->gev_in, 0));
  API_TRY((int)cudaMemcpyAsync(e->gx, x0, bt * F * 4, cudaMemcpyDeviceToDevice, gs));
  API_TRY((int)cudaMemcpyAsync(e->gtext, text, bt * Ft * 4, cudaMemcpyDeviceToDevice, gs));
  API_TRY((int)cudaMemcpyAsync(e->gspeech, speech, bt * F * 4, cudaMemcpyDeviceToDevice, gs));
  if (mask) API_TRY((int)cudaMemcpyAsync(e->gmask, mask, bt, cudaMemcpyDeviceToDevice, gs));
  const uint8_t* gm = mask ? e->gmask : nullptr;

  auto bits = [](double d) {
    unsigned long long u;
    memcpy(&u, &d, 8);
    return u;
  };
  const std::vector<unsigned long long> key = {(unsigned long long)variant, (unsigned long long)B, (unsigned long long)T,
                                               (unsigned long long)num_step, (unsigned long long)all_zero,
                                               (unsigned long long)guidance_len, bits(t_start), bits(t_end), bits(t_shift),
                                               (unsigned long long)(mask != nullptr), (unsigned long long)(e->gemm_impl * 2 + e->fm.stream),
                                               (unsigned long long)zv::g_attn_impl};
  Engine::SampleGraph* g = nullptr;
  for (auto& c : e->graphs)
    if (c.key == key) g = &c;
  ++e->graph_tick;
  if (g) g->last_use = e->graph_tick;
  int rc = 0;
  if (!g) {  // first sighting: eager
    if (e->graphs.size() >= Engine::kMaxGraphs) {  // bounded cache: evict the least recently used key
      size_t lru = 0;
      for (size_t i = 1; i < e->graphs.size(); ++i)
        if (e->graphs[i].last_use < e->graphs[lru].last_use) lru = i;
      if (e->graphs[lru].exec) {
        API_TRY((int)cudaStreamSynchronize(gs));  // a replay of it may still be running
        cudaGraphExecDestroy(e->graphs[lru].exec);
      }
      e->graphs.erase(e->graphs.begin() + lru);
    }
    Engine::SampleGraph c;
    c.key = key;
    c.last_use = e->graph_tick;
    e->graphs.push_back(c);
    rc = sample_loop(e, variant, ts, e->gx, e->gtext, e->gspeech, gm, B, T, num_step, all_zero, guidance_len, gs);
  } else if (g->unsupported) {
    rc = sample_loop(e, variant, ts, e->gx, e->gtext, e->gspeech, gm, B, T, num_step, all_zero, guidance_len, gs);
  } else {
    if (!g->exec) {
      const long long l0 = g_launch_count;
      cudaGraph_t graph = nullptr;
      cudaError_t ce = cudaStreamBeginCapture(gs, cudaStreamCaptureModeThreadLocal);
      if (ce == cudaSuccess) {
        rc = sample_loop(e, variant, ts, e->gx, e->gtext, e->gspeech, gm, B, T, num_step, all_zero, guidance_len, gs);
        ce = cudaStreamEndCapture(gs, &graph);
      }
      g->launches = g_launch_count - l0;
      g_launch_count = l0;  // captured, not executed
      if (rc == 0 && ce == cudaSuccess && graph) ce = cudaGraphInstantiate(&g->exec, graph, 0);
      if (graph) cudaGraphDestroy(graph);
      if (rc != 0 || ce != cudaSuccess || !g->exec) {  // not capturable here: remember and run the plain loop
        cudaGetLastError();
        g->exec = nullptr, g->unsupported = true;
        rc = sample_loop(e, variant, ts, e->gx, e->gtext, e->gspeech, gm, B, T, num_step, all_zero, guidance_len, gs);
      }
    }
    if (g->exec) {
      API_TRY((int)cudaGraphLaunch(g->exec, gs));
      g_launch_count += g->launches;
    }
  }
  if (rc) return rc;
  API_TRY((int)cudaMemcpyAsync(x1, e->gx, bt * F * 4, cudaMemcpyDeviceToDevice, gs));
  API_TRY((int)cudaEventRecord(e->gev_out, gs));
  API_TRY((int)cudaStreamWaitEvent(s, e->gev_out, 0));
  return 0;
}

extern "C" int zv_velocity(void* handle, int variant, float t, const float* x, const float* text,
                           const float* speech, const uint8_t* mask, int B, int T, const float* guidance,
                           int guidance_len, float* v_out, void* stream) {
  Engine* e = (Engine*)handle;
  if (!e) return 960;
  e->err.clear();
  cudaStream_t s = (cudaStream_t)stream;
  API_TRY(check_ready(e, variant, B, T, guidance, guidance_len));
  bool all_zero = true;
  for (int i = 0; i < guidance_len; ++i) all_zero = all_zero && guidance[i] == 0.0f;
  API_TRY(ws_reserve(*e, e->fm, e->ws_fm, 2 * B, T));
  API_TRY(upload_small(e, e->ws_fm.t_dev, &t, 1, s));
  API_TRY(upload_small(e, e->ws_fm.guid_dev, guidance, guidance_len, s));
  return velocity_step(e, variant, t, e->ws_fm.t_dev, x, text, speech, mask, B, T, all_zero, guidance_len, v_out,
                       nullptr, 0.f, s);
}

// ZipVoice.forward_fm_decoder (zipvoice.py:135-185): the raw network call -- no classifier-free-guidance logic.
extern "C" int zv_forward_fm_decoder(void* handle, const float* t, int t_len, const float* x, const float* text,
                                     const float* speech, const uint8_t* mask, int N, int T, const float* guidance,
                                     int guidance_len, float* v_out, void* stream) {
  Engine* e = (Engine*)handle;
  if (!e) return 960;
  e->err.clear();
  cudaStream_t s = (cudaStream_t)stream;
  if (!e->finalized) {
    e->err = "zv_finalize_weights has not been called";
    return 963;
  }
  if (N <= 0 || T <= 0 || !t || !(t_len == 1 || t_len == N) ||
      !(guidance_len == 0 || guidance_len == 1 || guidance_len == N) || (guidance_len > 0 && !guidance)) {
    e->err = "bad arguments (N, T > 0; t_len must be 1 or N; guidance_len 0, 1 or N)";
    return 964;
  }
  if (guidance_len > 0 && !e->cfg.use_guidance_scale_embed) {  // zipformer.py:270-272
    e->err = "guidance_scale needs a model built with use_guidance_scale_embed (ZipVoiceDistill)";
    return 965;
  }
  Net& net = e->fm;
  Workspace& ws = e->ws_fm;
  const int Ft = e->cfg.feat_dim, F = e->fm.out_dim;
  API_TRY(ws_reserve(*e, net, ws, N, T));
  API_TRY(upload_small(e, ws.t_dev, t, t_len, s));
  if (guidance_len > 0) API_TRY(upload_small(e, ws.guid_dev, guidance, guidance_len, s));
  API_K(KC_PACK, (double)N * T * (2 * F + Ft) * 4.0 + (double)N * T * net.in_pad * 2.0, s,
        API_TRY(pack_input(x, text, speech, mask, N, T, F, Ft, 0, 0, ws.x_in, net.in_pad, mask ? ws.mask[0] : nullptr,
                           s)));
  TimeArgs ta;
  ta.t = ws.t_dev, ta.t_stride = t_len == 1 ? 0 : 1;
  if (guidance_len > 0) ta.g = ws.guid_dev, ta.g_stride = guidance_len == 1 ? 0 : 1;
  API_TRY(net_forward(*e, net, ws, N, T, mask ? ws.mask[0] : nullptr, ta, ws.vel, s));
  API_K(KC_CFG_EULER, (double)N * T * F * 4.0 * 2, s,
        API_TRY(cfg_euler_update(nullptr, ws.vel, ws.guid_dev, 1, 1.0f, N, T, F, 0, 0.f, v_out, s)));
  return 0;
}

extern "C" int zv_sample(void* handle, int variant, const float* x0, const float* text, const float* speech,
                         const uint8_t* mask, int B, int T, int num_step, const float* guidance, int guidance_len,
                         double t_start, double t_end, double t_shift, float* x1, void* stream) {
  Engine* e = (Engine*)handle;
  if (!e) return 960;
  e->err.clear();
  cudaStream_t s = (cudaStream_t)stream;
  API_TRY(check_ready(e, variant, B, T, guidance, guidance_len));
  if (num_step < 1 || num_step > 1024) {
    e->err = "num_step out of range";
    return 964;
  }
  bool all_zero = true;
  for (int i = 0; i < guidance_len; ++i) all_zero = all_zero && guidance[i] == 0.0f;
  API_TRY(ws_reserve(*e, e->fm, e->ws_fm, 2 * B, T));
  std::vector<float> ts(num_step + 1);
  zv_time_steps(t_start, t_end, num_step, t_shift, ts.data());
  float* ts_dev = e->ws_fm.ts_dev;  // time table on the device: no host synchronisation inside the loop
  API_TRY(upload_small(e, ts_dev, ts.data(), (int)ts.size(), s));
  API_TRY(upload_small(e, e->ws_fm.guid_dev, guidance, guidance_len, s));
  const size_t n = (size_t)B * T * e->fm.out_dim;
  const bool cfg2 = (variant == 0) && !all_zero;
  static const bool graph_off = getenv("ZV_GRAPH") && atoi(getenv("ZV_GRAPH")) == 0;  // ZV_GRAPH=0: debugging
  if (!graph_off && e->graph_max_rows > 0 && !e->prof.on &&
      (size_t)(cfg2 ? 2 : 1) * B * T <= (size_t)e->graph_max_rows)
    return sample_graphed(e, variant, x0, text, speech, mask, B, T, num_step, all_zero, guidance_len, t_start, t_end,
                          t_shift, ts, x1, s);
  if (x1 != x0) API_TRY((int)cudaMemcpyAsync(x1, x0, n * 4, cudaMemcpyDeviceToDevice, s));
  return sample_loop(e, variant, ts, x1, text, speech, mask, B, T, num_step, all_zero, guidance_len, s);
}

extern "C" int zv_text_embed(void* handle, const int64_t* tokens_padded, const int64_t* tokens_lens, int B, int S,
                             float* embed_out, void* stream) {
  Engine* e = (Engine*)handle;
  if (!e) return 960;
  e->err.clear();
  cudaStream_t s = (cudaStream_t)stream;
  if (!e->finalized) {
    e->err = "zv_finalize_weights has not been called";
    return 963;
  }
  Net& net = e->text;
  Workspace& ws = e->ws_text;
  API_TRY(ws_reserve(*e, net, ws, B, S));
  const size_t total = (size_t)B * S * net.in_pad;
  API_K(KC_TEXT, (double)total * 6.0, s,
        (embed_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(tokens_padded, tokens_lens, e->embed, S,
                                                                      e->cfg.text_embed_dim, net.in_pad,
                                                                      e->cfg.vocab_size, ws.x_in, ws.mask[0], total)));
  API_TRY(net_forward(*e, net, ws, B, S, ws.mask[0], TimeArgs(), embed_out, s));
  if (e->cfg.dialog) {
    API_K(KC_TEXT, (double)B * S * e->cfg.feat_dim * 8.0, s,
          (spk_add_kernel<<<B, 128, 0, s>>>(tokens_padded, B, S, e->cfg.feat_dim, e->cfg.pad_id, e->cfg.spk_a_id,
                                            e->cfg.spk_b_id, e->spk_embed, embed_out)));
  }
  return 0;
}

extern "C" int zv_text_condition(void* handle, const float* embed, const int64_t* tokens_lens,
                                 const int64_t* features_lens, int B, int S, int T, float* out, uint8_t* mask_out,
                                 void* stream) {
  Engine* e = (Engine*)handle;
  if (!e) return 960;
  e->err.clear();
  const size_t total = (size_t)B * T * e->cfg.feat_dim;
  cudaStream_t s = (cudaStream_t)stream;
  API_K(KC_TEXT, (double)total * 8.0, s,
        (text_condition_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(embed, tokens_lens, features_lens, S, T,
                                                                               e->cfg.feat_dim, out, mask_out, total)));
  return 0;
}

