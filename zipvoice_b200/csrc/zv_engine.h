// Engine internals: packed weights, workspace and the per-NFE schedule of kernel launches.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/zipvoice_b200.h"
#include "zv_gemm.h"
#include "zv_kernels.h"

namespace zv {

struct HostTensor {
  std::vector<float> data;
  std::vector<int64_t> shape;
  size_t numel() const { return data.size(); }
};

// One packed Linear: bf16 weight [N_pad][ldw] (rows possibly permuted for fused epilogues) + fp32 bias.
struct Lin {
  __nv_bfloat16* w = nullptr;
  float* bias = nullptr;  // accumulator-column order, length N
  int N = 0, N_pad = 0, K = 0, K_valid = 0, ldw = 0, block_n = 0, mode = 0, gate_cols = 0;
};

struct Layer {
  Lin attn_in, ff_in[3], ff_out[3], na_in, na_out, sa_in[2], conv_in[2], conv_out[2];  // SelfAttention.out_proj lives inside conv_in / conv_out
  // feed_forward1.in_proj and the attention q | k | p projection read the same rows (the layer input): one launch,
  // weight rows [FF1 in_proj | attention in_proj], EPI_SWOOSHL_BF16 with gate_cols = FF1's width.  attn_in and
  // ff_in[0] above stay packed for the measurement switch ZV_MERGE_QKP=0.
  Lin ff1_qkp;
  float* w_pos = nullptr;  // [H*dp][pos_dim]
  float* dw_w[2] = {nullptr, nullptr};
  float* dw_b[2] = {nullptr, nullptr};
  float* norm_bias = nullptr;
  float log_scale = 0.f;
  float* bypass_scale = nullptr;
  float* bypass_mid_scale = nullptr;
  size_t te_off_ff1 = 0;          // column offset of this layer's FF1 row-bias in the te projection
  size_t te_off_conv[2] = {0, 0}; // ... of the two conv in_proj row-biases (W_in te_stack, accumulator order)
  std::map<int, float*> pos_tab;  // L -> E[H][2L-1][dp]
  std::map<int, void*> pos_tab16; // L -> pos_table_f16 of it (SelfAttention kernel's tensor-core position term)
};

struct Stack {
  int ds = 1, kernel = 0;
  float ds_w[4] = {1.f, 0.f, 0.f, 0.f};  // softmax(downsample.bias)
  float* out_comb_scale = nullptr;
  size_t te_off = 0;  // column offset of te_stack (D columns) in the te projection
  std::vector<Layer> layers;
};

struct Net {
  std::string prefix;
  int in_dim = 0, in_pad = 0, out_dim = 0, D = 0, F = 0, H = 0, hidden_na = 0;
  bool use_time = false, use_guidance = false;
  Lin in_proj, out_proj;
  // TTSZipformerTwoStream (zipformer_two_stream.py:139-150,238-244) keeps TWO projection pairs and picks one by the
  // width of its input.  in_proj / out_proj / in_dim / in_pad / out_dim above are the ACTIVE pair; the other one
  // waits here and zv_set_option("stream", i) swaps them.  Workspaces are sized for the wider pair.
  bool two_stream = false;
  int stream = 0;  // index of the active pair (0: two channels, 1: mono)
  Lin alt_in_proj, alt_out_proj;
  int alt_in_dim = 0, alt_in_pad = 0, alt_out_dim = 0;
  int in_pad_max = 0, out_dim_max = 0;
  float *te_w0 = nullptr, *te_b0 = nullptr, *te_w2 = nullptr, *te_b2 = nullptr, *te_wg = nullptr;
  float* te_proj_w = nullptr;  // [N_te][time_dim] folded (stack time_emb, FF1 in_proj . time_emb)
  float* te_proj_b = nullptr;  // [N_te]
  int N_te = 0;
  std::vector<Stack> stacks;
};

struct Workspace {
  size_t cap_rows = 0;  // nb * T capacity
  unsigned long long gen = 0;  // bumped whenever the buffers are reallocated (captured graphs hold their addresses)
  int cap_nb = 0;
  int D = 0;
  float *vel = nullptr, *xcur = nullptr;
  // residual streams in the s24 storage format (zv_common.cuh): X = stack input / output at full rate, Xs = the
  // same at the stack's frame rate, Y = working stream inside a layer.  The hi planes are the GEMM A operands.
  __nv_bfloat16 *X_hi = nullptr, *Xs_hi = nullptr, *Y_hi = nullptr;
  uint8_t *X_lo = nullptr, *Xs_lo = nullptr, *Y_lo = nullptr;
  __nv_bfloat16 *x_in = nullptr, *qkp = nullptr, *hidden = nullptr, *na_v = nullptr,
                *na_y = nullptr, *na_o = nullptr, *sa_v = nullptr, *conv_a = nullptr,
                *conv_b = nullptr;
  float* sa_part = nullptr;   // key-split partial states of the SelfAttention kernel (AttnParams::part)
  size_t sa_part_cap = 0;     // bytes
  uint8_t* mask[3] = {nullptr, nullptr, nullptr};  // per distinct downsampling rate (1, 2, 4)
  float *t_dev = nullptr, *ts_dev = nullptr /* time table of one sample(): 1025 floats */, *g_dev = nullptr, *te_a = nullptr, *te_out = nullptr, *guid_dev = nullptr;
  std::vector<void*> allocs;
};

// Per-kernel-class device timing (CUDA events on the launching stream), enabled by zv_prof_enable.
// work = algorithmic FLOPs (tensor-bound classes) or algorithmic bytes (HBM-bound classes) of a launch.
enum KClass : int {
  KC_GEMM_BIAS = 0, KC_GEMM_SWOOSHL, KC_GEMM_GLU, KC_GEMM_NONLIN, KC_GEMM_RESIDUAL, KC_GEMM_F32, KC_GEMM_GELU,
  KC_ATTN_SELF, KC_ATTN_NONLIN, KC_DWCONV, KC_BIASNORM, KC_RESAMPLE, KC_PACK, KC_CFG_EULER, KC_TIME_EMBED,
  KC_TEXT, KC_FBANK, KC_VOCODER_ELEM, KC_COUNT
};
const char* kclass_name(int c);
bool kclass_is_tensor(int c);
struct Prof {
  bool on = false;
  struct Rec { int cls; double work; cudaEvent_t a, b; };
  std::vector<Rec> recs;
  std::vector<cudaEvent_t> pool;
  size_t used = 0;
  cudaEvent_t get();
  void begin(int cls, double work, cudaStream_t s);
  void end(cudaStream_t s);
  // synchronises, accumulates elapsed time per class and resets the record list
  int read(double* ms, double* work, long long* launches);
  void release();
};

struct Engine {
  zv_config cfg;
  Prof prof;
  std::map<std::string, HostTensor> raw;
  bool finalized = false;
  Net fm, text;
  float* embed = nullptr;      // [vocab][text_embed_dim]
  float* spk_embed = nullptr;  // [2][feat_dim]
  Workspace ws_fm, ws_text;
  std::vector<void*> weight_allocs;
  std::string err;
  int gemm_impl = 0;
  // CUDA-graph replay of whole zv_sample calls for launch-bound batches (zv_api.cu): the NFE loop of a
  // (shape, schedule, guidance mode) key is captured once on `gstream` from engine-owned staging buffers
  // and replayed afterwards.  graph_max_rows: nb * T at or below which the graph path is taken (0 = off).
  struct SampleGraph {
    std::vector<unsigned long long> key;
    cudaGraphExec_t exec = nullptr;  // null: key seen once (ran eagerly), capture on the next call
    long long launches = 0;          // kernels inside the graph (added to the launch counter per replay)
    bool unsupported = false;        // capture failed once: stay on the eager path
    unsigned long long last_use = 0; // graph_tick of the last call with this key (least recently used is evicted)
  };
  static constexpr size_t kMaxGraphs = 16;  // instantiated graphs kept per handle (~350 kernel nodes per NFE each)
  unsigned long long graph_tick = 0;
  int graph_max_rows = 32768;
  std::vector<SampleGraph> graphs;
  unsigned long long graphs_ws_gen = 0;
  cudaStream_t gstream = nullptr;
  cudaEvent_t gev_in = nullptr, gev_out = nullptr;
  std::vector<void*> gallocs;
  size_t g_cap = 0;  // staging capacity in B * T elements
  float *gx = nullptr, *gtext = nullptr, *gspeech = nullptr;
  uint8_t* gmask = nullptr;
  // Small host -> device uploads (time table, guidance values) go through a ring of PINNED slots so that the copy is
  // truly asynchronous (a cudaMemcpyAsync from pageable memory stages synchronously).  A slot is reused only after
  // the event recorded behind its last copy has completed.
  static constexpr int kPinSlots = 8, kPinSlotFloats = 4096;
  float* pin_host = nullptr;
  cudaEvent_t pin_ev[kPinSlots] = {};
  int pin_next = 0;
  // audio front-end / vocoder state lives in zv_audio.cu
  void* audio = nullptr;
};

extern long long g_launch_count;

int engine_finalize(Engine& e);
// One TTSZipformer forward on nb sequences of T frames.  ws.x_in holds the bf16 input rows
// [nb*T][in_pad]; the result (fp32 [nb*T][out_dim]) is written to `out`.
struct TimeArgs {
  const float* t = nullptr;  // device; value of sequence b is t[b * t_stride]
  int t_stride = 0;
  const float* g = nullptr;  // device guidance scale for the distilled model (embedded), or null
  int g_stride = 0;
};
int net_forward(Engine& e, Net& net, Workspace& ws, int nb, int T, const uint8_t* mask_dev, const TimeArgs& ta,
                float* out, cudaStream_t s);
int ws_reserve(Engine& e, Net& net, Workspace& ws, int nb, int T);
// Position tables E[H][2L-1][4] are cached per (layer, L), at most kMaxPosTables lengths per layer; when a layer is
// full its tables are freed (after a device synchronise) and ws_fm.gen is bumped so that captured graphs, which hold
// their addresses, are dropped.  free_pos_tables releases them all (zv_destroy).
static constexpr size_t kMaxPosTables = 32;
void free_pos_tables(Net& net);
void audio_destroy(Engine& e);
// Shared with the vocoder (zv_audio.cu): pack one Linear from host fp32 (rows optionally permuted for the
// fused epilogue of `mode`), and launch it.  row_scale (optional, [N]) multiplies weight rows and bias.
int pack_linear_host(Engine& e, const float* w, const float* bias, int N, int K, int mode, Lin* lin, int k_pad_to,
                     const float* row_scale);
int run_gemm(Engine& e, const Lin& lin, const __nv_bfloat16* A, int lda, int M, EpiParams ep, cudaStream_t s,
             const __nv_bfloat16* A2 = nullptr, int lda2 = 0, int k_split = 0);
int dev_alloc(Engine& e, std::vector<void*>& pool, void** p, size_t bytes);

}  // namespace zv
