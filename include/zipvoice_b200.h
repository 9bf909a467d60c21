/* zipvoice_b200 -- C ABI of the B200-native ZipVoice hot path (libzipvoice_b200.so).
 *
 * Conventions: every function returns 0 on success, non-zero error code otherwise (never throws);
 * zv_last_error() gives the text.  All tensor pointers are DEVICE pointers to contiguous fp32 /
 * uint8 / int64 data unless stated; calls are asynchronous on `stream` (a cudaStream_t passed as
 * void*).  A handle is bound to the current CUDA device and is not thread-safe.
 * Citations are file:line in the reference repository (winlaic/ZipVoice).
 */
#ifndef ZIPVOICE_B200_H
#define ZIPVOICE_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ZV_MAX_STACKS 8

/* Model hyper-parameters: the "model" section of model.json (egs/zipvoice/conf/zipvoice_base.json:1-25)
 * = the keyword arguments of ZipVoice.__init__ (zipvoice/models/zipvoice.py:38-60). */
typedef struct zv_config {
  int feat_dim;
  int fm_num_stacks;
  int fm_downsampling_factor[ZV_MAX_STACKS];
  int fm_num_layers[ZV_MAX_STACKS];
  int fm_cnn_module_kernel[ZV_MAX_STACKS];
  int fm_feedforward_dim, fm_num_heads, fm_dim;
  int text_num_layers, text_feedforward_dim, text_cnn_module_kernel, text_num_heads, text_dim;
  int query_head_dim, value_head_dim, pos_head_dim, pos_dim;
  int time_embed_dim, text_embed_dim;
  int vocab_size, pad_id;
  int use_guidance_scale_embed; /* 1 for ZipVoiceDistill (zipvoice_distill.py:52-67) */
  int dialog;                   /* 1 for ZipVoiceDialog: spk_embed rows added (zipvoice_dialog.py:127-159) */
  int spk_a_id, spk_b_id;
  int stereo;                   /* 1 for ZipVoiceDialogStereo (zipvoice_dialog.py:218-256): the fm_decoder is a
                                 * TTSZipformerTwoStream (zipformer_two_stream.py:219-264) used through its first
                                 * stream: x and speech_condition are 2*feat_dim wide (two channels), the text
                                 * condition stays feat_dim wide; in_proj.0 (5*feat_dim -> D), out_proj.0 (D -> 2*feat_dim) */
} zv_config;

/* Replaces ZipVoice.__init__ (zipvoice.py:38-133): builds the fm_decoder + text_encoder skeleton. */
int zv_create(const zv_config* cfg, void** handle);
void zv_destroy(void* handle);
const char* zv_last_error(void* handle);

/* Weights enter by state_dict name (torch nn.Module.load_state_dict equivalent,
 * bin/infer_zipvoice.py:561-566).  `data` is fp32, host (on_device=0) or device (on_device=1). */
int zv_set_weight(void* handle, const char* name, const float* data, const int64_t* shape, int ndim, int on_device);
/* Packs weights into the device arena (bf16 GEMM operands, permuted GLU / gate layouts, folded
 * time-embedding projections).  Fails if a tensor of the architecture is missing. */
int zv_finalize_weights(void* handle);

/* options: "gemm_impl" 0 = tcgen05 kernel, CTA pairs (cta_group::2) for large problems (default), 1 = SIMT checker
 * kernel (tests only), 2 = tcgen05 single-CTA tiles only, 3 = tcgen05 CTA pairs only;
 * "attn_impl" 0 = NonlinAttention on the tcgen05 kernel (default), 1 = mma.sync flash kernel, 2 = as 0 with the
 * exact-maxima repeat of the NonlinAttention kernel forced (tests) (process-wide);
 * "stream" (two-stream / stereo models only) 0 = the two-channel projection pair in_proj.0 / out_proj.0 (default; x and
 * speech 2*feat_dim wide), 1 = the mono pair in_proj.1 / out_proj.1 (x and speech feat_dim wide) -- the choice
 * TTSZipformerTwoStream.forward makes from the input width (zipformer_two_stream.py:238-244);
 * "graph_max_rows" (default 32768; 0 = off): zv_sample calls whose network batch x frames is at or below this
 * value are launch-bound; from the third call of a (shape, schedule, guidance mode) key on they replay a CUDA
 * graph of the whole NFE loop captured on an engine-owned stream from engine-owned input copies (results are
 * bit-identical to the plain loop; the caller's stream waits on the replay). */
int zv_set_option(void* handle, const char* key, int value);

/* get_time_steps (models/modules/solver.py:256-281), fp32 arithmetic identical to torch
 * (python-double scalars, fp32 tensor ops, FMA linspace); host output. */
int zv_time_steps(double t_start, double t_end, int num_step, double t_shift, float* out_host /* num_step+1 */);

/* One call of DiffusionModel.forward (variant 0, solver.py:40-110: CFG batch doubling, the t>0.5
 * branch, (1+g) cond - g uncond) or DistillDiffusionModel.forward (variant 1, solver.py:127-165)
 * around ZipVoice.forward_fm_decoder (zipvoice.py:135-185) -> TTSZipformer.forward
 * (zipformer.py:242-293).  x,speech,v_out: (B,T,X) fp32 with X = feat_dim (2*feat_dim when cfg.stereo); text:
 * (B,T,feat_dim); padding_mask (B,T) uint8,
 * 1 = padded, may be NULL; guidance: HOST floats, guidance_len 1 or B. */
int zv_velocity(void* handle, int variant, float t, const float* x, const float* text_condition,
                const float* speech_condition, const uint8_t* padding_mask, int B, int T, const float* guidance,
                int guidance_len, float* v_out, void* stream);

/* ZipVoice.forward_fm_decoder (zipvoice.py:135-185) -> TTSZipformer.forward (zipformer.py:242-293): the network
 * itself, without the classifier-free-guidance logic of DiffusionModel.  t: HOST floats, t_len 1 (the 0-dim case,
 * :168-169) or N (the (N,1,1) case, :165-167: one time step per row); guidance: HOST floats embedded by the distilled
 * model (:171-178), guidance_len 0 (guidance_scale=None), 1 or N.  x, speech, v_out (N,T,X); text (N,T,feat_dim). */
int zv_forward_fm_decoder(void* handle, const float* t, int t_len, const float* x, const float* text_condition,
                          const float* speech_condition, const uint8_t* padding_mask, int N, int T,
                          const float* guidance, int guidance_len, float* v_out, void* stream);

/* EulerSolver.sample / DistillEulerSolver.sample (solver.py:182-240).  x0 is not modified; x1 receives
 * the solution at t_end.  No host synchronisation inside. */
int zv_sample(void* handle, int variant, const float* x0, const float* text_condition,
              const float* speech_condition, const uint8_t* padding_mask, int B, int T, int num_step,
              const float* guidance, int guidance_len, double t_start, double t_end, double t_shift, float* x1,
              void* stream);

/* ZipVoice.forward_text_embed (zipvoice.py:187-212; dialog variant zipvoice_dialog.py:127-159):
 * tokens_padded (B,S) int64 as produced by pad_labels (utils/common.py:255-268), tokens_lens (B) int64
 * -> embed (B,S,feat_dim). */
int zv_text_embed(void* handle, const int64_t* tokens_padded, const int64_t* tokens_lens, int B, int S,
                  float* embed_out, void* stream);
/* ZipVoice.forward_text_condition (zipvoice.py:214-251) with prepare_avg_tokens_durations /
 * get_tokens_index (utils/common.py:246-295) evaluated on the device.  Also writes the padding
 * mask (B,T) uint8 (make_pad_mask, utils/common.py:395-420). */
int zv_text_condition(void* handle, const float* embed, const int64_t* tokens_lens, const int64_t* features_lens,
                      int B, int S, int T, float* text_condition_out, uint8_t* padding_mask_out, void* stream);

/* VocosFbank.extract / _feature_fn (zipvoice/utils/feature.py:56-116): wav (B,N) fp32 24 kHz mono,
 * lens (B) int64 valid samples -> log-mel (B, T, 100), T = (N + 128) / 256 frames. */
int zv_fbank(void* handle, const float* wav, const int64_t* lens, int B, int N, float* mel_out, void* stream);
/* BigVGANFbank.extract (zipvoice/utils/feature.py:133-204 on zipvoice/utils/_bigvgan_mel_feature.py:42-111; the
 * alternative feature type of conf/zipvoice_base_bigvgan_v2.json, selected at bin/infer_zipvoice.py:585-586): reflect
 * padding 384, un-centred frames, sqrt(|X|^2 + 1e-9), librosa Slaney mel basis, log(clamp(., 1e-5)).  Same shapes
 * as zv_fbank. */
int zv_fbank_bigvgan(void* handle, const float* wav, const int64_t* lens, int B, int N, float* mel_out,
                     void* stream);

/* torchaudio.transforms.Resample(orig_freq, new_freq) as the reference applies it to the prompt recording
 * (bin/infer_zipvoice.py:334-339; torchaudio defaults: sinc_interp_hann, lowpass_filter_width 6, rolloff 0.99):
 * wav (B,N) fp32, lens (B) int64 valid samples per row or NULL -> out (B, n_out) with
 * n_out = ceil(new_freq * N / orig_freq); row b is 0 from sample ceil(new_freq * lens[b] / orig_freq) on. */
int zv_resample(void* handle, const float* wav, const int64_t* lens, int B, int N, int orig_freq, int new_freq,
                float* out, int n_out, void* stream);

/* zv_resample without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same arguments, results and error codes as zv_resample, which cudaMallocs its polyphase filter
 * bank ([new][K] fp32, new / K from the reduced rate ratio) the first time it sees a rate pair.  Here the bank is
 * SCRATCH: every call rebuilds it on the device (double arithmetic, the formula of zv_resample) into a file-scope
 * `static __device__` array of ZV_RESAMPLE_STATIC_MAX_FLOATS floats (256 KB, held by every GPU that loads the
 * library until the process exits) and then runs the resampling kernel on it.  The call makes no cudaMalloc /
 * cudaFree / allocating library call and no host<->device copy; the caller owns wav, lens and out.
 *  - Limit: new_r * K <= ZV_RESAMPLE_STATIC_MAX_FLOATS, with new_r = new_freq / gcd, orig_r = orig_freq / gcd,
 *    K = 2 * ceil(6 * orig_r / (0.99 * min(orig_r, new_r))) + orig_r (8 / 11.025 / 16 / 22.05 / 32 / 44.1 / 48 / 96 kHz
 *    <-> 24 kHz all fit; 44100 -> 24000 needs 13 680).  Larger banks and B > 65535 are rejected with error 967 and a
 *    message in zv_last_error; nothing is launched.
 *  - The array is shared by every call on that GPU (all handles, all streams) and keeps the last call's contents.
 *    Nothing in it has to start at zero: each call overwrites every entry it reads.  Calls that OVERLAP on the GPU
 *    (different streams, or graph branches without an edge between them) race on it; calls on one stream do not.
 *  - The first call on a device looks the array's device address up (cudaGetSymbolAddress, once per device, that
 *    device current); make that first call outside stream capture. */
#define ZV_RESAMPLE_STATIC_MAX_FLOATS 65536
int zv_resample_static(void* handle, const float* wav, const int64_t* lens, int B, int N, int orig_freq,
                       int new_freq, float* out, int n_out, void* stream);

/* Vocos.decode (third-party `vocos`; call site bin/infer_zipvoice.py:378): mel (B,100,T) -> wav (B, 256*(T-1)).
 * Weights are set through zv_set_weight with the "vocoder." prefix (vocos checkpoint names: backbone.*,
 * head.out.*).  lens (B) int64 valid frames per row, or NULL for full rows: frames >= lens[b] are treated
 * as absent (zero-padded convolutions, as if each utterance were decoded alone) and their samples are 0. */
int zv_vocos_decode(void* handle, const float* mel, const int64_t* lens, int B, int T, float* wav_out, void* stream);

/* Per-kernel-class device timing: when enabled every launch is bracketed by CUDA events on its stream.
 * zv_prof_read synchronises the device, fills ms / work / launches (arrays of zv_prof_num_classes()) and
 * resets the records.  work = algorithmic FLOPs for tensor-bound classes, algorithmic bytes otherwise. */
int zv_prof_enable(void* handle, int on);
int zv_prof_num_classes(void);
const char* zv_prof_class_name(int cls);
int zv_prof_class_is_tensor(int cls);
int zv_prof_read(void* handle, double* ms, double* work, long long* launches);
/* Launch counters: kernels launched by this library since the last reset. */
long long zv_launch_count(void);
void zv_launch_count_reset(void);

#ifdef __cplusplus
}
#endif
#endif
