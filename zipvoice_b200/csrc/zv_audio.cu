// VocosFbank mel front-end and Vocos vocoder head.
#include "zv_common.cuh"
#include "zv_engine.h"

using namespace zv;

extern "C" int zv_fbank(void* handle, const float* wav, const int64_t* lens, int B, int N, float* mel_out,
                        void* stream) {
  Engine* e = (Engine*)handle;
  if (!e) return 960;
  e->err = "zv_fbank: kernel not built yet in this revision";
  return 990;
}

extern "C" int zv_vocos_decode(void* handle, const float* mel, int B, int T, float* wav_out, void* stream) {
  Engine* e = (Engine*)handle;
  if (!e) return 960;
  e->err = "zv_vocos_decode: kernel not built yet in this revision";
  return 990;
}
