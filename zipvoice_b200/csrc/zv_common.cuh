// Common device helpers for the zipvoice_b200 kernels (sm_100a only).
// PTX wrappers for mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (MMA / TMEM) and
// small math used by several kernels.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define ZV_DEVINL __device__ __forceinline__

namespace zv {

// Function attributes (the dynamic shared-memory opt-in) and occupancy answers are PER DEVICE: a launcher keeps one
// bit per device ordinal and runs its set-up the first time it is called with that device current.
inline bool first_use_on_device(unsigned long long& mask) {
  int dev = 0;
  cudaGetDevice(&dev);
  const unsigned long long bit = 1ull << (dev & 63);
  if (mask & bit) return false;
  mask |= bit;
  return true;
}

typedef __nv_bfloat16 bf16;

// ------------------------------------------------------------------ math
// SwooshL / SwooshR as evaluated by the reference (scaling.py:1174-1191):
// log(1 + exp(x - off)) - 0.08 x - c, with "if the log overflowed to inf use x - off".
ZV_DEVINL float softplus_guard(float xo) {
  float e = __expf(xo);
  float ls = __logf(1.0f + e);
  return (xo > 80.0f) ? xo : ls;  // exp overflows to inf above ~88.7; reference falls back to xo
}
ZV_DEVINL float swoosh_l(float x) { return softplus_guard(x - 4.0f) - 0.08f * x - 0.035f; }
ZV_DEVINL float swoosh_r(float x) { return softplus_guard(x - 1.0f) - 0.08f * x - 0.313261687f; }
ZV_DEVINL float sigmoidf_(float x) { return 1.0f / (1.0f + __expf(-x)); }
ZV_DEVINL float tanhf_(float x) {
  // tanh(x) = 1 - 2/(exp(2x)+1); exact enough in fp32 and safe for large |x|
  float e = __expf(2.0f * x);
  return 1.0f - 2.0f / (e + 1.0f);
}


ZV_DEVINL uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
ZV_DEVINL uint32_t pack_f16x2(float lo, float hi) {
  __half2 v = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
// fp32 pair -> fp16 pair, round to nearest, SATURATING at +-65504 instead of overflowing to +-inf (NaN stays NaN):
// the attention q | k | p rows are stored in fp16 un-scaled, and an inf there would turn a whole softmax row into NaN.
ZV_DEVINL uint32_t pack_f16x2_sat(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
ZV_DEVINL __half f2h_sat(float x) {
  const uint32_t d = pack_f16x2_sat(x, 0.f);
  return __ushort_as_half((unsigned short)(d & 0xffffu));
}
ZV_DEVINL float bf16_lo(uint32_t v) { return __uint_as_float(v << 16); }
ZV_DEVINL float bf16_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

// ------------------------------------------------------------------ s24: the residual stream's storage format
// The residual stream of a Zipformer stack is an fp32 quantity that every module reads and rewrites.  It is kept in
// HBM as TWO planes of the same [rows][D] shape: `hi`, the value rounded to bf16 -- which is also the A operand of the
// next in_proj GEMM, so no separate bf16 copy of the stream exists -- and `lo`, one more byte of mantissa.  3 bytes per
// element instead of 4 (fp32) + 2 (bf16 copy); 16 significant bits (relative rounding error 2^-17 per rewrite, against
// 2^-9 for the bf16 operands of the GEMMs that produce the increments).
//   w  = bits(v) + 0x8080        (round the magnitude to 24 bits, half up, and pre-add half a bf16 ulp)
//   hi = w >> 16                 (bf16 of v, rounded half up)
//   lo = (w >> 8) & 0xff         (biased: the rounding of hi is undone by the - 0x8000 below)
//   v' = bits^-1(((hi << 16) | (lo << 8)) - 0x8000)  =  v rounded to 24 bits
ZV_DEVINL uint32_t s24_w(float v) { return __float_as_uint(v) + 0x8080u; }
ZV_DEVINL float s24_dec(uint32_t hi16, uint32_t lo8) {
  return __uint_as_float(((hi16 << 16) | (lo8 << 8)) - 0x8000u);
}
// 8 consecutive elements: hi = 8 bf16 (16 bytes), lo = 8 bytes
ZV_DEVINL void s24_dec8(const uint4& hi, const uint2& lo, float (&v)[8]) {
  const uint32_t h[4] = {hi.x, hi.y, hi.z, hi.w};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t lw = k < 2 ? lo.x : lo.y;
    // bytes [x, lo(2k), h.b0, h.b1] and [x, lo(2k+1), h.b2, h.b3]; byte 0 is cleared by the mask
    const uint32_t e = __byte_perm(h[k], lw, 0x1040u | (((2 * k) & 3) << 4));
    const uint32_t o = __byte_perm(h[k], lw, 0x3240u | (((2 * k + 1) & 3) << 4));
    v[2 * k] = __uint_as_float((e & 0xffffff00u) - 0x8000u);
    v[2 * k + 1] = __uint_as_float((o & 0xffffff00u) - 0x8000u);
  }
}
ZV_DEVINL void s24_enc8(const float (&v)[8], uint4& hi, uint2& lo) {
  uint32_t w[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) w[i] = s24_w(v[i]);
  hi.x = __byte_perm(w[0], w[1], 0x7632u), hi.y = __byte_perm(w[2], w[3], 0x7632u);
  hi.z = __byte_perm(w[4], w[5], 0x7632u), hi.w = __byte_perm(w[6], w[7], 0x7632u);
  lo.x = __byte_perm(__byte_perm(w[0], w[1], 0x0051u), __byte_perm(w[2], w[3], 0x0051u), 0x5410u);
  lo.y = __byte_perm(__byte_perm(w[4], w[5], 0x0051u), __byte_perm(w[6], w[7], 0x0051u), 0x5410u);
}

ZV_DEVINL float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
ZV_DEVINL float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ------------------------------------------------------------------ smem / mbarrier
ZV_DEVINL uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

ZV_DEVINL void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
ZV_DEVINL void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
ZV_DEVINL void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

ZV_DEVINL void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
ZV_DEVINL void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
ZV_DEVINL bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
ZV_DEVINL void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
// For producer warps that run several stages ahead: back off between polls so that the spinning warp does not
// take issue slots from the compute warps of its scheduler (the wake-up delay is hidden by the ring depth).
ZV_DEVINL void mbar_wait_backoff(uint64_t* bar, uint32_t parity, unsigned ns = 128) {
  while (!mbar_try_wait(bar, parity)) __nanosleep(ns);
}

// ------------------------------------------------------------------ programmatic dependent launch
// A kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may start while its predecessor in
// the stream is still running: everything before griddep_wait() must not touch global memory that another
// kernel writes (set-up of barriers, TMEM, descriptor prefetch is fine); griddep_wait() returns once the
// predecessor grid has completed and its writes are visible.  griddep_launch() lets the successor start its own
// set-up as soon as SM resources free up.  Both are no-ops for a normally launched kernel.
ZV_DEVINL void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
ZV_DEVINL void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ------------------------------------------------------------------ TMA
ZV_DEVINL void tma_prefetch_desc(const void* desc) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(desc)) : "memory");
}
// 2D tiled load: coordinates (c0 = innermost element index, c1 = row index).
ZV_DEVINL void tma_load_2d(void* smem_dst, const void* desc, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(desc)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}

// 3D tiled load: coordinates (c0 innermost, c1, c2); elements outside the tensor (also negative
// coordinates) are zero-filled and still counted in the transaction bytes.
ZV_DEVINL void tma_load_3d(void* smem_dst, const void* desc, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(desc)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}

// 2D tiled store smem -> global (bulk async group); rows / columns outside the tensor are clipped.
ZV_DEVINL void tma_store_2d(const void* desc, const void* smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(desc)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
ZV_DEVINL void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// wait until at most N of this thread's bulk groups still READ their shared-memory source
template <int N>
ZV_DEVINL void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
ZV_DEVINL void tma_store_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// named barrier among `count` threads (count multiple of 32), id 1..15
ZV_DEVINL void named_bar_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}

// ------------------------------------------------------------------ fast math (MUFU approximations)
ZV_DEVINL float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Exact (erf) GELU in 13 instructions with one MUFU op, for the GEMM epilogue that is bound by its issue
// rate: erf(t) = 1 - 2^(-t Q(t)) on t = min(|x| / sqrt 2, 4) with a degree-6 minimax Q (fit of
// -log2(erfc t) / t weighted by the sensitivity of erf; max |error| of erf 1.7e-7 evaluated in fp32, the
// size of erff's own 2 ulp), and gelu(x) = x/2 + |x/2| erf(|x| / sqrt 2). erfc(4) = 1.5e-8 < 2^-24, so
// the clamp changes nothing in fp32.
ZV_DEVINL float gelu_erf(float x) {
  const float t = fminf(fabsf(x) * 0.70710678118654752f, 4.0f);
  float q = -0.00010020830814028159f;
  q = fmaf(q, t, 0.00046149984700605273f);
  q = fmaf(q, t, 0.0023023956455290318f);
  q = fmaf(q, t, -0.029452674090862274f);
  q = fmaf(q, t, 0.14896374940872192f);
  q = fmaf(q, t, 0.9183286428451538f);
  q = fmaf(q, t, 1.6279137134552002f);
  const float e = 1.0f - ex2_approx(-q * t);
  const float h = 0.5f * x;
  return fmaf(fabsf(h), e, h);
}
ZV_DEVINL float lg2_approx(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
ZV_DEVINL float rcp_approx(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
ZV_DEVINL float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// Packed pairs of fp32 (Blackwell FFMA2 / FADD2 / FMUL2: two fp32 operations per issue slot).
typedef unsigned long long f32x2;
ZV_DEVINL f32x2 f2_pack(float a, float b) {
  f32x2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
ZV_DEVINL void f2_unpack(f32x2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
ZV_DEVINL f32x2 f2_fma(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
ZV_DEVINL f32x2 f2_add(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
ZV_DEVINL f32x2 f2_mul(f32x2 a, f32x2 b) {
  f32x2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
// SwooshL on a pair (same formula as swoosh_l_fast)
ZV_DEVINL void swoosh_l_fast2(float& x0, float& x1) {
  const float L2E = 1.4426950408889634f;
  const f32x2 x = f2_pack(x0, x1);
  float t0, t1;
  f2_unpack(f2_fma(x, f2_pack(L2E, L2E), f2_pack(-4.0f * L2E, -4.0f * L2E)), t0, t1);
  const f32x2 s = f2_add(f2_pack(ex2_approx(-fabsf(t0)), ex2_approx(-fabsf(t1))), f2_pack(1.0f, 1.0f));
  float s0, s1;
  f2_unpack(s, s0, s1);
  const f32x2 q = f2_add(f2_pack(fmaxf(t0, 0.f), fmaxf(t1, 0.f)), f2_pack(lg2_approx(s0), lg2_approx(s1)));
  const f32x2 lin = f2_fma(x, f2_pack(-0.08f, -0.08f), f2_pack(-0.035f, -0.035f));
  f2_unpack(f2_fma(q, f2_pack(0.6931471805599453f, 0.6931471805599453f), lin), x0, x1);
}

// SwooshR on a packed pair (same formula as swoosh_r_fast); returns the packed result
ZV_DEVINL f32x2 swoosh_r_fast2(f32x2 x) {
  const float L2E = 1.4426950408889634f;
  float t0, t1;
  f2_unpack(f2_fma(x, f2_pack(L2E, L2E), f2_pack(-L2E, -L2E)), t0, t1);
  const f32x2 s = f2_add(f2_pack(ex2_approx(-fabsf(t0)), ex2_approx(-fabsf(t1))), f2_pack(1.0f, 1.0f));
  float s0, s1;
  f2_unpack(s, s0, s1);
  const f32x2 q = f2_add(f2_pack(fmaxf(t0, 0.f), fmaxf(t1, 0.f)), f2_pack(lg2_approx(s0), lg2_approx(s1)));
  const f32x2 lin = f2_fma(x, f2_pack(-0.08f, -0.08f), f2_pack(-0.313261687f, -0.313261687f));
  return f2_fma(q, f2_pack(0.6931471805599453f, 0.6931471805599453f), lin);
}

// SwooshL on a pair with ONE MUFU op per element: ln(1 + z), z = 2^-|t| in (0, 1], is a degree-5
// polynomial z*q(z) on the FMA pipe (max abs error 1.5e-6, far below the bf16 rounding of the result)
// instead of a second MUFU (lg2).  The GEMM epilogue is MUFU-bound with two per element.
ZV_DEVINL void swoosh_l_poly2(float& x0, float& x1) {
  const float L2E = 1.4426950408889634f;
  const f32x2 x = f2_pack(x0, x1);
  float t0, t1;
  f2_unpack(f2_fma(x, f2_pack(L2E, L2E), f2_pack(-4.0f * L2E, -4.0f * L2E)), t0, t1);
  const f32x2 z = f2_pack(ex2_approx(-fabsf(t0)), ex2_approx(-fabsf(t1)));
  f32x2 q = f2_fma(z, f2_pack(-0.01834004558622837f, -0.01834004558622837f),
                   f2_pack(0.08557301759719849f, 0.08557301759719849f));
  q = f2_fma(q, z, f2_pack(-0.19376389682292938f, -0.19376389682292938f));
  q = f2_fma(q, z, f2_pack(0.31765028834342957f, 0.31765028834342957f));
  q = f2_fma(q, z, f2_pack(-0.49787530303001404f, -0.49787530303001404f));
  q = f2_fma(q, z, f2_pack(0.9999016523361206f, 0.9999016523361206f));
  const f32x2 lin = f2_fma(x, f2_pack(-0.08f, -0.08f), f2_pack(-0.035f, -0.035f));
  const f32x2 u = f2_fma(f2_pack(fmaxf(t0, 0.f), fmaxf(t1, 0.f)), f2_pack(0.6931471805599453f, 0.6931471805599453f), lin);
  f2_unpack(f2_fma(z, q, u), x0, x1);
}

// softplus(x - off) = ln2 * (max(t, 0) + lg2(1 + 2^-|t|)), t = (x - off) log2(e): no overflow, so the
// reference's "inf -> x - off" guard (scaling.py:1176-1178) is satisfied by construction.
ZV_DEVINL float softplus_shift(float x, float off_log2e) {
  const float t = fmaf(x, 1.4426950408889634f, -off_log2e);
  return 0.6931471805599453f * (fmaxf(t, 0.f) + lg2_approx(1.0f + ex2_approx(-fabsf(t))));
}
ZV_DEVINL float swoosh_l_fast(float x) {
  return softplus_shift(x, 4.0f * 1.4426950408889634f) + fmaf(x, -0.08f, -0.035f);
}
ZV_DEVINL float swoosh_r_fast(float x) {
  return softplus_shift(x, 1.4426950408889634f) + fmaf(x, -0.08f, -0.313261687f);
}
ZV_DEVINL float sigmoid_fast(float x) { return rcp_approx(1.0f + ex2_approx(-1.4426950408889634f * x)); }

// ------------------------------------------------------------------ tcgen05 / TMEM
// One lane of a converged warp (always the same one for the full mask).  The single-thread tcgen05 / TMA instructions
// take their operands from UNIFORM registers: issued from `if (lane == 0)` code every operand is a per-lane value for
// the compiler and each instruction is preceded by seven R2UR broadcasts inside a waterfall loop (~90 cycles per
// tcgen05.mma, measured with ZV_TRACE); issued by an elected lane of warp-uniform code the operands stay in uniform
// registers.
ZV_DEVINL bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
ZV_DEVINL void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
ZV_DEVINL void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// Whole-warp: allocate `cols` TMEM columns (power of two >= 32); base address lands in *smem_out.
ZV_DEVINL void tmem_alloc(uint32_t* smem_out, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_out)),
               "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
ZV_DEVINL void tmem_dealloc(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}

// D[tmem] (+)= A[smem] * B[smem]^T, bf16 inputs, fp32 accumulate; issued by ONE thread.
ZV_DEVINL void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Make all previously issued MMAs arrive on an mbarrier when they complete (implies fence::before).
ZV_DEVINL void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// ------------------------------------------------------------------ CTA pairs (cluster of 2, cta_group::2)
ZV_DEVINL uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
ZV_DEVINL void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cta address of this CTA -> shared::cluster address of the same offset in CTA `rank`
ZV_DEVINL uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
// arrive on an mbarrier that lives in another CTA of the cluster (address from mapa_shared)
ZV_DEVINL void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// The same without release semantics: for hand-overs that carry no generic-proxy memory (a TMEM accumulator stage
// whose reads are ordered by tcgen05.wait::ld + tcgen05.fence::before_thread_sync).  The cluster-scope release of the
// form above costs ~1000 cycles per arrive inside a busy epilogue warp (measured, ZV_TRACE).
ZV_DEVINL void mbar_arrive_cluster_relaxed(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// 2D tiled load into THIS CTA's shared memory whose completion bytes are signalled on an mbarrier of
// either CTA of the pair (cluster address): both CTAs of a pair feed the leader's full barrier.
ZV_DEVINL void tma_load_2d_pair(void* smem_dst, const void* desc, uint32_t bar_cluster_addr, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], "
      "[%2];" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(desc)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
// TMEM allocation for a CTA pair: one warp of EACH CTA executes it with the same shared-memory offset.
ZV_DEVINL void tmem_alloc_pair(uint32_t* smem_out, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_out)),
               "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
ZV_DEVINL void tmem_dealloc_pair(uint32_t addr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
// D[tmem of both CTAs, 256 rows] (+)= A[128 rows per CTA] * B[N/2 rows per CTA]^T; issued by ONE thread of
// the leader CTA (rank 0); shared-memory descriptors are offsets valid in both CTAs.
ZV_DEVINL void umma_bf16_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// all previously issued pair MMAs arrive on the mbarrier at this offset in every CTA of `cta_mask`
ZV_DEVINL void umma_commit_pair(uint64_t* bar, uint16_t cta_mask) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(cta_mask)
      : "memory");
}

// 32 lanes x 16 consecutive fp32 columns -> 16 registers per thread (thread = TMEM lane).
ZV_DEVINL void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
// 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread
ZV_DEVINL void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
ZV_DEVINL void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Instruction descriptor for kind::f16, A/B = bf16 (K-major both), D = fp32, M x N tile
// (cute/arch/mma_sm100_desc.hpp InstrDescriptor: c_format [4,6), a_format [7,10),
//  b_format [10,13), a_major bit 15, b_major bit 16, n>>3 at [17,23), m>>4 at [24,29)).
__host__ __device__ inline uint32_t make_idesc_bf16(uint32_t M, uint32_t N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);
}
// Shared-memory matrix descriptor, K-major operand with 128-byte swizzle: rows are 128 B
// (64 bf16), 8-row swizzle atoms 1024 B apart (SBO); LBO unused for swizzled K-major;
// version = 1 (sm_100), layout_type = 2 (SWIZZLE_128B) at bits [61,64).
ZV_DEVINL uint64_t make_sdesc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr & 0x3FFFFu) >> 4);       // start address  [0,14)
  d |= static_cast<uint64_t>(1) << 16;                            // LBO (ignored)  [16,30)
  d |= static_cast<uint64_t>(1024 >> 4) << 32;                    // SBO            [32,46)
  d |= static_cast<uint64_t>(1) << 46;                            // version        [46,48)
  d |= static_cast<uint64_t>(2) << 61;                            // SWIZZLE_128B   [61,64)
  return d;
}

}  // namespace zv
