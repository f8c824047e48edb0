// LIF / LI step arithmetic shared by the stand-alone scans (neuron_scan.cu) and the fused
// convolution epilogues (conv_tc3.cu, conv_direct_lif.cu).  Op-for-op the fp32 restatement of norse 1.1.0's
// lif_feed_forward_step / li_feed_forward_step (oracle/norse_restated.py; reference call sites
// models/modules/layer_gen.py:233,253): explicit round-to-nearest intrinsics, no FMA contraction.
#pragma once

#include "../../include/soda_b200.h"

namespace soda {

struct NP {
  float a, s, v_leak, v_th, v_reset, alpha;
  int jump_first, detach_reset;
};

__device__ __forceinline__ void lif_step(const NP& p, float x, float& v, float& i, float& z,
                                         float& v_dec) {
  float i_src;
  if (p.jump_first) {
    i_src = __fadd_rn(i, x);  // i_new = i + x
  } else {
    i_src = i;
  }
  const float dv = __fmul_rn(p.a, __fadd_rn(__fsub_rn(p.v_leak, v), i_src));
  v_dec = __fadd_rn(v, dv);
  const float di = __fmul_rn(-p.s, i_src);
  float i_dec = __fadd_rn(i_src, di);
  if (!p.jump_first) i_dec = __fadd_rn(i_dec, x);
  z = (__fsub_rn(v_dec, p.v_th) > 0.0f) ? 1.0f : 0.0f;
  // (1 - z) * v_dec + z * v_reset with z in {0, 1}: one of the two products is an exact zero, so the sum IS
  // the other operand (up to the sign of zero) - a select gives the same fp32 value as the four-op form.
  v = (z != 0.0f) ? p.v_reset : v_dec;
  i = i_dec;
}

// ---- two neurons per instruction: Blackwell's packed fp32 pipe (SASS FADD2 / FMUL2 / FFMA2) -----------------
// add/sub.rn.f32x2 are two IEEE fp32 operations, each rounded on its own: lif_step2 is bit-identical to two lif_step
// calls (scripts/microbench/f32x2_lif.cu checks it on the GPU).  The two PRODUCTS stay scalar on purpose: ptxas
// (12.9) contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 - even with --fmad=false - which rounds once instead of
// twice (measured: 40 % of membranes off by an ulp after 32 steps).  The fp32 pipe issues one 3-register instruction
// per two clocks and SMSP; where a kernel keeps only a few warps resident and does its neuron arithmetic between
// tensor-core stages (conv_tc4's fused LIF epilogue), that pipe is the bound.
typedef unsigned long long f32x2_t;
__device__ __forceinline__ f32x2_t pk2(float lo, float hi) {
  f32x2_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk2(f32x2_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2_t add2(f32x2_t a, f32x2_t b) {
  f32x2_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2_t sub2(f32x2_t a, f32x2_t b) {
  f32x2_t r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2_t mul2(f32x2_t a, f32x2_t b) {
  f32x2_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f32x2_t fma2(f32x2_t a, f32x2_t b, f32x2_t c) {
  f32x2_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}

struct NP2 {   // the constants of lif_step, duplicated into both halves once per kernel
  f32x2_t v_leak, v_th;
  float a, ms, v_reset;
  int jump_first;
};
__device__ __forceinline__ NP2 np_pairs(const NP& p) {
  NP2 q;
  q.v_leak = pk2(p.v_leak, p.v_leak), q.v_th = pk2(p.v_th, p.v_th);
  q.a = p.a, q.ms = -p.s, q.v_reset = p.v_reset, q.jump_first = p.jump_first;
  return q;
}
// x, v, i: two neurons; z0 / z1: their spikes as 0.0f / 1.0f
__device__ __forceinline__ void lif_step2(const NP2& p, f32x2_t x, f32x2_t& v, f32x2_t& i, float& z0, float& z1) {
  const f32x2_t i_src = p.jump_first ? add2(i, x) : i;
  float u0, u1, s0, s1;
  upk2(add2(sub2(p.v_leak, v), i_src), u0, u1);
  upk2(i_src, s0, s1);
  const f32x2_t v_dec = add2(v, pk2(__fmul_rn(p.a, u0), __fmul_rn(p.a, u1)));
  f32x2_t i_dec = add2(i_src, pk2(__fmul_rn(p.ms, s0), __fmul_rn(p.ms, s1)));
  if (!p.jump_first) i_dec = add2(i_dec, x);
  float d0, d1, vd0, vd1;
  upk2(sub2(v_dec, p.v_th), d0, d1);
  upk2(v_dec, vd0, vd1);
  z0 = d0 > 0.0f ? 1.0f : 0.0f;
  z1 = d1 > 0.0f ? 1.0f : 0.0f;
  v = pk2(d0 > 0.0f ? p.v_reset : vd0, d1 > 0.0f ? p.v_reset : vd1);
  i = i_dec;
}

__device__ __forceinline__ void li_step(const NP& p, float x, float& v, float& i) {
  const float i_src = p.jump_first ? __fadd_rn(i, x) : i;
  const float dv = __fmul_rn(p.a, __fadd_rn(__fsub_rn(p.v_leak, v), i_src));
  v = __fadd_rn(v, dv);
  const float di = __fmul_rn(-p.s, i_src);
  float i_dec = __fadd_rn(i_src, di);
  if (!p.jump_first) i_dec = __fadd_rn(i_dec, x);
  i = i_dec;
}

// Saturable leaky integrator of the reference's own models/modules/sli.py:110-126 (sli_feed_forward_step):
//   i_jump = i + x * sigmoid(v_st - |v|);  v' = v + dt*tau_mem*((v_leak - v) + i_jump);  i' = i_jump - dt*tau_syn*i_jump
// v_st travels in the v_th field.  sigmoid(z) = 1 / (1 + exp(-z)) as ATen evaluates it.
__device__ __forceinline__ void sli_step(const NP& p, float x, float& v, float& i) {
  const float z = __fsub_rn(p.v_th, fabsf(v));
  const float gate = __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-z)));
  const float i_jump = __fadd_rn(i, __fmul_rn(x, gate));
  const float dv = __fmul_rn(p.a, __fadd_rn(__fsub_rn(p.v_leak, v), i_jump));
  v = __fadd_rn(v, dv);
  i = __fadd_rn(i_jump, __fmul_rn(-p.s, i_jump));
}

inline NP to_np(const soda_neuron_params_t* q) {
  NP p;
  p.a = q->dt_tau_mem, p.s = q->dt_tau_syn, p.v_leak = q->v_leak, p.v_th = q->v_th;
  p.v_reset = q->v_reset, p.alpha = q->alpha, p.jump_first = q->jump_first;
  p.detach_reset = q->detach_reset;
  return p;
}

}  // namespace soda
