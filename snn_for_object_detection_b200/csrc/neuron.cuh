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
