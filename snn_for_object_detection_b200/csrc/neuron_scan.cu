// LIF / LI forward scans over T and the fused surrogate-gradient reverse scans.
//
// Reference arithmetic: norse 1.1.0 lif_feed_forward_step / li_feed_forward_step as restated in
// oracle/norse_restated.py (reference call sites models/modules/layer_gen.py:233,253).  The
// forward uses explicit round-to-nearest intrinsics op-for-op (no FMA contraction) so that for
// identical inputs the state and the spikes are bit-identical to the fp32 restatement.
//
// HBM-bound elementwise work: the state (v, i) lives in registers for all T steps; per
// neuron-step the forward moves 4 B in + 4 B out (+4 B per optional trace).  The time loop is
// unrolled so that kUnroll independent 16-byte loads are in flight per thread.
#include "common.cuh"
#include "neuron.cuh"

namespace soda {

constexpr int kUnroll = 4;

template <int V>
struct Vec;
template <>
struct Vec<4> {
  using T = float4;
};
template <>
struct Vec<1> {
  using T = float;
};

template <int V>
__device__ __forceinline__ void vload(float (&r)[V], const float* p) {
  if constexpr (V == 4) {
    const float4 t = __ldcs(reinterpret_cast<const float4*>(p));
    r[0] = t.x, r[1] = t.y, r[2] = t.z, r[3] = t.w;
  } else {
    r[0] = __ldcs(p);
  }
}
template <int V>
__device__ __forceinline__ void vstore(float* p, const float (&r)[V]) {
  if constexpr (V == 4) {
    *reinterpret_cast<float4*>(p) = make_float4(r[0], r[1], r[2], r[3]);
  } else {
    *p = r[0];
  }
}

// kMode: 0 = LIF, 1 = LI, 2 = LI + tanh, 3 = SLI (saturable leaky integrator)
template <int V, int kMode>
__global__ void __launch_bounds__(256)
neuron_fwd_kernel(NP p, const float* __restrict__ x, const float* __restrict__ v0,
                  const float* __restrict__ i0, int T, int64_t n, float* __restrict__ out,
                  float* __restrict__ v_out, float* __restrict__ i_out,
                  float* __restrict__ v_dec_trace, float* __restrict__ v_trace,
                  float* __restrict__ i_trace) {
  pdl_trigger();   // lets a dependent tensor-core conv start its prologue in this grid's tail
  const int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * V;
  if (e >= n) return;
  float v[V], i[V];
  if (v0 != nullptr) {
    vload<V>(v, v0 + e);
  } else {
#pragma unroll
    for (int k = 0; k < V; ++k) v[k] = p.v_leak;
  }
  if (i0 != nullptr) {
    vload<V>(i, i0 + e);
  } else {
#pragma unroll
    for (int k = 0; k < V; ++k) i[k] = 0.0f;
  }

  int t = 0;
  for (; t + kUnroll <= T; t += kUnroll) {
    float xs[kUnroll][V];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) vload<V>(xs[u], x + (int64_t)(t + u) * n + e);
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      float o[V], vd[V];
#pragma unroll
      for (int k = 0; k < V; ++k) {
        if constexpr (kMode == 0) {
          lif_step(p, xs[u][k], v[k], i[k], o[k], vd[k]);
        } else if constexpr (kMode == 3) {
          sli_step(p, xs[u][k], v[k], i[k]);
          o[k] = v[k];
        } else {
          li_step(p, xs[u][k], v[k], i[k]);
          o[k] = (kMode == 2) ? tanhf(v[k]) : v[k];
        }
      }
      const int64_t off = (int64_t)(t + u) * n + e;
      vstore<V>(out + off, o);
      if constexpr (kMode == 0) {
        if (v_dec_trace != nullptr) vstore<V>(v_dec_trace + off, vd);
      }
      if (v_trace != nullptr) vstore<V>(v_trace + off, v);
      if (i_trace != nullptr) vstore<V>(i_trace + off, i);
    }
  }
  for (; t < T; ++t) {
    float xs[V], o[V], vd[V];
    vload<V>(xs, x + (int64_t)t * n + e);
#pragma unroll
    for (int k = 0; k < V; ++k) {
      if constexpr (kMode == 0) {
        lif_step(p, xs[k], v[k], i[k], o[k], vd[k]);
      } else if constexpr (kMode == 3) {
        sli_step(p, xs[k], v[k], i[k]);
        o[k] = v[k];
      } else {
        li_step(p, xs[k], v[k], i[k]);
        o[k] = (kMode == 2) ? tanhf(v[k]) : v[k];
      }
    }
    const int64_t off = (int64_t)t * n + e;
    vstore<V>(out + off, o);
    if constexpr (kMode == 0) {
      if (v_dec_trace != nullptr) vstore<V>(v_dec_trace + off, vd);
    }
    if (v_trace != nullptr) vstore<V>(v_trace + off, v);
    if (i_trace != nullptr) vstore<V>(i_trace + off, i);
  }
  vstore<V>(v_out + e, v);
  vstore<V>(i_out + e, i);
}

// Reverse scan.  kLif: uses v_dec_trace, surrogate gradient and the (non-detached) reset path.
template <int V, bool kLif>
__global__ void __launch_bounds__(256)
neuron_bwd_kernel(NP p, const float* __restrict__ gout, const float* __restrict__ gvT,
                  const float* __restrict__ giT, const float* __restrict__ v_dec_trace, int T,
                  int64_t n, float* __restrict__ gx, float* __restrict__ gv0,
                  float* __restrict__ gi0, int C, int64_t g_ps) {
  pdl_trigger();   // lets a dependent tensor-core conv start its prologue in this grid's tail
  const int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * V;
  if (e >= n) return;
  // grad_out may be a channel slice of a wider pixel-major tensor (the gradient of a Dense concat): element e of step t
  // is channel e % C of pixel t * n / C + e / C, pixels g_ps floats apart.  g_ps == 0: dense [T, n].
  const int64_t g_step = g_ps ? (n / C) * g_ps : n;
  const int64_t g_e = g_ps ? (e / C) * g_ps + (e % C) : e;
  float gv[V], gi[V];
  if (gvT != nullptr) {
    vload<V>(gv, gvT + e);
  } else {
#pragma unroll
    for (int k = 0; k < V; ++k) gv[k] = 0.0f;
  }
  if (giT != nullptr) {
    vload<V>(gi, giT + e);
  } else {
#pragma unroll
    for (int k = 0; k < V; ++k) gi[k] = 0.0f;
  }
  const float one_m_a = 1.0f - p.a, one_m_s = 1.0f - p.s;

  for (int t0 = T; t0 > 0; t0 -= kUnroll) {
    const int cnt = t0 >= kUnroll ? kUnroll : t0;
    float go[kUnroll][V], vd[kUnroll][V];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      if (u < cnt) {
        const int64_t off = (int64_t)(t0 - 1 - u) * n + e;
        vload<V>(go[u], gout + (int64_t)(t0 - 1 - u) * g_step + g_e);
        if constexpr (kLif) vload<V>(vd[u], v_dec_trace + off);
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      if (u < cnt) {
        float g[V];
#pragma unroll
        for (int k = 0; k < V; ++k) {
          float gvd;
          if constexpr (kLif) {
            const float u_th = vd[u][k] - p.v_th;
            const float z = u_th > 0.0f ? 1.0f : 0.0f;
            float gz = go[u][k];
            if (!p.detach_reset) gz += gv[k] * (p.v_reset - vd[u][k]);
            const float den = p.alpha * fabsf(u_th) + 1.0f;
            gvd = gv[k] * (1.0f - z) + gz / (den * den);
          } else {
            gvd = go[u][k] + gv[k];
          }
          gv[k] = gvd * one_m_a;
          if (p.jump_first) {
            gi[k] = gvd * p.a + gi[k] * one_m_s;
            g[k] = gi[k];
          } else {
            g[k] = gi[k];
            gi[k] = gvd * p.a + gi[k] * one_m_s;
          }
        }
        vstore<V>(gx + (int64_t)(t0 - 1 - u) * n + e, g);
      }
    }
  }
  if (gv0 != nullptr) vstore<V>(gv0 + e, gv);
  if (gi0 != nullptr) vstore<V>(gi0 + e, gi);
}

// LIF scan whose output goes through a pixel stride (a channel slice of a Dense concat buffer) with
// an optional residual add: out = z (+ residual).  One pass instead of scan + add_n + slice copy
// (Residual sum models/generator.py:146, Dense cat :158).  x and the state are dense [T][P][C].
__global__ void __launch_bounds__(256)
lif_res_fwd_kernel(NP p, const float* __restrict__ x, const float* __restrict__ v0,
                   const float* __restrict__ i0, int T, int64_t P, int C,
                   const float* __restrict__ res, int64_t res_ps, float* __restrict__ out, int64_t out_ps,
                   float* __restrict__ v_out, float* __restrict__ i_out) {
  pdl_trigger();   // lets a dependent tensor-core conv start its prologue in this grid's tail
  const int64_t n = P * C;
  const int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (e >= n) return;
  const int64_t q = e / C;
  const int c = (int)(e - q * C);
  float v[4], i[4];
  if (v0 != nullptr) {
    vload<4>(v, v0 + e);
    vload<4>(i, i0 + e);
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = p.v_leak, i[k] = 0.0f;
  }
  for (int t0 = 0; t0 < T; t0 += kUnroll) {
    float xs[kUnroll][4], rs[kUnroll][4];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      if (t0 + u < T) {
        vload<4>(xs[u], x + (int64_t)(t0 + u) * n + e);
        if (res != nullptr) {
          const float4 r4 = __ldg(reinterpret_cast<const float4*>(res + ((int64_t)(t0 + u) * P + q) * res_ps + c));
          rs[u][0] = r4.x, rs[u][1] = r4.y, rs[u][2] = r4.z, rs[u][3] = r4.w;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      if (t0 + u < T) {
        float o[4], vd;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          lif_step(p, xs[u][k], v[k], i[k], o[k], vd);
          if (res != nullptr) o[k] = __fadd_rn(o[k], rs[u][k]);
        }
        vstore<4>(out + ((int64_t)(t0 + u) * P + q) * out_ps + c, o);
      }
    }
  }
  vstore<4>(v_out + e, v);
  vstore<4>(i_out + e, i);
}

// LIF scan with a fused non-overlapping 2x2 pooling of the spikes (generator block Conv/Norm/LIF/Pool,
// reference models/modules/layer_gen.py:139-173 + common.py:45-49): one thread owns the 4 pixels of a pooling
// window (x 4 channels) for all T steps - membrane state of the 16 neurons in registers - and stores only the
// pooled value, so the full-resolution spike map never exists in HBM.  Pixels of an odd last row / column belong
// to no window (floor mode) but still integrate: their state is part of v_out / i_out.
// mode: 0 = sum (avg * 2 * 2 like the reference), 1 = avg, 2 = max.  Same op order as pool2d_kernel.
__global__ void __launch_bounds__(256)
lif_pool2_fwd_kernel(NP p, const float* __restrict__ x, const float* __restrict__ v0, const float* __restrict__ i0,
                     int T, int B, int H, int W, int C, int mode, float* __restrict__ out, int64_t out_ps,
                     float* __restrict__ v_out, float* __restrict__ i_out) {
  pdl_trigger();   // lets a dependent tensor-core conv start its prologue in this grid's tail
  const int cv = C / 4;
  const int Hw = (H + 1) / 2, Ww = (W + 1) / 2, Ho = H / 2, Wo = W / 2;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (int64_t)B * Hw * Ww * cv) return;
  const int c = (int)(idx % cv) * 4;
  int64_t r = idx / cv;
  const int wx = (int)(r % Ww);
  r /= Ww;
  const int wy = (int)(r % Hw);
  const int b = (int)(r / Hw);
  const bool full = wy < Ho && wx < Wo;
  bool have[4];
  int64_t poff[4];   // element offset of the window's pixels inside one frame
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int iy = wy * 2 + (q >> 1), ix = wx * 2 + (q & 1);
    have[q] = iy < H && ix < W;
    poff[q] = ((int64_t)iy * W + ix) * C + c;
  }
  const int64_t frame = (int64_t)H * W * C;
  float v[4][4], cur[4][4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
#pragma unroll
    for (int k = 0; k < 4; ++k) v[q][k] = p.v_leak, cur[q][k] = 0.0f;
    if (have[q] && v0 != nullptr) {
      vload<4>(v[q], v0 + (int64_t)b * frame + poff[q]);
      vload<4>(cur[q], i0 + (int64_t)b * frame + poff[q]);
    }
  }
  for (int t = 0; t < T; ++t) {
    const float* xf = x + ((int64_t)t * B + b) * frame;
    float xs[4][4];
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (have[q]) vload<4>(xs[q], xf + poff[q]);
    float acc[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) acc[k] = (mode == 2) ? -3.402823466e+38f : 0.0f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      if (have[q]) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float z, vd;
          lif_step(p, xs[q][k], v[q][k], cur[q][k], z, vd);
          acc[k] = (mode == 2) ? fmaxf(acc[k], z) : __fadd_rn(acc[k], z);
        }
      }
    }
    if (full) {
      float o[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (mode == 2) o[k] = acc[k];
        else if (mode == 1) o[k] = __fmul_rn(acc[k], 0.25f);
        else o[k] = __fmul_rn(__fmul_rn(__fmul_rn(acc[k], 0.25f), 2.0f), 2.0f);
      }
      vstore<4>(out + ((((int64_t)t * B + b) * Ho + wy) * Wo + wx) * out_ps + c, o);
    }
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    if (have[q]) {
      vstore<4>(v_out + (int64_t)b * frame + poff[q], v[q]);
      vstore<4>(i_out + (int64_t)b * frame + poff[q], cur[q]);
    }
  }
}

static bool all_aligned16(std::initializer_list<const void*> ptrs) {
  for (const void* q : ptrs)
    if (q != nullptr && (reinterpret_cast<uintptr_t>(q) & 15) != 0) return false;
  return true;
}

template <int kMode>
static int launch_fwd(const soda_neuron_params_t* q, const float* x, const float* v0,
                      const float* i0, int T, int64_t n, float* out, float* v_out, float* i_out,
                      float* v_dec_trace, float* v_trace, float* i_trace, void* stream,
                      const char* name) {
  if (!q || !x || !out || !v_out || !i_out) return fail(SODA_ERR_BAD_ARG, "neuron fwd: null pointer");
  if (T <= 0 || n <= 0) return fail(SODA_ERR_BAD_ARG, "neuron fwd: non-positive size");
  const NP p = to_np(q);
  cudaStream_t st = (cudaStream_t)stream;
  const bool vec = (n % 4 == 0) &&
                   all_aligned16({x, v0, i0, out, v_out, i_out, v_dec_trace, v_trace, i_trace});
  if (vec) {
    const int64_t blocks = ceil_div(n / 4, 256);
    neuron_fwd_kernel<4, kMode><<<(unsigned)blocks, 256, 0, st>>>(
        p, x, v0, i0, T, n, out, v_out, i_out, v_dec_trace, v_trace, i_trace);
  } else {
    const int64_t blocks = ceil_div(n, 256);
    neuron_fwd_kernel<1, kMode><<<(unsigned)blocks, 256, 0, st>>>(
        p, x, v0, i0, T, n, out, v_out, i_out, v_dec_trace, v_trace, i_trace);
  }
  return after_launch(name);
}

template <bool kLif>
static int launch_bwd(const soda_neuron_params_t* q, const float* gout, const float* gvT,
                      const float* giT, const float* v_dec_trace, int T, int64_t n, float* gx,
                      float* gv0, float* gi0, void* stream, const char* name, int C = 0, int64_t g_ps = 0) {
  if (!q || !gout || !gx || (kLif && !v_dec_trace))
    return fail(SODA_ERR_BAD_ARG, "neuron bwd: null pointer");
  if (T <= 0 || n <= 0) return fail(SODA_ERR_BAD_ARG, "neuron bwd: non-positive size");
  if (g_ps != 0 && (C <= 0 || n % C != 0 || g_ps < C))
    return fail(SODA_ERR_BAD_ARG, "neuron bwd: strided grad_out needs C > 0, n % C == 0 and pixel stride >= C");
  if (g_ps == C) g_ps = 0;   // dense after all
  const NP p = to_np(q);
  cudaStream_t st = (cudaStream_t)stream;
  const bool vec = (n % 4 == 0) && all_aligned16({gout, gvT, giT, v_dec_trace, gx, gv0, gi0}) &&
                   (g_ps == 0 || (C % 4 == 0 && g_ps % 4 == 0));
  if (vec) {
    neuron_bwd_kernel<4, kLif><<<(unsigned)ceil_div(n / 4, 256), 256, 0, st>>>(
        p, gout, gvT, giT, v_dec_trace, T, n, gx, gv0, gi0, C, g_ps);
  } else {
    neuron_bwd_kernel<1, kLif><<<(unsigned)ceil_div(n, 256), 256, 0, st>>>(
        p, gout, gvT, giT, v_dec_trace, T, n, gx, gv0, gi0, C, g_ps);
  }
  return after_launch(name);
}

}  // namespace soda

extern "C" int soda_lif_fwd(const soda_neuron_params_t* p, const float* x, const float* v0,
                            const float* i0, int T, int64_t n, float* out, float* v_out,
                            float* i_out, float* v_dec_trace, float* v_trace, float* i_trace,
                            void* stream) {
  return soda::launch_fwd<0>(p, x, v0, i0, T, n, out, v_out, i_out, v_dec_trace, v_trace, i_trace,
                             stream, "lif_fwd");
}

extern "C" int soda_lif_res_fwd(const soda_neuron_params_t* q, const float* x, const float* v0,
                                const float* i0, int T, int64_t n_pix, int C, const float* residual,
                                int64_t res_pix_stride, float* out, int64_t out_pix_stride,
                                float* v_out, float* i_out, void* stream) {
  using namespace soda;
  if (!q || !x || !out || !v_out || !i_out) return fail(SODA_ERR_BAD_ARG, "lif_res_fwd: null pointer");
  if (T <= 0 || n_pix <= 0 || C <= 0) return fail(SODA_ERR_BAD_ARG, "lif_res_fwd: non-positive size");
  if ((v0 == nullptr) != (i0 == nullptr)) return fail(SODA_ERR_BAD_ARG, "lif_res_fwd: v0 / i0 must come together");
  if (C % 4 != 0 || out_pix_stride % 4 != 0 || (residual != nullptr && res_pix_stride % 4 != 0) ||
      !all_aligned16({x, v0, i0, out, v_out, i_out, residual}))
    return fail(SODA_ERR_UNSUPPORTED, "lif_res_fwd: needs C % 4 == 0, strides % 4 == 0 and 16-byte aligned buffers");
  const int64_t blocks = ceil_div(n_pix * C / 4, 256);
  lif_res_fwd_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      to_np(q), x, v0, i0, T, n_pix, C, residual, res_pix_stride, out, out_pix_stride, v_out, i_out);
  return after_launch("lif_res_fwd");
}

extern "C" int soda_lif_pool_fwd(const soda_neuron_params_t* q, const float* x, const float* v0, const float* i0,
                                 int T, int B, int H, int W, int C, int mode, float* out, int64_t out_pix_stride,
                                 float* v_out, float* i_out, void* stream) {
  using namespace soda;
  if (!q || !x || !out || !v_out || !i_out) return fail(SODA_ERR_BAD_ARG, "lif_pool_fwd: null pointer");
  if (T <= 0 || B <= 0 || H < 2 || W < 2 || C <= 0 || mode < 0 || mode > 2)
    return fail(SODA_ERR_BAD_ARG, "lif_pool_fwd: bad size or mode");
  if ((v0 == nullptr) != (i0 == nullptr)) return fail(SODA_ERR_BAD_ARG, "lif_pool_fwd: v0 / i0 must come together");
  if (C % 4 != 0 || out_pix_stride % 4 != 0 || !all_aligned16({x, v0, i0, out, v_out, i_out}))
    return fail(SODA_ERR_UNSUPPORTED, "lif_pool_fwd: needs C % 4 == 0, stride % 4 == 0 and 16-byte aligned buffers");
  const int64_t threads = (int64_t)B * ((H + 1) / 2) * ((W + 1) / 2) * (C / 4);
  lif_pool2_fwd_kernel<<<(unsigned)ceil_div(threads, 256), 256, 0, (cudaStream_t)stream>>>(
      to_np(q), x, v0, i0, T, B, H, W, C, mode, out, out_pix_stride, v_out, i_out);
  return after_launch("lif_pool_fwd");
}

extern "C" int soda_li_fwd(const soda_neuron_params_t* p, const float* x, const float* v0,
                           const float* i0, int T, int64_t n, int apply_tanh, float* out,
                           float* v_out, float* i_out, float* i_trace, void* stream) {
  if (apply_tanh)
    return soda::launch_fwd<2>(p, x, v0, i0, T, n, out, v_out, i_out, nullptr, nullptr, i_trace,
                               stream, "li_tanh_fwd");
  return soda::launch_fwd<1>(p, x, v0, i0, T, n, out, v_out, i_out, nullptr, nullptr, i_trace,
                             stream, "li_fwd");
}

namespace soda {

// Synaptic transmission model of the reference's own models/modules/synapse.py:73-103 (synapse_feed_forward_step):
//   tau = x > 0 ? tau_secretion : tau_dissociation;  p' = p + ((x - p) * tau) * dt
//   g = sigma != 0 ? (4 * sigma) * (p' - sigma * p'^2) : p';  out = max(g, 0)          one state (p) per synapse
template <int V>
__global__ void __launch_bounds__(256)
synapse_fwd_kernel(float tau_sec, float tau_dis, float sigma, float dt, const float* __restrict__ x,
                   const float* __restrict__ p0, int T, int64_t n, float* __restrict__ out, float* __restrict__ p_out) {
  pdl_trigger();
  const int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * V;
  if (e >= n) return;
  float p[V];
  if (p0 != nullptr) {
    vload<V>(p, p0 + e);
  } else {
#pragma unroll
    for (int k = 0; k < V; ++k) p[k] = 0.0f;
  }
  const float s4 = __fmul_rn(4.0f, sigma);
  for (int t = 0; t < T; ++t) {
    float xs[V], o[V];
    vload<V>(xs, x + (int64_t)t * n + e);
#pragma unroll
    for (int k = 0; k < V; ++k) {
      const float tau = xs[k] > 0.0f ? tau_sec : tau_dis;
      p[k] = __fadd_rn(p[k], __fmul_rn(__fmul_rn(__fsub_rn(xs[k], p[k]), tau), dt));
      const float g = sigma != 0.0f ? __fmul_rn(s4, __fsub_rn(p[k], __fmul_rn(sigma, __fmul_rn(p[k], p[k])))) : p[k];
      o[k] = fmaxf(g, 0.0f);
    }
    vstore<V>(out + (int64_t)t * n + e, o);
  }
  vstore<V>(p_out + e, p);
}

}  // namespace soda

extern "C" int soda_synapse_fwd(float tau_secretion, float tau_dissociation, float sigma_inhibition, float dt,
                                const float* x, const float* p0, int T, int64_t n, float* out, float* p_out,
                                void* stream) {
  using namespace soda;
  if (!x || !out || !p_out) return fail(SODA_ERR_BAD_ARG, "synapse_fwd: null pointer");
  if (T <= 0 || n <= 0) return fail(SODA_ERR_BAD_ARG, "synapse_fwd: non-positive size");
  cudaStream_t st = (cudaStream_t)stream;
  if (n % 4 == 0 && all_aligned16({x, p0, out, p_out})) {
    synapse_fwd_kernel<4><<<(unsigned)ceil_div(n / 4, 256), 256, 0, st>>>(tau_secretion, tau_dissociation,
                                                                         sigma_inhibition, dt, x, p0, T, n, out, p_out);
  } else {
    synapse_fwd_kernel<1><<<(unsigned)ceil_div(n, 256), 256, 0, st>>>(tau_secretion, tau_dissociation, sigma_inhibition,
                                                                     dt, x, p0, T, n, out, p_out);
  }
  return after_launch("synapse_fwd");
}

extern "C" int soda_sli_fwd(const soda_neuron_params_t* p, const float* x, const float* v0, const float* i0, int T,
                            int64_t n, float* out, float* v_out, float* i_out, float* i_trace, void* stream) {
  return soda::launch_fwd<3>(p, x, v0, i0, T, n, out, v_out, i_out, nullptr, nullptr, i_trace, stream, "sli_fwd");
}

extern "C" int soda_lif_bwd(const soda_neuron_params_t* p, const float* grad_out,
                            const float* grad_v_T, const float* grad_i_T, const float* v_dec_trace,
                            int T, int64_t n, float* grad_x, float* grad_v0, float* grad_i0,
                            void* stream) {
  return soda::launch_bwd<true>(p, grad_out, grad_v_T, grad_i_T, v_dec_trace, T, n, grad_x,
                                grad_v0, grad_i0, stream, "lif_bwd");
}

extern "C" int soda_lif_bwd_strided(const soda_neuron_params_t* p, const float* grad_out, int C,
                                    int64_t grad_out_pix_stride, const float* grad_v_T, const float* grad_i_T,
                                    const float* v_dec_trace, int T, int64_t n, float* grad_x, float* grad_v0,
                                    float* grad_i0, void* stream) {
  if (C <= 0 || grad_out_pix_stride < C) return soda::fail(SODA_ERR_BAD_ARG, "lif_bwd_strided: bad C / pixel stride");
  return soda::launch_bwd<true>(p, grad_out, grad_v_T, grad_i_T, v_dec_trace, T, n, grad_x, grad_v0, grad_i0,
                                stream, "lif_bwd_strided", C, grad_out_pix_stride);
}

extern "C" int soda_li_bwd(const soda_neuron_params_t* p, const float* grad_out,
                           const float* grad_v_T, const float* grad_i_T, int T, int64_t n,
                           float* grad_x, float* grad_v0, float* grad_i0, void* stream) {
  return soda::launch_bwd<false>(p, grad_out, grad_v_T, grad_i_T, nullptr, T, n, grad_x, grad_v0,
                                 grad_i0, stream, "li_bwd");
}
