// Direct small-K convolution + folded BatchNorm + LIF over T time steps, state in registers
// (soda_conv_lif_fwd, kind SODA_CONV_FP32).  This is the detector's FIRST layer: the binned event
// frames (plain NCHW, Cin = 2 polarities) -> Conv(64, 3, stride 2) -> Norm -> LIF
// (reference models/tiny_yolo.py:16,53-58; nn.Conv2d models/modules/layer_gen.py:129-136, eval
// BatchNorm :211-214, norse LIFCell :233, run once per time step by models/generator.py:184-198).
//
// K = Cin*k*k = 18: far too small for an MMA tile, and the layer is HBM-bound on its OUTPUT
// (64 fp32 spikes per pixel-step; the 93 %-zero input frames stay L2-resident), so it is a plain
// FP32-pipe kernel: one thread owns CPT output channels of one output pixel of one sample for ALL T
// steps; membrane potential and synaptic current never leave its registers.  Weights (BN scale
// folded, [K][Cout_pad] = conv_ffma's packed layout) and the BN shift sit in shared memory.
// Accumulation is a sequential fmaf chain in k = (ky*k + kx)*Cin + ci order - the order of
// conv_ffma_kernel - so results are bit-identical to soda_conv2d_fwd(FP32) + soda_lif_fwd.
#include <cuda_fp16.h>

#include "common.cuh"
#include "neuron.cuh"

namespace soda {

struct DirectP {
  int T, B, H, W, Cin, Cout, Cout_pad, k, stride, pad, Ho, Wo, Ktot;
  int64_t out_bs, out_ps;   // BYTES
  int out_fmt;              // SODA_ACT_F32 / PAIR / HALF (spikes are exact in all of them)
  NP np;
};

// byte offset of channel c (c % 4 == 0) inside a pixel of the output format
__device__ __forceinline__ int64_t direct_ch_off(int fmt, int c) {
  if (fmt == SODA_ACT_F32) return (int64_t)c * 4;
  if (fmt == SODA_ACT_HALF) return (int64_t)c * 2;
  return (int64_t)(c & ~31) * 4 + (int64_t)(c & 31) * 2;
}
__device__ __forceinline__ void direct_store4(uint8_t* dst, int fmt, float a, float b, float c, float d) {
  if (fmt == SODA_ACT_F32) {
    __stcs(reinterpret_cast<float4*>(dst), make_float4(a, b, c, d));
    return;
  }
  const __half2 h0 = __floats2half2_rn(a, b), h1 = __floats2half2_rn(c, d);
  uint2 u;
  u.x = *reinterpret_cast<const uint32_t*>(&h0), u.y = *reinterpret_cast<const uint32_t*>(&h1);
  __stcs(reinterpret_cast<uint2*>(dst), u);
  if (fmt == SODA_ACT_PAIR) __stcs(reinterpret_cast<uint2*>(dst + 64), make_uint2(0u, 0u));   // lo halves: exact zero
}

template <int CPT>
__device__ __forceinline__ void direct_neuron_store(const NP& np, const float (&acc)[CPT], const float (&sh)[CPT],
                                                    float (&v)[CPT], float (&cur)[CPT], uint8_t* __restrict__ dst,
                                                    int64_t off_hi, int fmt) {
  float z[CPT];
#pragma unroll
  for (int c = 0; c < CPT; ++c) {
    float vd;
    lif_step(np, __fadd_rn(acc[c], sh[c]), v[c], cur[c], z[c], vd);
  }
#pragma unroll
  for (int c = 0; c < CPT; c += 4)
    direct_store4(dst + (c ? off_hi : 0), fmt, z[c], z[c + 1], z[c + 2], z[c + 3]);
}

// KS / CIN > 0: compile-time filter size and input channels (fully unrolled taps, input offsets
// hoisted out of the time loop, next step's inputs prefetched while this step is computed);
// KS = 0: run-time sizes.
template <int CPT, int KS, int CIN>
__global__ void __launch_bounds__(256, 2)
conv_direct_lif_kernel(DirectP p, const float* __restrict__ x, const float* __restrict__ wp,
                       const float* __restrict__ shift, const float* __restrict__ v0,
                       const float* __restrict__ i0, uint8_t* __restrict__ y, float* __restrict__ v_out,
                       float* __restrict__ i_out) {
  pdl_trigger();   // lets a dependent tensor-core conv start its prologue in this grid's tail
  extern __shared__ float smem_w[];  // [Ktot][Cout] weights, then [Cout] shift
  float* smem_shift = smem_w + p.Ktot * p.Cout;
  for (int e = threadIdx.x; e < p.Ktot * p.Cout; e += blockDim.x) {
    const int kk = e / p.Cout, co = e - kk * p.Cout;
    smem_w[e] = wp[(int64_t)kk * p.Cout_pad + co];
  }
  for (int e = threadIdx.x; e < p.Cout; e += blockDim.x) smem_shift[e] = shift ? shift[e] : 0.0f;
  __syncthreads();

  const int G = p.Cout / CPT;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int g = (int)(gid % G);
  const int64_t pix = gid / G;
  const int64_t hw = (int64_t)p.Ho * p.Wo;
  if (pix >= (int64_t)p.B * hw) return;
  const unsigned amask = __activemask();
  const int b = (int)(pix / hw);
  const int rr = (int)(pix - (int64_t)b * hw);
  const int oy = rr / p.Wo, ox = rr - oy * p.Wo;
  // Thread g owns channels [4g, 4g+4) and [Cout/2 + 4g, Cout/2 + 4g + 4): the 8 lanes of a pixel then read
  // 128 contiguous bytes of a weight row per LDS.128 (no bank conflicts) and store 128 contiguous bytes.
  static_assert(CPT == 8, "channel mapping below is for 8 channels per thread");
  const int c0 = g * 4;
  const int chalf = p.Cout / 2;
  const int iy0 = oy * p.stride - p.pad, ix0 = ox * p.stride - p.pad;

  float v[CPT], cur[CPT];
  const int64_t soff = pix * p.Cout + c0;
#pragma unroll
  for (int c = 0; c < CPT; c += 4) {
    float4 a = make_float4(p.np.v_leak, p.np.v_leak, p.np.v_leak, p.np.v_leak);
    float4 bq = make_float4(0.f, 0.f, 0.f, 0.f);
    if (v0 != nullptr) {
      a = __ldg(reinterpret_cast<const float4*>(v0 + soff + (c ? chalf : 0)));
      bq = __ldg(reinterpret_cast<const float4*>(i0 + soff + (c ? chalf : 0)));
    }
    v[c] = a.x, v[c + 1] = a.y, v[c + 2] = a.z, v[c + 3] = a.w;
    cur[c] = bq.x, cur[c + 1] = bq.y, cur[c + 2] = bq.z, cur[c + 3] = bq.w;
  }
  float sh[CPT];
#pragma unroll
  for (int c = 0; c < CPT; ++c) sh[c] = smem_shift[c0 + (c < 4 ? c : chalf + c - 4)];

  const int64_t frame_elems = (int64_t)p.Cin * p.H * p.W;
  const int64_t step_elems = frame_elems * p.B;
  const float* xb = x + (int64_t)b * frame_elems;          // frame (t, b) = xb + t * step_elems
  uint8_t* yb = y + (int64_t)b * p.out_bs + (int64_t)rr * p.out_ps + direct_ch_off(p.out_fmt, c0);
  const int64_t y_step = (int64_t)p.B * p.out_bs;
  const int64_t off_hi = direct_ch_off(p.out_fmt, chalf + c0) - direct_ch_off(p.out_fmt, c0);

  if constexpr (KS > 0) {
    constexpr int NT = KS * KS * CIN;
    int off[NT];   // element offset of tap kk inside a frame, or -1 = zero padding
#pragma unroll
    for (int ky = 0; ky < KS; ++ky)
#pragma unroll
      for (int kx = 0; kx < KS; ++kx)
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci) {
          const int iy = iy0 + ky, ix = ix0 + kx;
          const bool inb = iy >= 0 && iy < p.H && ix >= 0 && ix < p.W;
          off[(ky * KS + kx) * CIN + ci] = inb ? (ci * p.H + iy) * p.W + ix : -1;
        }
    const uint32_t w_base = (uint32_t)__cvta_generic_to_shared(smem_w + c0);
    float xv[NT], xn[NT];
#pragma unroll
    for (int kk = 0; kk < NT; ++kk) xn[kk] = off[kk] >= 0 ? __ldg(xb + off[kk]) : 0.0f;
#pragma unroll 1
    for (int t = 0; t < p.T; ++t) {
#pragma unroll
      for (int kk = 0; kk < NT; ++kk) xv[kk] = xn[kk];
      if (t + 1 < p.T) {
        const float* xf = xb + (int64_t)(t + 1) * step_elems;
#pragma unroll
        for (int kk = 0; kk < NT; ++kk) xn[kk] = off[kk] >= 0 ? __ldg(xf + off[kk]) : 0.0f;
      }
      float acc[CPT];
#pragma unroll
      for (int c = 0; c < CPT; ++c) acc[c] = 0.0f;
#pragma unroll
      for (int kk = 0; kk < NT; ++kk) {
        // Event frames are ~93 % zeros: a tap whose input is zero for all 4 pixels of the warp contributes
        // exactly nothing (fmaf(0, w, acc) == acc), so the warp skips its weight loads and FMAs (warp-uniform
        // branch; results are bit-identical to the dense evaluation).
        if (__any_sync(amask, xv[kk] != 0.0f)) {
          const uint32_t wr = w_base + (uint32_t)(kk * p.Cout) * 4u;
#pragma unroll
          for (int c = 0; c < CPT; c += 4) {
            float4 w4;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(w4.x), "=f"(w4.y), "=f"(w4.z), "=f"(w4.w)
                         : "r"(wr + (uint32_t)(c ? chalf : 0) * 4u));
            acc[c] = fmaf(xv[kk], w4.x, acc[c]);
            acc[c + 1] = fmaf(xv[kk], w4.y, acc[c + 1]);
            acc[c + 2] = fmaf(xv[kk], w4.z, acc[c + 2]);
            acc[c + 3] = fmaf(xv[kk], w4.w, acc[c + 3]);
          }
        }
      }
      direct_neuron_store<CPT>(p.np, acc, sh, v, cur, yb + (int64_t)t * y_step, off_hi, p.out_fmt);
    }
  } else {
    for (int t = 0; t < p.T; ++t) {
      const float* xf = xb + (int64_t)t * step_elems;
      float acc[CPT];
#pragma unroll
      for (int c = 0; c < CPT; ++c) acc[c] = 0.0f;
      int kk = 0;
      for (int ky = 0; ky < p.k; ++ky) {
        const int iy = iy0 + ky;
        for (int kx = 0; kx < p.k; ++kx) {
          const int ix = ix0 + kx;
          const bool inb = iy >= 0 && iy < p.H && ix >= 0 && ix < p.W;
          for (int ci = 0; ci < p.Cin; ++ci, ++kk) {
            const float xv = inb ? __ldg(xf + ((int64_t)ci * p.H + iy) * p.W + ix) : 0.0f;
            const float* wr = smem_w + kk * p.Cout + c0;
#pragma unroll
            for (int c = 0; c < CPT; c += 4) {
              const float4 w4 = *reinterpret_cast<const float4*>(wr + (c ? chalf : 0));
              acc[c] = fmaf(xv, w4.x, acc[c]);
              acc[c + 1] = fmaf(xv, w4.y, acc[c + 1]);
              acc[c + 2] = fmaf(xv, w4.z, acc[c + 2]);
              acc[c + 3] = fmaf(xv, w4.w, acc[c + 3]);
            }
          }
        }
      }
      direct_neuron_store<CPT>(p.np, acc, sh, v, cur, yb + (int64_t)t * y_step, off_hi, p.out_fmt);
    }
  }
#pragma unroll
  for (int c = 0; c < CPT; c += 4) {
    *reinterpret_cast<float4*>(v_out + soff + (c ? chalf : 0)) = make_float4(v[c], v[c + 1], v[c + 2], v[c + 3]);
    *reinterpret_cast<float4*>(i_out + soff + (c ? chalf : 0)) = make_float4(cur[c], cur[c + 1], cur[c + 2], cur[c + 3]);
  }
}

static inline int direct_out_fmt(const soda_conv_desc_t* d) {
  if (d->flags & SODA_CONV_FLAG_OUT_HALF) return SODA_ACT_HALF;
  if (d->flags & SODA_CONV_FLAG_OUT_PAIR) return SODA_ACT_PAIR;
  return SODA_ACT_F32;
}

bool conv_direct_lif_supported(const soda_conv_desc_t* d, const float* y) {
  if (!d->in_nchw || d->Cout % 8 != 0) return false;
  if ((int64_t)d->Cin * d->k * d->k * d->Cout > 11000) return false;  // weights + shift <= 48 KB of smem
  const int fmt = direct_out_fmt(d);
  if (fmt == SODA_ACT_PAIR) {
    if (d->Cout % 32 != 0 || d->out_pix_stride % 32 != 0 || d->out_batch_stride % 32 != 0) return false;
    if ((reinterpret_cast<uintptr_t>(y) & 127) != 0) return false;
    return true;
  }
  if (d->out_pix_stride % 4 != 0 || d->out_batch_stride % 4 != 0) return false;
  if ((reinterpret_cast<uintptr_t>(y) & (fmt == SODA_ACT_HALF ? 7 : 15)) != 0) return false;
  return true;
}

int conv_direct_lif_dispatch(const soda_conv_desc_t* d, const float* x, const float* wp, int cout_pad,
                             const float* shift, int T, const soda_neuron_params_t* np, const float* v0,
                             const float* i0, float* y, float* v_out, float* i_out, void* stream) {
  if (!conv_direct_lif_supported(d, y))
    return fail(SODA_ERR_UNSUPPORTED,
                "conv_lif(fp32): needs NCHW frames, Cout % 8 == 0, Cin*k*k*Cout <= 11000, 16-byte aligned output");
  if (T <= 0 || d->N % T != 0) return fail(SODA_ERR_BAD_ARG, "conv_lif: N must be T * batch");
  if ((v0 == nullptr) != (i0 == nullptr)) return fail(SODA_ERR_BAD_ARG, "conv_lif: v0 / i0 must come together");
  DirectP p;
  p.T = T, p.B = d->N / T, p.H = d->H, p.W = d->W, p.Cin = d->Cin, p.Cout = d->Cout, p.Cout_pad = cout_pad;
  p.k = d->k, p.stride = d->stride, p.pad = d->k / 2;
  p.Ho = (d->H + 2 * p.pad - d->k) / d->stride + 1;
  p.Wo = (d->W + 2 * p.pad - d->k) / d->stride + 1;
  if (p.Ho <= 0 || p.Wo <= 0) return fail(SODA_ERR_BAD_ARG, "conv: empty output");
  p.Ktot = d->Cin * d->k * d->k;
  p.out_fmt = direct_out_fmt(d);
  const int ob = p.out_fmt == SODA_ACT_HALF ? 2 : 4;
  p.out_bs = d->out_batch_stride * ob, p.out_ps = d->out_pix_stride * ob;
  p.np = to_np(np);
  constexpr int CPT = 8;
  const int64_t threads = (int64_t)p.B * p.Ho * p.Wo * (p.Cout / CPT);
  const size_t smem = ((size_t)p.Ktot * p.Cout + p.Cout) * sizeof(float);
  const unsigned grid = (unsigned)ceil_div(threads, 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (d->k == 3 && d->Cin == 2 && (int64_t)d->Cin * d->H * d->W < (1LL << 31))
    conv_direct_lif_kernel<CPT, 3, 2><<<grid, 256, smem, st>>>(p, x, wp, shift, v0, i0, reinterpret_cast<uint8_t*>(y), v_out, i_out);
  else
    conv_direct_lif_kernel<CPT, 0, 0><<<grid, 256, smem, st>>>(p, x, wp, shift, v0, i0, reinterpret_cast<uint8_t*>(y), v_out, i_out);
  return after_launch("conv_direct_lif_kernel");
}

}  // namespace soda
