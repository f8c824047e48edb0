// Direct small-K 3x3 convolution + folded BatchNorm + LIF over T steps + 2x2 sum pooling of the spikes, one kernel
// (soda_conv_lif_pool_fwd): the generator block  Conv, Norm, LIF, Pool("S")  of the reference's README.md:57-64
// (vgg_block(8)/Pool/vgg_block(32)/Pool/...; nn.Conv2d models/modules/layer_gen.py:129-136, eval BatchNorm
// :211-214, norse LIFCell :233, SumPool2d models/modules/common.py:45-49), run once per time step by
// models/generator.py:184-198.
//
// K = Cin * 9 is 18 or 72 here: far below an MMA tile, and the block is HBM-bound when nothing but its input frame
// and its POOLED spike counts touch memory (row N1 of the scope table: 2 bytes per pooled count instead of a 4-byte
// spike per neuron).  One thread owns one 2x2 pooling window x 8 output channels for ALL T steps: 32 membrane
// potentials + 32 synaptic currents in registers, the 4x4 input patch of a step streamed through registers row by
// row, weights (BatchNorm scale folded, [tap][ci][Cout]) and shift in shared memory.  The full-resolution spike
// map never exists; the pooled counts {0..4} are written in the consumer's format (fp16, or the fp16 hi|lo pair
// layout with zero lo halves for a tensor-core consumer, or fp32).
#include <cuda_fp16.h>

#include "common.cuh"
#include "neuron.cuh"

namespace soda {

struct ClpP {
  int T, B, H, W, Cin, Cout, Cout_pad, Hp, Wp;
  int in_fmt;                 // SODA_ACT_F32 with nchw = 1 (event frames), or SODA_ACT_HALF pixel-major
  int64_t in_bs, in_ps;       // elements (pixel-major input)
  int out_fmt;
  int64_t out_bs, out_ps;     // BYTES
  NP np;
};

__device__ __forceinline__ int64_t clp_ch_off(int fmt, int c) {
  if (fmt == SODA_ACT_F32) return (int64_t)c * 4;
  if (fmt == SODA_ACT_HALF) return (int64_t)c * 2;
  return (int64_t)(c & ~31) * 4 + (int64_t)(c & 31) * 2;
}

// CIN = 2: plain NCHW fp32 frames [T*B, 2, H, W];  CIN = 8: pixel-major fp16 [T*B, H, W, 8]
template <int CIN>
__global__ void __launch_bounds__(128)
conv3_lif_pool2_kernel(ClpP p, const void* __restrict__ xin, const float* __restrict__ wp, const float* __restrict__ shift,
                       const float* __restrict__ v0, const float* __restrict__ i0, uint8_t* __restrict__ y,
                       float* __restrict__ v_out, float* __restrict__ i_out) {
  pdl_trigger();
  extern __shared__ float smem_w[];   // [9 * CIN][Cout] weights, then [Cout] shift
  const int Ktot = 9 * CIN;
  float* smem_shift = smem_w + Ktot * p.Cout;
  for (int e = threadIdx.x; e < Ktot * p.Cout; e += blockDim.x) {
    const int kk = e / p.Cout, co = e - kk * p.Cout;
    smem_w[e] = wp[(int64_t)kk * p.Cout_pad + co];
  }
  for (int e = threadIdx.x; e < p.Cout; e += blockDim.x) smem_shift[e] = shift ? shift[e] : 0.0f;
  __syncthreads();

  const int G = p.Cout / 8;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int g = (int)(gid % G);
  const int64_t win = gid / G;
  const int64_t wins = (int64_t)p.Hp * p.Wp;
  if (win >= (int64_t)p.B * wins) return;
  const int b = (int)(win / wins);
  const int wr = (int)(win - (int64_t)b * wins);
  const int py = wr / p.Wp, px = wr - py * p.Wp;
  const int c0 = g * 8;
  const int iy0 = 2 * py - 1, ix0 = 2 * px - 1;   // top-left of the 4x4 input patch (pad 1)

  float v[4][8], cur[4][8];   // window pixel q = oy * 2 + ox
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int64_t soff = (((int64_t)b * p.H + (2 * py + (q >> 1))) * p.W + (2 * px + (q & 1))) * p.Cout + c0;
    if (v0 != nullptr) {
      const float4 a0 = __ldg(reinterpret_cast<const float4*>(v0 + soff)), a1 = __ldg(reinterpret_cast<const float4*>(v0 + soff + 4));
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(i0 + soff)), b1 = __ldg(reinterpret_cast<const float4*>(i0 + soff + 4));
      v[q][0] = a0.x, v[q][1] = a0.y, v[q][2] = a0.z, v[q][3] = a0.w, v[q][4] = a1.x, v[q][5] = a1.y, v[q][6] = a1.z, v[q][7] = a1.w;
      cur[q][0] = b0.x, cur[q][1] = b0.y, cur[q][2] = b0.z, cur[q][3] = b0.w, cur[q][4] = b1.x, cur[q][5] = b1.y, cur[q][6] = b1.z, cur[q][7] = b1.w;
    } else {
#pragma unroll
      for (int c = 0; c < 8; ++c) v[q][c] = p.np.v_leak, cur[q][c] = 0.0f;
    }
  }
  float sh[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) sh[c] = smem_shift[c0 + c];
  const uint32_t w_base = (uint32_t)__cvta_generic_to_shared(smem_w + c0);
  uint8_t* yb = y + (int64_t)b * p.out_bs + (int64_t)wr * p.out_ps + clp_ch_off(p.out_fmt, c0);
  const int64_t y_step = (int64_t)p.B * p.out_bs;

  for (int t = 0; t < p.T; ++t) {
    float acc[4][8];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int c = 0; c < 8; ++c) acc[q][c] = 0.0f;
    const int64_t frame = (int64_t)t * p.B + b;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int iy = iy0 + r;
      float in[4][CIN];
      bool any = false;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int ix = ix0 + c;
        const bool inb = iy >= 0 && iy < p.H && ix >= 0 && ix < p.W;
        if constexpr (CIN == 2) {
          const float* xf = reinterpret_cast<const float*>(xin) + frame * (2LL * p.H * p.W);
          in[c][0] = inb ? __ldg(xf + (int64_t)iy * p.W + ix) : 0.0f;
          in[c][1] = inb ? __ldg(xf + ((int64_t)p.H + iy) * p.W + ix) : 0.0f;
        } else {
          uint4 u = make_uint4(0u, 0u, 0u, 0u);
          if (inb)
            u = __ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const __half*>(xin) + frame * p.in_bs +
                                                     ((int64_t)iy * p.W + ix) * p.in_ps));
          const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&uw[k]));
            in[c][2 * k] = f.x, in[c][2 * k + 1] = f.y;
          }
        }
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci) any = any || in[c][ci] != 0.0f;
      }
      // event frames and spike counts are mostly zeros: a patch row that is zero for the whole warp adds nothing
      if (!__any_sync(__activemask(), any)) continue;
#pragma unroll
      for (int oy = 0; oy < 2; ++oy) {
        const int ky = r - oy;
        if (ky < 0 || ky > 2) continue;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
#pragma unroll
          for (int ci = 0; ci < CIN; ++ci) {
            const uint32_t wa = w_base + (uint32_t)(((ky * 3 + kx) * CIN + ci) * p.Cout) * 4u;
            float4 w0, w1;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(w0.x), "=f"(w0.y), "=f"(w0.z), "=f"(w0.w) : "r"(wa));
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(w1.x), "=f"(w1.y), "=f"(w1.z), "=f"(w1.w) : "r"(wa + 16u));
            const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
            for (int ox = 0; ox < 2; ++ox) {
              const float a = in[ox + kx][ci];
#pragma unroll
              for (int c = 0; c < 8; ++c) acc[oy * 2 + ox][c] = fmaf(a, ww[c], acc[oy * 2 + ox][c]);
            }
          }
        }
      }
    }
    float pooled[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) pooled[c] = 0.0f;
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        float z, vd;
        lif_step(p.np, __fadd_rn(acc[q][c], sh[c]), v[q][c], cur[q][c], z, vd);
        pooled[c] += z;
      }
    uint8_t* dst = yb + (int64_t)t * y_step;
    if (p.out_fmt == SODA_ACT_F32) {
      __stcs(reinterpret_cast<float4*>(dst), make_float4(pooled[0], pooled[1], pooled[2], pooled[3]));
      __stcs(reinterpret_cast<float4*>(dst + 16), make_float4(pooled[4], pooled[5], pooled[6], pooled[7]));
    } else {
      uint4 u;
      const __half2 h0 = __floats2half2_rn(pooled[0], pooled[1]), h1 = __floats2half2_rn(pooled[2], pooled[3]);
      const __half2 h2 = __floats2half2_rn(pooled[4], pooled[5]), h3 = __floats2half2_rn(pooled[6], pooled[7]);
      u.x = *reinterpret_cast<const uint32_t*>(&h0), u.y = *reinterpret_cast<const uint32_t*>(&h1);
      u.z = *reinterpret_cast<const uint32_t*>(&h2), u.w = *reinterpret_cast<const uint32_t*>(&h3);
      __stcs(reinterpret_cast<uint4*>(dst), u);
      if (p.out_fmt == SODA_ACT_PAIR) __stcs(reinterpret_cast<uint4*>(dst + 64), make_uint4(0u, 0u, 0u, 0u));
    }
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int64_t soff = (((int64_t)b * p.H + (2 * py + (q >> 1))) * p.W + (2 * px + (q & 1))) * p.Cout + c0;
    *reinterpret_cast<float4*>(v_out + soff) = make_float4(v[q][0], v[q][1], v[q][2], v[q][3]);
    *reinterpret_cast<float4*>(v_out + soff + 4) = make_float4(v[q][4], v[q][5], v[q][6], v[q][7]);
    *reinterpret_cast<float4*>(i_out + soff) = make_float4(cur[q][0], cur[q][1], cur[q][2], cur[q][3]);
    *reinterpret_cast<float4*>(i_out + soff + 4) = make_float4(cur[q][4], cur[q][5], cur[q][6], cur[q][7]);
  }
}

}  // namespace soda

extern "C" int soda_conv_lif_pool_supported(int Cin, int Cout, int k, int stride, int H, int W, int in_nchw, int in_fmt) {
  if (k != 3 || stride != 1 || (H & 1) || (W & 1) || Cout % 8 != 0 || 9 * Cin * Cout + Cout > 11000) return 0;
  if (Cin == 2 && in_nchw && in_fmt == SODA_ACT_F32) return 1;
  if (Cin == 8 && !in_nchw && in_fmt == SODA_ACT_HALF) return 1;
  return 0;
}

extern "C" int soda_conv_lif_pool_fwd(const void* x, int in_fmt, int in_nchw, int64_t in_batch_stride, int64_t in_pix_stride,
                                      const float* packed, const float* shift, int T, int B, int H, int W, int Cin, int Cout,
                                      const soda_neuron_params_t* np, const float* v0, const float* i0, void* y, int out_fmt,
                                      int64_t out_batch_stride, int64_t out_pix_stride, float* v_out, float* i_out,
                                      void* stream) {
  using namespace soda;
  if (!x || !packed || !np || !y || !v_out || !i_out) return fail(SODA_ERR_BAD_ARG, "conv_lif_pool: null pointer");
  if ((v0 == nullptr) != (i0 == nullptr)) return fail(SODA_ERR_BAD_ARG, "conv_lif_pool: v0 / i0 must come together");
  if (T <= 0 || B <= 0) return fail(SODA_ERR_BAD_ARG, "conv_lif_pool: non-positive size");
  if (!soda_conv_lif_pool_supported(Cin, Cout, 3, 1, H, W, in_nchw, in_fmt))
    return fail(SODA_ERR_UNSUPPORTED,
                "conv_lif_pool: 3x3 stride-1 convolutions on even maps with Cin = 2 (NCHW fp32 frames) or 8 (fp16 pixel-major), "
                "Cout % 8 == 0");
  if (out_fmt < SODA_ACT_F32 || out_fmt > SODA_ACT_HALF) return fail(SODA_ERR_BAD_ARG, "conv_lif_pool: unknown output format");
  const int ob = out_fmt == SODA_ACT_HALF ? 2 : 4;
  const int al = out_fmt == SODA_ACT_PAIR ? 128 : 16;
  if ((reinterpret_cast<uintptr_t>(y) % al) || (out_batch_stride * ob) % al || (out_pix_stride * ob) % al ||
      (out_fmt == SODA_ACT_PAIR && Cout % 32))
    return fail(SODA_ERR_BAD_ARG, "conv_lif_pool: misaligned output");
  if (!in_nchw && ((reinterpret_cast<uintptr_t>(x) & 15) || (in_pix_stride * 2) % 16 || (in_batch_stride * 2) % 16))
    return fail(SODA_ERR_BAD_ARG, "conv_lif_pool: misaligned input");
  ClpP p;
  p.T = T, p.B = B, p.H = H, p.W = W, p.Cin = Cin, p.Cout = Cout, p.Cout_pad = (Cout + 31) / 32 * 32;
  p.Hp = H / 2, p.Wp = W / 2;
  p.in_fmt = in_fmt, p.in_bs = in_batch_stride, p.in_ps = in_pix_stride;
  p.out_fmt = out_fmt, p.out_bs = out_batch_stride * ob, p.out_ps = out_pix_stride * ob;
  p.np = to_np(np);
  const int64_t threads = (int64_t)B * p.Hp * p.Wp * (Cout / 8);
  const size_t smem = ((size_t)9 * Cin * Cout + Cout) * sizeof(float);
  const unsigned grid = (unsigned)ceil_div(threads, 128);
  cudaStream_t st = (cudaStream_t)stream;
  if (Cin == 2)
    conv3_lif_pool2_kernel<2><<<grid, 128, smem, st>>>(p, x, packed, shift, v0, i0, reinterpret_cast<uint8_t*>(y), v_out, i_out);
  else
    conv3_lif_pool2_kernel<8><<<grid, 128, smem, st>>>(p, x, packed, shift, v0, i0, reinterpret_cast<uint8_t*>(y), v_out, i_out);
  return after_launch("conv3_lif_pool2_kernel");
}
