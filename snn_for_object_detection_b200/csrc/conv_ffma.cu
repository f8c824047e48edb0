// fp32 implicit-GEMM convolution on the FP32 pipe (exact-fp32 path; soda_conv2d_fwd kind 0).
//
// Reference arithmetic: nn.Conv2d(bias=False, padding=k/2) models/modules/layer_gen.py:129-136,
// followed by the eval BatchNorm affine (layer_gen.py:211-214), the Residual sum
// (models/generator.py:146) and the Dense concat slice store (:158) in the epilogue.
//
//   D[m, co] = sum_k A[m, k] * Wp[k, co],  m = (n, oy, ox),  k = (ky*K + kx)*Cin + ci
//
// A is gathered on the fly from the pixel-major activation tensor (or the NCHW event frames);
// Wp is the packed weight (soda_conv_pack_weights: [k*k*Cin][Cout_pad], BN scale folded in).
// CTA tile BM x BN, K chunk 16, 8x8 register micro-tile, double-buffered shared memory with the
// next chunk's global loads held in registers while the current one is consumed.
//
// This is the bit-faithful fp32 route: it serves the shapes the tensor-core kernel does not
// take (Cin not a multiple of 32, NCHW frames) and is the on-device cross-check for it.
#include "common.cuh"

namespace soda {

constexpr int BK = 16;
constexpr int TM = 8, TN = 8;

struct ConvP {
  int N, H, W, Cin, Cout, Cout_pad, k, stride, pad, Ho, Wo;
  int in_nchw;
  int64_t M;  // N*Ho*Wo
  int Ktot;
  int64_t in_bs, in_ps, out_bs, out_ps, res_bs, res_ps;
};

template <int BM, int BN, bool kFastA>
__global__ void __launch_bounds__((BM / TM) * (BN / TN))
conv_ffma_kernel(ConvP p, const float* __restrict__ x, const float* __restrict__ wp,
                 const float* __restrict__ shift, const float* __restrict__ residual,
                 float* __restrict__ y) {
  constexpr int NT = (BM / TM) * (BN / TN);
  constexpr int LDA = BM + 4;
  __shared__ __align__(16) float As[2][BK][LDA];
  __shared__ __align__(16) float Bs[2][BK][BN];

  const int tid = threadIdx.x;
  const int tx = tid % (BN / TN);  // n group
  const int ty = tid / (BN / TN);  // m group
  const int64_t m0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int HoWo = p.Ho * p.Wo;

  // ---- A loader set-up -------------------------------------------------------------------
  // fast: 4 lanes per pixel, one float4 (4 channels) each; generic: lane -> m, loop over k.
  constexpr int A_F4 = BM * (BK / 4) / NT;    // float4 loads per thread (fast path)
  constexpr int A_SC = BM * BK / NT;          // scalar loads per thread (generic path)
  constexpr int A_REGS = kFastA ? A_F4 * 4 : A_SC;
  float a_stage[A_REGS];

  int a_oy[kFastA ? A_F4 : 1], a_ox[kFastA ? A_F4 : 1];
  const float* a_base[kFastA ? A_F4 : 1];
  bool a_ok[kFastA ? A_F4 : 1];
  if constexpr (kFastA) {
#pragma unroll
    for (int j = 0; j < A_F4; ++j) {
      const int pm = (tid >> 2) + j * (NT >> 2);
      const int64_t m = m0 + pm;
      a_ok[j] = m < p.M;
      const int64_t mm = a_ok[j] ? m : 0;
      const int n = (int)(mm / HoWo);
      const int r = (int)(mm - (int64_t)n * HoWo);
      a_oy[j] = (r / p.Wo) * p.stride - p.pad;
      a_ox[j] = (r % p.Wo) * p.stride - p.pad;
      a_base[j] = x + (int64_t)n * p.in_bs + ((tid & 3) << 2);
    }
  } else {
    const int pm = tid % BM;
    const int64_t m = m0 + pm;
    a_ok[0] = m < p.M;
    const int64_t mm = a_ok[0] ? m : 0;
    const int n = (int)(mm / HoWo);
    const int r = (int)(mm - (int64_t)n * HoWo);
    a_oy[0] = (r / p.Wo) * p.stride - p.pad;
    a_ox[0] = (r % p.Wo) * p.stride - p.pad;
    a_base[0] = x + (int64_t)n * (p.in_nchw ? (int64_t)p.Cin * p.H * p.W : p.in_bs);
  }

  // ---- B loader set-up: BK x BN floats as float4, row-major in Wp ---------------------------
  constexpr int B_F4 = BK * BN / 4 / NT > 0 ? BK * BN / 4 / NT : 1;
  constexpr int B_THREADS = BK * BN / 4;  // threads that carry a float4 when B_F4 == 1
  float4 b_stage[B_F4];

  auto load_chunk = [&](int kc) {
    const int k0 = kc * BK;
    if constexpr (kFastA) {
      const int tap = k0 / p.Cin;
      const int ci0 = k0 - tap * p.Cin;
      const int ky = tap / p.k, kx = tap - ky * p.k;
#pragma unroll
      for (int j = 0; j < A_F4; ++j) {
        const int iy = a_oy[j] + ky, ix = a_ox[j] + kx;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (a_ok[j] && iy >= 0 && iy < p.H && ix >= 0 && ix < p.W)
          v = __ldg(reinterpret_cast<const float4*>(a_base[j] + ((int64_t)iy * p.W + ix) * p.in_ps + ci0));
        a_stage[j * 4 + 0] = v.x, a_stage[j * 4 + 1] = v.y;
        a_stage[j * 4 + 2] = v.z, a_stage[j * 4 + 3] = v.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < A_SC; ++j) {
        const int k = k0 + tid / BM + j * (NT / BM);
        float v = 0.f;
        if (a_ok[0] && k < p.Ktot) {
          const int tap = k / p.Cin;
          const int ci = k - tap * p.Cin;
          const int ky = tap / p.k, kx = tap - ky * p.k;
          const int iy = a_oy[0] + ky, ix = a_ox[0] + kx;
          if (iy >= 0 && iy < p.H && ix >= 0 && ix < p.W) {
            if (p.in_nchw)
              v = __ldg(a_base[0] + ((int64_t)ci * p.H + iy) * p.W + ix);
            else
              v = __ldg(a_base[0] + ((int64_t)iy * p.W + ix) * p.in_ps + ci);
          }
        }
        a_stage[j] = v;
      }
    }
#pragma unroll
    for (int j = 0; j < B_F4; ++j) {
      const int f = tid + j * NT;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (f < B_THREADS) {
        const int kr = f / (BN / 4), nc = (f % (BN / 4)) * 4;
        if (k0 + kr < p.Ktot && n0 + nc < p.Cout_pad)
          v = __ldg(reinterpret_cast<const float4*>(wp + (int64_t)(k0 + kr) * p.Cout_pad + n0 + nc));
      }
      b_stage[j] = v;
    }
  };

  auto store_chunk = [&](int buf) {
    if constexpr (kFastA) {
#pragma unroll
      for (int j = 0; j < A_F4; ++j) {
        const int pm = (tid >> 2) + j * (NT >> 2);
        const int kq = (tid & 3) << 2;
#pragma unroll
        for (int q = 0; q < 4; ++q) As[buf][kq + q][pm] = a_stage[j * 4 + q];
      }
    } else {
#pragma unroll
      for (int j = 0; j < A_SC; ++j) As[buf][tid / BM + j * (NT / BM)][tid % BM] = a_stage[j];
    }
#pragma unroll
    for (int j = 0; j < B_F4; ++j) {
      const int f = tid + j * NT;
      if (f < B_THREADS) {
        const int kr = f / (BN / 4), nc = (f % (BN / 4)) * 4;
        *reinterpret_cast<float4*>(&Bs[buf][kr][nc]) = b_stage[j];
      }
    }
  };

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  const int n_chunks = (p.Ktot + BK - 1) / BK;
  load_chunk(0);
  store_chunk(0);
  __syncthreads();

  for (int kc = 0; kc < n_chunks; ++kc) {
    const int buf = kc & 1;
    if (kc + 1 < n_chunks) load_chunk(kc + 1);
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[TM], b[TN];
      const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * TM]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * TM + 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * TN]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * TN + 4]);
      a[0] = a0.x, a[1] = a0.y, a[2] = a0.z, a[3] = a0.w;
      a[4] = a1.x, a[5] = a1.y, a[6] = a1.z, a[7] = a1.w;
      b[0] = b0.x, b[1] = b0.y, b[2] = b0.z, b[3] = b0.w;
      b[4] = b1.x, b[5] = b1.y, b[6] = b1.z, b[7] = b1.w;
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (kc + 1 < n_chunks) {
      store_chunk(buf ^ 1);
      __syncthreads();
    }
  }

  // ---- epilogue: + shift, + residual, store into the (possibly strided) output slice ---------
  const int co0 = n0 + tx * TN;
  float sh[TN];
#pragma unroll
  for (int j = 0; j < TN; ++j) sh[j] = (shift != nullptr && co0 + j < p.Cout) ? __ldg(shift + co0 + j) : 0.f;
  const bool vec_ok = (co0 + TN <= p.Cout) && (p.out_ps % 4 == 0) && (p.out_bs % 4 == 0) &&
                      aligned16(y) && (residual == nullptr ||
                                       (p.res_ps % 4 == 0 && p.res_bs % 4 == 0 && aligned16(residual)));
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int64_t m = m0 + ty * TM + i;
    if (m >= p.M) continue;
    const int n = (int)(m / HoWo);
    const int64_t r = m - (int64_t)n * HoWo;
    float* dst = y + (int64_t)n * p.out_bs + r * p.out_ps + co0;
    const float* rs = residual ? residual + (int64_t)n * p.res_bs + r * p.res_ps + co0 : nullptr;
    float o[TN];
#pragma unroll
    for (int j = 0; j < TN; ++j) o[j] = acc[i][j] + sh[j];
    if (vec_ok) {
      if (rs != nullptr) {
        const float4 r0 = __ldg(reinterpret_cast<const float4*>(rs));
        const float4 r1 = __ldg(reinterpret_cast<const float4*>(rs + 4));
        o[0] += r0.x, o[1] += r0.y, o[2] += r0.z, o[3] += r0.w;
        o[4] += r1.x, o[5] += r1.y, o[6] += r1.z, o[7] += r1.w;
      }
      *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(o[4], o[5], o[6], o[7]);
    } else {
#pragma unroll
      for (int j = 0; j < TN; ++j) {
        if (co0 + j < p.Cout) dst[j] = o[j] + (rs ? __ldg(rs + j) : 0.f);
      }
    }
  }
}

// [Cout, Cin, k, k] -> [(ky*k + kx)*Cin + ci][Cout_pad], scale folded, zero padded columns.
__global__ void pack_weights_kn_kernel(const float* __restrict__ w, const float* __restrict__ scale,
                                       int Cout, int Cin, int kk, int Cout_pad,
                                       float* __restrict__ wp) {
  const int64_t total = (int64_t)kk * Cin * Cout_pad;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int co = (int)(i % Cout_pad);
    const int64_t k = i / Cout_pad;
    const int tap = (int)(k / Cin), ci = (int)(k % Cin);
    float v = 0.f;
    if (co < Cout) {
      v = w[((int64_t)co * Cin + ci) * kk + tap];
      if (scale != nullptr) v *= scale[co];
    }
    wp[i] = v;
  }
}

int conv_ffma_pack(const float* w, const float* scale, int Cout, int Cin, int k, int cout_pad,
                   float* packed, cudaStream_t st) {
  const int64_t total = (int64_t)k * k * Cin * cout_pad;
  pack_weights_kn_kernel<<<(unsigned)std::min<int64_t>(ceil_div(total, 256), 4096), 256, 0, st>>>(
      w, scale, Cout, Cin, k * k, cout_pad, packed);
  return after_launch("pack_weights_kn_kernel");
}

template <int BM, int BN>
static int launch_ffma(const ConvP& p, bool fast, const float* x, const float* wp,
                       const float* shift, const float* residual, float* y, cudaStream_t st) {
  dim3 grid((unsigned)ceil_div(p.M, BM), (unsigned)ceil_div(p.Cout, BN));
  constexpr int NT = (BM / TM) * (BN / TN);
  if (fast)
    conv_ffma_kernel<BM, BN, true><<<grid, NT, 0, st>>>(p, x, wp, shift, residual, y);
  else
    conv_ffma_kernel<BM, BN, false><<<grid, NT, 0, st>>>(p, x, wp, shift, residual, y);
  return after_launch("conv_ffma_kernel");
}

int conv_ffma_dispatch(const soda_conv_desc_t* d, const float* x, const float* wp, int cout_pad,
                       const float* shift, const float* residual, float* y, void* stream) {
  ConvP p;
  p.N = d->N, p.H = d->H, p.W = d->W, p.Cin = d->Cin, p.Cout = d->Cout, p.Cout_pad = cout_pad;
  p.k = d->k, p.stride = d->stride, p.pad = d->k / 2;
  p.Ho = (d->H + 2 * p.pad - d->k) / d->stride + 1;
  p.Wo = (d->W + 2 * p.pad - d->k) / d->stride + 1;
  p.in_nchw = d->in_nchw;
  p.M = (int64_t)p.N * p.Ho * p.Wo;
  p.Ktot = d->k * d->k * d->Cin;
  p.in_bs = d->in_batch_stride, p.in_ps = d->in_pix_stride;
  p.out_bs = d->out_batch_stride, p.out_ps = d->out_pix_stride;
  p.res_bs = d->res_batch_stride, p.res_ps = d->res_pix_stride;
  if (p.Ho <= 0 || p.Wo <= 0) return fail(SODA_ERR_BAD_ARG, "conv: empty output");
  cudaStream_t st = (cudaStream_t)stream;
  const bool fast = !d->in_nchw && (d->Cin % BK == 0) && (p.in_ps % 4 == 0) && (p.in_bs % 4 == 0) &&
                    ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
  if (d->Cout > 64) return launch_ffma<128, 128>(p, fast, x, wp, shift, residual, y, st);
  // the generic gather maps one thread per tile row and needs NT >= BM, i.e. BN >= 64
  if (d->Cout > 32 || !fast) return launch_ffma<128, 64>(p, fast, x, wp, shift, residual, y, st);
  dim3 grid((unsigned)ceil_div(p.M, 128), (unsigned)ceil_div(p.Cout, 32));
  conv_ffma_kernel<128, 32, true><<<grid, 64, 0, st>>>(p, x, wp, shift, residual, y);
  return after_launch("conv_ffma_kernel");
}

}  // namespace soda
