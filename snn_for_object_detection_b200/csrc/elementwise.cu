// Pooling, per-channel affine and n-ary add on pixel-major fp32 tensors (HBM-bound glue kernels).
//
// Reference arithmetic: models/modules/common.py:45-49 (SumPool2d = avg_pool2d * k * k),
// layer_gen.py:163-168 (Avg/Max pool), layer_gen.py:211-214 (eval BatchNorm as a stand-alone
// layer), models/generator.py:146 (Residual sum) / :158 (Dense concat via strided copy).
#include <float.h>

#include "common.cuh"

namespace soda {

// mode: 0 = sum (computed as avg * k * k like the reference), 1 = avg, 2 = max
template <int V>
__global__ void __launch_bounds__(256)
pool2d_kernel(const float* __restrict__ x, int N, int H, int W, int C, int k, int stride, int mode,
              int Ho, int Wo, int64_t in_bs, int64_t in_ps, float* __restrict__ y, int64_t out_bs,
              int64_t out_ps) {
  pdl_trigger();   // lets a dependent tensor-core conv start its prologue in this grid's tail
  const int cv = C / V;
  const int64_t total = (int64_t)N * Ho * Wo * cv;
  const float inv = 1.0f / (float)(k * k);
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % cv) * V;
    int64_t r = idx / cv;
    const int ox = (int)(r % Wo);
    r /= Wo;
    const int oy = (int)(r % Ho);
    const int n = (int)(r / Ho);
    float acc[V];
#pragma unroll
    for (int q = 0; q < V; ++q) acc[q] = (mode == 2) ? -FLT_MAX : 0.0f;
    for (int ky = 0; ky < k; ++ky)
      for (int kx = 0; kx < k; ++kx) {
        const float* src = x + (int64_t)n * in_bs + ((int64_t)(oy * stride + ky) * W + ox * stride + kx) * in_ps + c;
        float v[V];
        if constexpr (V == 4) {
          const float4 t = __ldg(reinterpret_cast<const float4*>(src));
          v[0] = t.x, v[1] = t.y, v[2] = t.z, v[3] = t.w;
        } else {
          v[0] = __ldg(src);
        }
#pragma unroll
        for (int q = 0; q < V; ++q) acc[q] = (mode == 2) ? fmaxf(acc[q], v[q]) : acc[q] + v[q];
      }
    float o[V];
#pragma unroll
    for (int q = 0; q < V; ++q) {
      if (mode == 2) o[q] = acc[q];
      else if (mode == 1) o[q] = acc[q] * inv;
      else o[q] = __fmul_rn(__fmul_rn(__fmul_rn(acc[q], inv), (float)k), (float)k);
    }
    float* dst = y + (int64_t)n * out_bs + ((int64_t)oy * Wo + ox) * out_ps + c;
    if constexpr (V == 4) *reinterpret_cast<float4*>(dst) = make_float4(o[0], o[1], o[2], o[3]);
    else *dst = o[0];
  }
}

template <int V>
__global__ void __launch_bounds__(256)
channel_affine_kernel(const float* __restrict__ x, int64_t n_pix, int C, int64_t in_ps,
                      const float* __restrict__ scale, const float* __restrict__ shift,
                      float* __restrict__ y, int64_t out_ps) {
  pdl_trigger();   // lets a dependent tensor-core conv start its prologue in this grid's tail
  const int cv = C / V;
  const int64_t total = n_pix * cv;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(idx % cv) * V;
    const int64_t pix = idx / cv;
    float v[V];
    if constexpr (V == 4) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(x + pix * in_ps + c));
      v[0] = t.x, v[1] = t.y, v[2] = t.z, v[3] = t.w;
    } else {
      v[0] = __ldg(x + pix * in_ps + c);
    }
#pragma unroll
    for (int q = 0; q < V; ++q) {
      const float s = scale ? __ldg(scale + c + q) : 1.0f;
      const float b = shift ? __ldg(shift + c + q) : 0.0f;
      v[q] = __fadd_rn(__fmul_rn(v[q], s), b);
    }
    float* dst = y + pix * out_ps + c;
    if constexpr (V == 4) *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
    else *dst = v[0];
  }
}

struct AddPtrs {
  const float* p[8];
};

template <int V>
__global__ void __launch_bounds__(256) add_n_kernel(AddPtrs xs, int k, int64_t n, float* __restrict__ y) {
  pdl_trigger();   // lets a dependent tensor-core conv start its prologue in this grid's tail
  for (int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * V; e < n;
       e += (int64_t)gridDim.x * blockDim.x * V) {
    float acc[V];
#pragma unroll
    for (int q = 0; q < V; ++q) acc[q] = 0.0f;
    for (int j = 0; j < k; ++j) {
      if constexpr (V == 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(xs.p[j] + e));
        acc[0] = j ? __fadd_rn(acc[0], t.x) : t.x, acc[1] = j ? __fadd_rn(acc[1], t.y) : t.y;
        acc[2] = j ? __fadd_rn(acc[2], t.z) : t.z, acc[3] = j ? __fadd_rn(acc[3], t.w) : t.w;
      } else {
        const float t = __ldg(xs.p[j] + e);
        acc[0] = j ? __fadd_rn(acc[0], t) : t;
      }
    }
    if constexpr (V == 4) *reinterpret_cast<float4*>(y + e) = make_float4(acc[0], acc[1], acc[2], acc[3]);
    else y[e] = acc[0];
  }
}

static unsigned grid_for(int64_t work_items) {
  int64_t b = ceil_div(work_items, 256);
  const int64_t cap = (int64_t)kNumSMs * 16;
  return (unsigned)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace soda

extern "C" int soda_pool2d_fwd(const float* x, int N, int H, int W, int C, int k, int stride,
                               int mode, int64_t in_bs, int64_t in_ps, float* y, int64_t out_bs,
                               int64_t out_ps, void* stream) {
  using namespace soda;
  if (!x || !y) return fail(SODA_ERR_BAD_ARG, "pool2d: null pointer");
  if (N <= 0 || C <= 0 || k <= 0 || stride <= 0 || mode < 0 || mode > 2)
    return fail(SODA_ERR_BAD_ARG, "pool2d: bad size or mode");
  const int Ho = (H - k) / stride + 1, Wo = (W - k) / stride + 1;
  if (H < k || W < k || Ho <= 0 || Wo <= 0) return fail(SODA_ERR_BAD_ARG, "pool2d: window larger than input");
  cudaStream_t st = (cudaStream_t)stream;
  const bool vec = (C % 4 == 0) && (in_ps % 4 == 0) && (in_bs % 4 == 0) && (out_ps % 4 == 0) &&
                   (out_bs % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0) &&
                   ((reinterpret_cast<uintptr_t>(y) & 15) == 0);
  if (vec)
    pool2d_kernel<4><<<grid_for((int64_t)N * Ho * Wo * (C / 4)), 256, 0, st>>>(
        x, N, H, W, C, k, stride, mode, Ho, Wo, in_bs, in_ps, y, out_bs, out_ps);
  else
    pool2d_kernel<1><<<grid_for((int64_t)N * Ho * Wo * C), 256, 0, st>>>(
        x, N, H, W, C, k, stride, mode, Ho, Wo, in_bs, in_ps, y, out_bs, out_ps);
  return after_launch("pool2d_kernel");
}

extern "C" int soda_channel_affine(const float* x, int64_t n_pix, int C, int64_t in_ps,
                                   const float* scale, const float* shift, float* y,
                                   int64_t out_ps, void* stream) {
  using namespace soda;
  if (!x || !y) return fail(SODA_ERR_BAD_ARG, "channel_affine: null pointer");
  if (n_pix <= 0 || C <= 0) return fail(SODA_ERR_BAD_ARG, "channel_affine: non-positive size");
  cudaStream_t st = (cudaStream_t)stream;
  const bool vec = (C % 4 == 0) && (in_ps % 4 == 0) && (out_ps % 4 == 0) &&
                   ((reinterpret_cast<uintptr_t>(x) & 15) == 0) &&
                   ((reinterpret_cast<uintptr_t>(y) & 15) == 0);
  if (vec)
    channel_affine_kernel<4><<<grid_for(n_pix * (C / 4)), 256, 0, st>>>(x, n_pix, C, in_ps, scale,
                                                                      shift, y, out_ps);
  else
    channel_affine_kernel<1><<<grid_for(n_pix * C), 256, 0, st>>>(x, n_pix, C, in_ps, scale, shift,
                                                                y, out_ps);
  return after_launch("channel_affine_kernel");
}

extern "C" int soda_add_n(const float* const* xs, int k, int64_t n, float* y, void* stream) {
  using namespace soda;
  if (!xs || !y) return fail(SODA_ERR_BAD_ARG, "add_n: null pointer");
  if (k <= 0 || k > 8 || n <= 0) return fail(SODA_ERR_BAD_ARG, "add_n: need 1..8 inputs and n > 0");
  AddPtrs ap;
  bool vec = (n % 4 == 0) && ((reinterpret_cast<uintptr_t>(y) & 15) == 0);
  for (int j = 0; j < k; ++j) {
    if (!xs[j]) return fail(SODA_ERR_BAD_ARG, "add_n: null input");
    ap.p[j] = xs[j];
    vec = vec && ((reinterpret_cast<uintptr_t>(xs[j]) & 15) == 0);
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (vec) add_n_kernel<4><<<grid_for(n / 4), 256, 0, st>>>(ap, k, n, y);
  else add_n_kernel<1><<<grid_for(n), 256, 0, st>>>(ap, k, n, y);
  return after_launch("add_n_kernel");
}

// ConvLSTM gates (reference models/modules/conv_lstm.py:62-75): pre [N, H, W, 4*hid] pixel-major holds the conv output
// (input | forget | out gate | in node, each hid channels);  c' = sigmoid(f) * c + sigmoid(i) * tanh(g),
// h' = sigmoid(o) * tanh(c').  One thread = 4 channels of one pixel.
namespace soda {
__device__ __forceinline__ float sigmoidf_(float z) { return __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-z))); }

__global__ void __launch_bounds__(256)
lstm_gate_kernel(const float* __restrict__ pre, const float* __restrict__ c_prev, int64_t n_pix, int hid,
                 float* __restrict__ h_out, float* __restrict__ c_out) {
  pdl_trigger();
  const int h4 = hid >> 2;
  const int64_t total = n_pix * h4;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t pix = idx / h4;
    const int cg = (int)(idx % h4);
    const float4* row = reinterpret_cast<const float4*>(pre + pix * 4 * hid);
    const float4 gi = __ldg(row + cg), gf = __ldg(row + h4 + cg), go = __ldg(row + 2 * h4 + cg), gg = __ldg(row + 3 * h4 + cg);
    float4 c = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c_prev != nullptr) c = __ldg(reinterpret_cast<const float4*>(c_prev + pix * hid) + cg);
    float4 cn, hn;
    cn.x = __fadd_rn(__fmul_rn(sigmoidf_(gf.x), c.x), __fmul_rn(sigmoidf_(gi.x), tanhf(gg.x)));
    cn.y = __fadd_rn(__fmul_rn(sigmoidf_(gf.y), c.y), __fmul_rn(sigmoidf_(gi.y), tanhf(gg.y)));
    cn.z = __fadd_rn(__fmul_rn(sigmoidf_(gf.z), c.z), __fmul_rn(sigmoidf_(gi.z), tanhf(gg.z)));
    cn.w = __fadd_rn(__fmul_rn(sigmoidf_(gf.w), c.w), __fmul_rn(sigmoidf_(gi.w), tanhf(gg.w)));
    hn.x = __fmul_rn(sigmoidf_(go.x), tanhf(cn.x)), hn.y = __fmul_rn(sigmoidf_(go.y), tanhf(cn.y));
    hn.z = __fmul_rn(sigmoidf_(go.z), tanhf(cn.z)), hn.w = __fmul_rn(sigmoidf_(go.w), tanhf(cn.w));
    reinterpret_cast<float4*>(c_out + pix * hid)[cg] = cn;
    reinterpret_cast<float4*>(h_out + pix * hid)[cg] = hn;
  }
}
}  // namespace soda

extern "C" int soda_lstm_gates(const float* pre, const float* c_prev, int64_t n_pix, int hidden, float* h_out,
                               float* c_out, void* stream) {
  using namespace soda;
  if (!pre || !h_out || !c_out) return fail(SODA_ERR_BAD_ARG, "lstm_gates: null pointer");
  if (n_pix <= 0 || hidden <= 0) return fail(SODA_ERR_BAD_ARG, "lstm_gates: non-positive size");
  if (hidden % 4 != 0 || (reinterpret_cast<uintptr_t>(pre) & 15) || (reinterpret_cast<uintptr_t>(h_out) & 15) ||
      (reinterpret_cast<uintptr_t>(c_out) & 15) || (reinterpret_cast<uintptr_t>(c_prev) & 15))
    return fail(SODA_ERR_UNSUPPORTED, "lstm_gates: hidden % 4 == 0 and 16-byte aligned buffers");
  const int64_t total = n_pix * (hidden / 4);
  const int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)kNumSMs * 16);
  lstm_gate_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(pre, c_prev, n_pix, hidden, h_out, c_out);
  return after_launch("lstm_gate_kernel");
}

// Stateless LayerGens of the reference (models/modules/layer_gen.py:176-194 Up, :255-284 ReLU / SiLU / Tanh) as
// elementwise kernels: activation over a dense tensor, nearest-neighbour upsampling of pixel-major frames.
namespace soda {
template <int MODE>   // 0 relu, 1 silu (x * sigmoid(x)), 2 tanh
__global__ void __launch_bounds__(256) unary_kernel(const float* __restrict__ x, int64_t n, float* __restrict__ y) {
  pdl_trigger();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = x[i];
    float o;
    if (MODE == 0) o = v > 0.0f ? v : 0.0f;
    else if (MODE == 1) o = __fmul_rn(v, __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-v))));
    else o = tanhf(v);
    y[i] = o;
  }
}

// x [N, H, W, C] pixel-major -> y [N, H*s, W*s, C]: y(n, oy, ox, :) = x(n, oy / s, ox / s, :)
__global__ void __launch_bounds__(256)
upsample_nearest_kernel(const float* __restrict__ x, int64_t N, int H, int W, int C, int s, float* __restrict__ y) {
  pdl_trigger();
  const int Ho = H * s, Wo = W * s;
  const int64_t total = N * Ho * Wo * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    int64_t r = i / C;
    const int ox = (int)(r % Wo);
    r /= Wo;
    const int oy = (int)(r % Ho);
    const int64_t n = r / Ho;
    y[i] = __ldg(x + ((n * H + oy / s) * W + ox / s) * C + c);
  }
}
}  // namespace soda

extern "C" int soda_unary_fwd(const float* x, int64_t n, int mode, float* y, void* stream) {
  using namespace soda;
  if (!x || !y) return fail(SODA_ERR_BAD_ARG, "unary_fwd: null pointer");
  if (n <= 0 || mode < 0 || mode > 2) return fail(SODA_ERR_BAD_ARG, "unary_fwd: bad size or mode");
  const int grid = (int)std::min<int64_t>((n + 255) / 256, (int64_t)kNumSMs * 32);
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == 0) unary_kernel<0><<<grid, 256, 0, st>>>(x, n, y);
  else if (mode == 1) unary_kernel<1><<<grid, 256, 0, st>>>(x, n, y);
  else unary_kernel<2><<<grid, 256, 0, st>>>(x, n, y);
  return after_launch("unary_kernel");
}

extern "C" int soda_upsample_nearest(const float* x, int64_t N, int H, int W, int C, int scale, float* y, void* stream) {
  using namespace soda;
  if (!x || !y) return fail(SODA_ERR_BAD_ARG, "upsample_nearest: null pointer");
  if (N <= 0 || H <= 0 || W <= 0 || C <= 0 || scale < 1) return fail(SODA_ERR_BAD_ARG, "upsample_nearest: bad size");
  const int64_t total = N * H * scale * W * scale * C;
  const int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)kNumSMs * 32);
  upsample_nearest_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, N, H, W, C, scale, y);
  return after_launch("upsample_nearest_kernel");
}

