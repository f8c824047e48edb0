// Packed activation formats (include/soda_b200.h: SODA_ACT_F32 / PAIR / HALF): conversion between them
// (soda_act_convert) and the LIF scan whose residual input and output are in any of them (soda_lif_fmt_fwd).
//
// Reference call sites: the tensor handed from one LayerGen module to the next by models/generator.py:184-198;
// torch.cat into a Dense buffer (:158), the Residual sum (:146), norse LIFCell (models/modules/layer_gen.py:233).
// Both kernels are pure HBM streams: one thread owns one pixel's 8-channel piece (a 16- or 32-byte vector in every
// format), consecutive threads take consecutive pieces, so a warp reads and writes whole 128-byte lines.
#include <cuda_fp16.h>
#include <stdlib.h>

#include <algorithm>

#include "common.cuh"
#include "neuron.cuh"

namespace soda {

// Streams read once / written once: `make CFLAGS=-DSODA_ACT_STREAM_CS` turns the accesses of this file into cache-
// streaming ones (ld.global.cs / st.global.cs) - an A/B switch.  Measured on B200: the scans alone are unchanged
// (4.3 - 5.9 TB/s either way); inside the step they gain 4 % and the convolutions next to them lose as much (5.753 ->
// 5.738 ms summed, inside the run-to-run spread), so the default stays the read-only path.
#ifdef SODA_ACT_STREAM_CS
#define ACT_LD(p) __ldcs(p)
#define ACT_ST4(p, v) __stcs(p, v)
#else
#define ACT_LD(p) __ldg(p)
#define ACT_ST4(p, v) (*(p) = (v))
#endif

__device__ __forceinline__ uint32_t h2_pack(float lo, float hi) {
  uint32_t d;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
__device__ __forceinline__ float2 h2_unpack(uint32_t v) {
  return __half22float2(*reinterpret_cast<const __half2*>(&v));
}

// byte offset of the 8-channel piece starting at channel c (c % 8 == 0) inside a pixel
__device__ __forceinline__ int64_t piece_off(int fmt, int c) {
  if (fmt == SODA_ACT_F32) return (int64_t)c * 4;
  if (fmt == SODA_ACT_HALF) return (int64_t)c * 2;
  return (int64_t)(c & ~31) * 4 + (int64_t)(c & 31) * 2;   // PAIR: hi piece; lo piece 64 bytes further
}

__device__ __forceinline__ void load8(const uint8_t* px, int fmt, int c, float (&v)[8]) {
  const uint8_t* q = px + piece_off(fmt, c);
  if (fmt == SODA_ACT_F32) {
    const float4 a = ACT_LD(reinterpret_cast<const float4*>(q)), b = ACT_LD(reinterpret_cast<const float4*>(q + 16));
    v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w, v[4] = b.x, v[5] = b.y, v[6] = b.z, v[7] = b.w;
  } else {
    const uint4 h = ACT_LD(reinterpret_cast<const uint4*>(q));
    const uint32_t hw[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 f = h2_unpack(hw[k]);
      v[2 * k] = f.x, v[2 * k + 1] = f.y;
    }
    if (fmt == SODA_ACT_PAIR) {
      const uint4 l = ACT_LD(reinterpret_cast<const uint4*>(q + 64));
      const uint32_t lw[4] = {l.x, l.y, l.z, l.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 f = h2_unpack(lw[k]);
        v[2 * k] += f.x, v[2 * k + 1] += f.y;
      }
    }
  }
}

__device__ __forceinline__ unsigned int store8(uint8_t* px, int fmt, int c, const float (&v)[8]) {
  uint8_t* q = px + piece_off(fmt, c);
  if (fmt == SODA_ACT_F32) {
    ACT_ST4(reinterpret_cast<float4*>(q), make_float4(v[0], v[1], v[2], v[3]));
    ACT_ST4(reinterpret_cast<float4*>(q + 16), make_float4(v[4], v[5], v[6], v[7]));
    return 0;
  }
  unsigned int over = 0;
  uint32_t hi[4], lo[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    hi[k] = h2_pack(v[2 * k], v[2 * k + 1]);
    const float2 f = h2_unpack(hi[k]);
    lo[k] = h2_pack(v[2 * k] - f.x, v[2 * k + 1] - f.y);
    over += (fabsf(v[2 * k]) > 65504.f) + (fabsf(v[2 * k + 1]) > 65504.f);
  }
  ACT_ST4(reinterpret_cast<uint4*>(q), make_uint4(hi[0], hi[1], hi[2], hi[3]));
  if (fmt == SODA_ACT_PAIR) ACT_ST4(reinterpret_cast<uint4*>(q + 64), make_uint4(lo[0], lo[1], lo[2], lo[3]));
  return over;
}

__global__ void __launch_bounds__(256)
act_convert_kernel(const uint8_t* __restrict__ src, int src_fmt, int64_t src_bs, int64_t src_ps, uint8_t* __restrict__ dst,
                   int dst_fmt, int64_t dst_bs, int64_t dst_ps, int64_t n_frames, int64_t pix_per_frame, int C,
                   unsigned int* status) {
  pdl_trigger();
  const int pieces = C >> 3;
  const int64_t total = n_frames * pix_per_frame * pieces;
  unsigned int over = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int pc = (int)(i % pieces);
    const int64_t pix = i / pieces;
    const int64_t f = pix / pix_per_frame, q = pix - f * pix_per_frame;
    float v[8];
    load8(src + f * src_bs + q * src_ps, src_fmt, pc * 8, v);
    over += store8(dst + f * dst_bs + q * dst_ps, dst_fmt, pc * 8, v);
  }
  if (status != nullptr && over != 0) atomicAdd(status, over);
}

// CH-channel variants of the piece accessors (CH = 4: 16 / 8 bytes per access)
template <int CH>
__device__ __forceinline__ void loadp(const uint8_t* px, int fmt, int c, float (&v)[CH]) {
  if constexpr (CH == 8) {
    load8(px, fmt, c, v);
  } else {
    const uint8_t* q = px + piece_off(fmt, c);
    if (fmt == SODA_ACT_F32) {
      const float4 a = ACT_LD(reinterpret_cast<const float4*>(q));
      v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w;
    } else {
      const uint2 h = ACT_LD(reinterpret_cast<const uint2*>(q));
      const float2 f0 = h2_unpack(h.x), f1 = h2_unpack(h.y);
      v[0] = f0.x, v[1] = f0.y, v[2] = f1.x, v[3] = f1.y;
      if (fmt == SODA_ACT_PAIR) {
        const uint2 l = ACT_LD(reinterpret_cast<const uint2*>(q + 64));
        const float2 g0 = h2_unpack(l.x), g1 = h2_unpack(l.y);
        v[0] += g0.x, v[1] += g0.y, v[2] += g1.x, v[3] += g1.y;
      }
    }
  }
}
template <int CH>
__device__ __forceinline__ unsigned int storep(uint8_t* px, int fmt, int c, const float (&v)[CH]) {
  if constexpr (CH == 8) {
    return store8(px, fmt, c, v);
  } else {
    uint8_t* q = px + piece_off(fmt, c);
    if (fmt == SODA_ACT_F32) {
      ACT_ST4(reinterpret_cast<float4*>(q), make_float4(v[0], v[1], v[2], v[3]));
      return 0;
    }
    const uint32_t h0 = h2_pack(v[0], v[1]), h1 = h2_pack(v[2], v[3]);
    ACT_ST4(reinterpret_cast<uint2*>(q), make_uint2(h0, h1));
    if (fmt == SODA_ACT_PAIR) {
      const float2 f0 = h2_unpack(h0), f1 = h2_unpack(h1);
      ACT_ST4(reinterpret_cast<uint2*>(q + 64), make_uint2(h2_pack(v[0] - f0.x, v[1] - f0.y), h2_pack(v[2] - f1.x, v[3] - f1.y)));
    }
    return (fabsf(v[0]) > 65504.f) + (fabsf(v[1]) > 65504.f) + (fabsf(v[2]) > 65504.f) + (fabsf(v[3]) > 65504.f);
  }
}

// One thread: CH channels of one pixel for all T steps, state in registers; the loads of UNR steps are issued before the
// first of them is consumed.  CH = 8 for large maps (32-byte accesses), CH = 4 where the map alone would not fill the
// GPU with threads (small maps at small batch: twice the threads, half the registers).
template <int CH, int UNR>
__global__ void __launch_bounds__(256)
lif_fmt_fwd_kernel(NP p, const float* __restrict__ x, const float* __restrict__ v0, const float* __restrict__ i0, int T,
                   int64_t n_pix, int C, const uint8_t* __restrict__ residual, int res_fmt, int64_t res_ps,
                   uint8_t* __restrict__ out, int out_fmt, int64_t out_ps, float* __restrict__ v_out,
                   float* __restrict__ i_out, unsigned int* status) {
  pdl_trigger();
  const int pieces = C / CH;
  const int64_t total = n_pix * pieces;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c = (int)(i % pieces) * CH;
  const int64_t q = i / pieces;
  const int64_t soff = q * C + c;
  float v[CH], cur[CH];
#pragma unroll
  for (int k = 0; k < CH; k += 4) {
    if (v0 != nullptr) {
      const float4 a = ACT_LD(reinterpret_cast<const float4*>(v0 + soff + k));
      const float4 e = ACT_LD(reinterpret_cast<const float4*>(i0 + soff + k));
      v[k] = a.x, v[k + 1] = a.y, v[k + 2] = a.z, v[k + 3] = a.w;
      cur[k] = e.x, cur[k + 1] = e.y, cur[k + 2] = e.z, cur[k + 3] = e.w;
    } else {
      v[k] = v[k + 1] = v[k + 2] = v[k + 3] = p.v_leak;
      cur[k] = cur[k + 1] = cur[k + 2] = cur[k + 3] = 0.f;
    }
  }
  const int64_t step = n_pix * C;   // floats between two time steps of x
  unsigned int over = 0;
  for (int t0 = 0; t0 < T; t0 += UNR) {
    float xin[UNR][CH], rin[UNR][CH];
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      if (t0 + u < T) {
        const float* xp = x + (int64_t)(t0 + u) * step + soff;
#pragma unroll
        for (int k = 0; k < CH; k += 4) {
          const float4 a = ACT_LD(reinterpret_cast<const float4*>(xp + k));
          xin[u][k] = a.x, xin[u][k + 1] = a.y, xin[u][k + 2] = a.z, xin[u][k + 3] = a.w;
        }
        if (residual != nullptr) loadp<CH>(residual + ((int64_t)(t0 + u) * n_pix + q) * res_ps, res_fmt, c, rin[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      if (t0 + u < T) {
        float o[CH];
#pragma unroll
        for (int k = 0; k < CH; ++k) {
          float z, vd;
          lif_step(p, xin[u][k], v[k], cur[k], z, vd);
          o[k] = residual != nullptr ? z + rin[u][k] : z;
        }
        over += storep<CH>(out + ((int64_t)(t0 + u) * n_pix + q) * out_ps, out_fmt, c, o);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < CH; k += 4) {
    *reinterpret_cast<float4*>(v_out + soff + k) = make_float4(v[k], v[k + 1], v[k + 2], v[k + 3]);
    *reinterpret_cast<float4*>(i_out + soff + k) = make_float4(cur[k], cur[k + 1], cur[k + 2], cur[k + 3]);
  }
  if (status != nullptr && over != 0) atomicAdd(status, over);
}

// im2col of the binned event frames for the detector's first layer (models/tiny_yolo.py:16 -> Conv(64, 3, stride 2) on
// [T*B, 2, H, W]): K = Cin * 9 = 18 is below one MMA k-block, so the frames are unfolded ONCE into a pixel-major
// SODA_ACT_PAIR tensor with 32 "channels" per output pixel - channel (ky * 3 + kx) * Cin + ci = x[ci, oy*s+ky-1,
// ox*s+kx-1], zeros beyond Cin * 9 and in the lo halves (occupancy {0,1} is exact in fp16) - and the convolution
// becomes a 1x1 layer on the tensor-core kernel.  One thread writes one pixel's 128-byte row.
template <int C>
__global__ void __launch_bounds__(256)
im2col3_frames_kernel(const float* __restrict__ x, int N, int H, int W, int stride, int Ho, int Wo,
                      uint8_t* __restrict__ out, int64_t out_bs, int64_t out_ps) {
  pdl_trigger();
  const int64_t total = (int64_t)N * Ho * Wo;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int ox = (int)(i % Wo);
    const int64_t r = i / Wo;
    const int oy = (int)(r % Ho);
    const int64_t n = r / Ho;
    const float* xf = x + n * (int64_t)C * H * W;
    float v[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) v[k] = 0.0f;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int iy = oy * stride + ky - 1;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int ix = ox * stride + kx - 1;
        const bool inb = iy >= 0 && iy < H && ix >= 0 && ix < W;
#pragma unroll
        for (int ci = 0; ci < C; ++ci)
          v[(ky * 3 + kx) * C + ci] = inb ? ACT_LD(xf + ((int64_t)ci * H + iy) * W + ix) : 0.0f;
      }
    }
    uint8_t* dst = out + n * out_bs + ((int64_t)oy * Wo + ox) * out_ps;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint4 u;
      u.x = h2_pack(v[8 * j], v[8 * j + 1]), u.y = h2_pack(v[8 * j + 2], v[8 * j + 3]);
      u.z = h2_pack(v[8 * j + 4], v[8 * j + 5]), u.w = h2_pack(v[8 * j + 6], v[8 * j + 7]);
      *reinterpret_cast<uint4*>(dst + j * 16) = u;
      *reinterpret_cast<uint4*>(dst + 64 + j * 16) = make_uint4(0u, 0u, 0u, 0u);
    }
  }
}

static inline int fmt_bytes(int fmt) { return fmt == SODA_ACT_HALF ? 2 : 4; }
static inline bool fmt_ok(int fmt) { return fmt >= SODA_ACT_F32 && fmt <= SODA_ACT_HALF; }
// pointer / stride (in elements) alignment a format needs for the vector accesses above
static inline bool fmt_aligned(const void* p, int fmt, int64_t s0, int64_t s1) {
  const int al = fmt == SODA_ACT_PAIR ? 128 : 16;
  const int b = fmt_bytes(fmt);
  return (reinterpret_cast<uintptr_t>(p) % al) == 0 && (s0 * b) % al == 0 && (s1 * b) % al == 0;
}

}  // namespace soda

extern "C" int soda_act_convert(const void* src, int src_fmt, int64_t src_batch_stride, int64_t src_pix_stride, void* dst,
                                int dst_fmt, int64_t dst_batch_stride, int64_t dst_pix_stride, int64_t n_frames,
                                int64_t pix_per_frame, int C, unsigned int* status, void* stream) {
  using namespace soda;
  if (!src || !dst) return fail(SODA_ERR_BAD_ARG, "act_convert: null pointer");
  if (!fmt_ok(src_fmt) || !fmt_ok(dst_fmt)) return fail(SODA_ERR_BAD_ARG, "act_convert: unknown format");
  if (n_frames <= 0 || pix_per_frame <= 0 || C <= 0) return fail(SODA_ERR_BAD_ARG, "act_convert: non-positive size");
  const bool pair = src_fmt == SODA_ACT_PAIR || dst_fmt == SODA_ACT_PAIR;
  if (C % (pair ? 32 : 8) != 0) return fail(SODA_ERR_UNSUPPORTED, "act_convert: C must be a multiple of 32 (PAIR) / 8");
  if (!fmt_aligned(src, src_fmt, src_batch_stride, src_pix_stride) || !fmt_aligned(dst, dst_fmt, dst_batch_stride, dst_pix_stride))
    return fail(SODA_ERR_BAD_ARG, "act_convert: misaligned pointer or stride");
  const int64_t total = n_frames * pix_per_frame * (C / 8);
  const unsigned grid = (unsigned)std::min<int64_t>(ceil_div(total, 256), 148 * 16);
  act_convert_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const uint8_t*>(src), src_fmt, src_batch_stride * fmt_bytes(src_fmt), src_pix_stride * fmt_bytes(src_fmt),
      reinterpret_cast<uint8_t*>(dst), dst_fmt, dst_batch_stride * fmt_bytes(dst_fmt), dst_pix_stride * fmt_bytes(dst_fmt),
      n_frames, pix_per_frame, C, status);
  return after_launch("act_convert_kernel");
}

extern "C" int soda_lif_fmt_fwd(const soda_neuron_params_t* p, const float* x, const float* v0, const float* i0, int T,
                                int64_t n_pix, int C, const void* residual, int res_fmt, int64_t res_pix_stride, void* out,
                                int out_fmt, int64_t out_pix_stride, float* v_out, float* i_out, unsigned int* status,
                                void* stream) {
  using namespace soda;
  if (!p || !x || !out || !v_out || !i_out) return fail(SODA_ERR_BAD_ARG, "lif_fmt: null pointer");
  if ((v0 == nullptr) != (i0 == nullptr)) return fail(SODA_ERR_BAD_ARG, "lif_fmt: v0 / i0 must come together");
  if (T <= 0 || n_pix <= 0 || C <= 0) return fail(SODA_ERR_BAD_ARG, "lif_fmt: non-positive size");
  if (!fmt_ok(out_fmt) || (residual != nullptr && !fmt_ok(res_fmt))) return fail(SODA_ERR_BAD_ARG, "lif_fmt: unknown format");
  if (C % 32 != 0) return fail(SODA_ERR_UNSUPPORTED, "lif_fmt: C must be a multiple of 32");
  if ((reinterpret_cast<uintptr_t>(x) & 15) || !fmt_aligned(out, out_fmt, out_pix_stride, out_pix_stride) ||
      (residual != nullptr && !fmt_aligned(residual, res_fmt, res_pix_stride, res_pix_stride)))
    return fail(SODA_ERR_BAD_ARG, "lif_fmt: misaligned pointer or stride");
  static const int env_ch = []() { const char* e = getenv("SODA_SCAN_CH"); return e ? atoi(e) : 0; }();
  const int64_t total8 = n_pix * (C / 8);
  const bool ch4 = env_ch == 4 || (env_ch != 8 && total8 < 148LL * 1024);   // fewer than one full wave of 8-channel threads
  const int64_t total = ch4 ? n_pix * (C / 4) : total8;
  const uint8_t* r8 = reinterpret_cast<const uint8_t*>(residual);
  uint8_t* o8 = reinterpret_cast<uint8_t*>(out);
  const int64_t rps = res_pix_stride * fmt_bytes(res_fmt), ops_ = out_pix_stride * fmt_bytes(out_fmt);
  const unsigned grid = (unsigned)ceil_div(total, 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (ch4)
    lif_fmt_fwd_kernel<4, 4><<<grid, 256, 0, st>>>(to_np(p), x, v0, i0, T, n_pix, C, r8, res_fmt, rps, o8, out_fmt, ops_, v_out,
                                                   i_out, status);
  else
    lif_fmt_fwd_kernel<8, 4><<<grid, 256, 0, st>>>(to_np(p), x, v0, i0, T, n_pix, C, r8, res_fmt, rps, o8, out_fmt, ops_, v_out,
                                                   i_out, status);
  return after_launch("lif_fmt_fwd_kernel");
}

extern "C" int soda_im2col3_frames(const float* x, int N, int C, int H, int W, int stride, void* out, int64_t out_batch_stride,
                                   int64_t out_pix_stride, void* stream) {
  using namespace soda;
  if (!x || !out) return fail(SODA_ERR_BAD_ARG, "im2col3_frames: null pointer");
  if (N <= 0 || C <= 0 || H <= 0 || W <= 0 || stride <= 0) return fail(SODA_ERR_BAD_ARG, "im2col3_frames: non-positive size");
  if (C * 9 > 32) return fail(SODA_ERR_UNSUPPORTED, "im2col3_frames: Cin * 9 must fit one 32-channel group");
  if (!fmt_aligned(out, SODA_ACT_PAIR, out_batch_stride, out_pix_stride))
    return fail(SODA_ERR_BAD_ARG, "im2col3_frames: misaligned output");
  const int Ho = (H + 2 - 3) / stride + 1, Wo = (W + 2 - 3) / stride + 1;
  const int64_t total = (int64_t)N * Ho * Wo;
  const unsigned grid = (unsigned)std::min<int64_t>(ceil_div(total, 256), 148 * 32);
  uint8_t* o8 = reinterpret_cast<uint8_t*>(out);
  cudaStream_t st = (cudaStream_t)stream;
  if (C == 1) im2col3_frames_kernel<1><<<grid, 256, 0, st>>>(x, N, H, W, stride, Ho, Wo, o8, out_batch_stride * 4, out_pix_stride * 4);
  else if (C == 2) im2col3_frames_kernel<2><<<grid, 256, 0, st>>>(x, N, H, W, stride, Ho, Wo, o8, out_batch_stride * 4, out_pix_stride * 4);
  else im2col3_frames_kernel<3><<<grid, 256, 0, st>>>(x, N, H, W, stride, Ho, Wo, o8, out_batch_stride * 4, out_pix_stride * 4);
  return after_launch("im2col3_frames_kernel");
}
