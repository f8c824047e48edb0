// Train-mode BatchNorm over T time steps (soda_bn_train_fwd / soda_bn_train_bwd): SURVEY.md 8(a)-7.
//
// Reference: nn.BatchNorm2d(C) in train mode (models/modules/layer_gen.py:197-214) is called once PER TIME STEP inside
// the loop of models/soda.py:142, so every step t normalises its B frames with that step's own batch statistics over
// (B, H, W) and updates the running statistics once:
//     y = (x - mean_t) * rsqrt(var_t + eps) * gamma + beta                   var_t biased (divide by n = B*H*W)
//     running_mean = (1 - m) * running_mean + m * mean_t
//     running_var  = (1 - m) * running_var  + m * var_t * n / (n - 1)          for t = 0 .. T-1 in order
// and autograd gives, per time step,
//     dx = gamma * invstd_t * (gy - mean(gy) - xhat * mean(gy * xhat)),  dgamma = sum_t sum gy * xhat,  dbeta = sum_t sum gy.
// Layer-major execution hands us all T steps at once: x is [T*B, H, W, C] pixel-major (frame t*B + b), so the B frames
// of a step are one contiguous run of n * C floats.  Three HBM-bound kernels per direction instead of T library calls:
//   bn_partial_kernel   grid (S, T): per-channel partial sums of (x, x*x) or (gy, gy*xhat) of one slice of one step,
//                       4 channels per thread as one 16-byte load, fixed-order block tree -> deterministic
//   bn_finalize_*       one thread per (t, c): partials combined in fp64 in slice order, per-(t, c) coefficients; the
//                       running-statistics chain / the sum over t of the parameter gradients by one row per channel
//   bn_apply_*          elementwise  y = x*a + b   /   dx = gy*c1 + x*c2 + c3   with per-(t, c) coefficients
// Algorithmic bytes: forward 12 B per element (x twice, y once), backward 20 B (gy, x twice each, dx once).
#include "common.cuh"

namespace soda {

constexpr int BN_THREADS = 256;
constexpr int BN_MAX_SLICES = 64;

// MODE 0: sums of (x, x^2).  MODE 1: sums of (g, g * (x - mean) * invstd) with mean / invstd of step t.
template <int MODE>
__global__ void __launch_bounds__(BN_THREADS)
bn_partial_kernel(const float* __restrict__ x, const float* __restrict__ g, const float* __restrict__ mean,
                  const float* __restrict__ invstd, int64_t n_pix, int C, float* __restrict__ partial /*[T][S][2][C]*/) {
  pdl_trigger();
  const int t = blockIdx.y, s = blockIdx.x, S = gridDim.x;
  const int c4n = C >> 2;                              // float4 channel groups
  const int cg = threadIdx.x % c4n;
  const int prow = threadIdx.x / c4n, pstride = BN_THREADS / c4n;   // host guarantees c4n divides BN_THREADS
  const int64_t p0 = n_pix * s / S, p1 = n_pix * (s + 1) / S;
  const float4* xb = reinterpret_cast<const float4*>(x + (int64_t)t * n_pix * C);
  const float4* gb = MODE == 1 ? reinterpret_cast<const float4*>(g + (int64_t)t * n_pix * C) : nullptr;
  float4 mu = make_float4(0.f, 0.f, 0.f, 0.f), is = mu;
  if (MODE == 1) {
    mu = reinterpret_cast<const float4*>(mean + (int64_t)t * C)[cg];
    is = reinterpret_cast<const float4*>(invstd + (int64_t)t * C)[cg];
  }
  float a0[4] = {0.f, 0.f, 0.f, 0.f}, a1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
  for (int64_t p = p0 + prow; p < p1; p += pstride) {   // same summation order as the rolled loop, 4 loads in flight
    const float4 xv = __ldg(xb + p * c4n + cg);
    if (MODE == 0) {
      a0[0] += xv.x, a0[1] += xv.y, a0[2] += xv.z, a0[3] += xv.w;
      a1[0] = fmaf(xv.x, xv.x, a1[0]), a1[1] = fmaf(xv.y, xv.y, a1[1]);
      a1[2] = fmaf(xv.z, xv.z, a1[2]), a1[3] = fmaf(xv.w, xv.w, a1[3]);
    } else {
      const float4 gv = __ldg(gb + p * c4n + cg);
      a0[0] += gv.x, a0[1] += gv.y, a0[2] += gv.z, a0[3] += gv.w;
      a1[0] = fmaf(gv.x, (xv.x - mu.x) * is.x, a1[0]), a1[1] = fmaf(gv.y, (xv.y - mu.y) * is.y, a1[1]);
      a1[2] = fmaf(gv.z, (xv.z - mu.z) * is.z, a1[2]), a1[3] = fmaf(gv.w, (xv.w - mu.w) * is.w, a1[3]);
    }
  }
  // fixed-order tree over the pixel rows of the block (deterministic)
  __shared__ float red[BN_THREADS * 8];
#pragma unroll
  for (int e = 0; e < 4; ++e) red[threadIdx.x * 8 + e] = a0[e], red[threadIdx.x * 8 + 4 + e] = a1[e];
  __syncthreads();
  for (int half = pstride >> 1; half > 0; half >>= 1) {
    if (prow < half) {
#pragma unroll
      for (int e = 0; e < 8; ++e) red[threadIdx.x * 8 + e] += red[(threadIdx.x + half * c4n) * 8 + e];
    }
    __syncthreads();
  }
  if (prow == 0) {
    float* dst = partial + (((int64_t)t * S + s) * 2) * C + cg * 4;
#pragma unroll
    for (int e = 0; e < 4; ++e) dst[e] = red[threadIdx.x * 8 + e], dst[C + e] = red[threadIdx.x * 8 + 4 + e];
  }
}

// Finalize kernels: block = 32 channels x BN_FT time rows; thread (ty, c) combines the S partials of its (t, c) in
// fp64 in slice order; the sequential part (running-statistics chain / sum over t) is done by row 0 from shared memory.
constexpr int BN_FT = 32;

// forward: mean / invstd per (t, c), affine coefficients, running statistics chained over t
__global__ void __launch_bounds__(32 * BN_FT)
bn_finalize_fwd_kernel(const float* __restrict__ partial, const double* __restrict__ sums /* [T][2][C] or NULL */,
                       int T, int S, int C, int64_t n_pix, float eps,
                       float momentum, const float* __restrict__ gamma, const float* __restrict__ beta,
                       float* __restrict__ mean, float* __restrict__ invstd, float* __restrict__ scale,
                       float* __restrict__ shift, float* __restrict__ running_mean, float* __restrict__ running_var) {
  __shared__ float s_m[BN_FT][32], s_v[BN_FT][32];
  const int cl = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cl;
  const bool live = c < C;
  const double n = (double)n_pix;
  const float gm = (live && gamma) ? gamma[c] : 1.0f, bt = (live && beta) ? beta[c] : 0.0f;
  float rm = (live && running_mean) ? running_mean[c] : 0.f, rv = (live && running_var) ? running_var[c] : 0.f;
  for (int t0 = 0; t0 < T; t0 += BN_FT) {
    const int t = t0 + ty;
    if (live && t < T) {
      double s0 = 0.0, s1 = 0.0;
      if (sums != nullptr) {
        s0 = sums[((int64_t)t * 2) * C + c], s1 = sums[((int64_t)t * 2 + 1) * C + c];
      } else {
        for (int s = 0; s < S; ++s) {
          const float* p = partial + (((int64_t)t * S + s) * 2) * C;
          s0 += (double)p[c], s1 += (double)p[C + c];
        }
      }
      const double m = s0 / n;
      double var = s1 / n - m * m;
      var = var < 0.0 ? 0.0 : var;
      const float is = (float)(1.0 / sqrt(var + (double)eps));
      mean[(int64_t)t * C + c] = (float)m;
      invstd[(int64_t)t * C + c] = is;
      const float a = gm * is;
      scale[(int64_t)t * C + c] = a;
      shift[(int64_t)t * C + c] = bt - (float)m * a;
      s_m[ty][cl] = (float)m;
      s_v[ty][cl] = (float)(n > 1.0 ? var * n / (n - 1.0) : var);
    }
    __syncthreads();
    if (ty == 0 && live && running_mean) {
      for (int k = 0; k < BN_FT && t0 + k < T; ++k) {
        rm = (1.0f - momentum) * rm + momentum * s_m[k][cl];
        rv = (1.0f - momentum) * rv + momentum * s_v[k][cl];
      }
    }
    __syncthreads();
  }
  if (ty == 0 && live && running_mean) running_mean[c] = rm, running_var[c] = rv;
}

// backward: dgamma / dbeta and the three per-(t, c) coefficients of dx = gy*c1 + x*c2 + c3
__global__ void __launch_bounds__(32 * BN_FT)
bn_finalize_bwd_kernel(const float* __restrict__ partial, const double* __restrict__ sums /* global, or NULL */,
                       const double* __restrict__ sums_local /* for dgamma / dbeta, or NULL = same */,
                       int T, int S, int C, int64_t n_pix,
                       const float* __restrict__ gamma, const float* __restrict__ mean,
                       const float* __restrict__ invstd, float* __restrict__ coef /*[T][3][C]*/,
                       float* __restrict__ dgamma, float* __restrict__ dbeta) {
  __shared__ double s_g[BN_FT][32], s_b[BN_FT][32];
  const int cl = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cl;
  const bool live = c < C;
  const double n = (double)n_pix;
  const double gm = (live && gamma) ? (double)gamma[c] : 1.0;
  double dg = 0.0, db = 0.0;
  for (int t0 = 0; t0 < T; t0 += BN_FT) {
    const int t = t0 + ty;
    double s0 = 0.0, s1 = 0.0, l0 = 0.0, l1 = 0.0;
    if (live && t < T) {
      if (sums != nullptr) {
        s0 = sums[((int64_t)t * 2) * C + c], s1 = sums[((int64_t)t * 2 + 1) * C + c];
      } else {
        for (int s = 0; s < S; ++s) {
          const float* p = partial + (((int64_t)t * S + s) * 2) * C;
          s0 += (double)p[c], s1 += (double)p[C + c];
        }
      }
      l0 = s0, l1 = s1;
      if (sums_local != nullptr) l0 = sums_local[((int64_t)t * 2) * C + c], l1 = sums_local[((int64_t)t * 2 + 1) * C + c];
      const double is = (double)invstd[(int64_t)t * C + c], mu = (double)mean[(int64_t)t * C + c];
      const double a = gm * is, m1 = s0 / n, m2 = s1 / n;
      float* k = coef + (int64_t)t * 3 * C;
      k[c] = (float)a;                              // * gy
      k[C + c] = (float)(-a * m2 * is);             // * x
      k[2 * C + c] = (float)(-a * m1 + a * m2 * is * mu);
    }
    s_b[ty][cl] = l0, s_g[ty][cl] = l1;
    __syncthreads();
    if (ty == 0) {
      for (int k = 0; k < BN_FT; ++k) db += s_b[k][cl], dg += s_g[k][cl];   // fixed order over t
    }
    __syncthreads();
  }
  if (ty == 0 && live) {
    if (dgamma) dgamma[c] = (float)dg;
    if (dbeta) dbeta[c] = (float)db;
  }
}

// y = x * a[t, c] + b[t, c]  (HAS_G = false)   /   dx = g * k1 + x * k2 + k3  (HAS_G = true, coef [T][3][C])
template <bool HAS_G>
__global__ void __launch_bounds__(256)
bn_apply_kernel(const float* __restrict__ x, const float* __restrict__ g, const float* __restrict__ ca,
                const float* __restrict__ cb, int64_t n_vec /* float4 per time step */, int C, int64_t total_vec,
                float* __restrict__ y) {
  pdl_trigger();
  const int c4n = C >> 2;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_vec; i += (int64_t)gridDim.x * blockDim.x) {
    const int t = (int)(i / n_vec);
    const int cg = (int)(i % c4n);
    const float4 xv = __ldg(reinterpret_cast<const float4*>(x) + i);
    float4 o;
    if (HAS_G) {
      const float4 gv = __ldg(reinterpret_cast<const float4*>(g) + i);
      const float4 k1 = reinterpret_cast<const float4*>(ca + (int64_t)t * 3 * C)[cg];
      const float4 k2 = reinterpret_cast<const float4*>(ca + (int64_t)t * 3 * C + C)[cg];
      const float4 k3 = reinterpret_cast<const float4*>(ca + (int64_t)t * 3 * C + 2 * C)[cg];
      o.x = fmaf(gv.x, k1.x, fmaf(xv.x, k2.x, k3.x)), o.y = fmaf(gv.y, k1.y, fmaf(xv.y, k2.y, k3.y));
      o.z = fmaf(gv.z, k1.z, fmaf(xv.z, k2.z, k3.z)), o.w = fmaf(gv.w, k1.w, fmaf(xv.w, k2.w, k3.w));
    } else {
      const float4 a = reinterpret_cast<const float4*>(ca + (int64_t)t * C)[cg];
      const float4 b = reinterpret_cast<const float4*>(cb + (int64_t)t * C)[cg];
      o.x = fmaf(xv.x, a.x, b.x), o.y = fmaf(xv.y, a.y, b.y), o.z = fmaf(xv.z, a.z, b.z), o.w = fmaf(xv.w, a.w, b.w);
    }
    reinterpret_cast<float4*>(y)[i] = o;
  }
}

// partial sums [T][S][2][C] -> double sums [T][2][C], slices added in order (deterministic)
__global__ void bn_reduce_kernel(const float* __restrict__ partial, int T, int S, int C, double* __restrict__ sums) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;   // over T * 2 * C
  if (i >= (int64_t)T * 2 * C) return;
  const int t = (int)(i / (2 * C));
  const int r = (int)(i % (2 * C));
  double a = 0.0;
  for (int s = 0; s < S; ++s) a += (double)partial[((int64_t)t * S + s) * 2 * C + r];
  sums[i] = a;
}

static int bn_slices(int64_t n_pix, int T) {
  // 8 resident CTAs per SM (28 registers, 8 KB of shared memory each): ncu showed 4 per SM at 52 % of the DRAM rate;
  // at least 64 pixels per slice row
  int S = (int)std::min<int64_t>(BN_MAX_SLICES, std::max<int64_t>(1, (8 * kNumSMs + T - 1) / T));
  while (S > 1 && n_pix / S < 64) S >>= 1;
  return S;
}

static int bn_check(int T, int64_t n_pix, int C, const void* a, const void* b, const void* c) {
  if (T <= 0 || n_pix <= 0 || C <= 0) return fail(SODA_ERR_BAD_ARG, "bn_train: non-positive size");
  if (C % 4 != 0 || BN_THREADS % (C / 4) != 0 || C / 4 > BN_THREADS)
    return fail(SODA_ERR_UNSUPPORTED, "bn_train: C must be 4 * a divisor of 256 (4, 8, 16, 32, 64, 128, 256, 512, 1024)");
  if ((reinterpret_cast<uintptr_t>(a) & 15) || (reinterpret_cast<uintptr_t>(b) & 15) || (reinterpret_cast<uintptr_t>(c) & 15))
    return fail(SODA_ERR_BAD_ARG, "bn_train: tensors must be 16-byte aligned");
  return 0;
}

}  // namespace soda

extern "C" int64_t soda_bn_train_workspace_bytes(int T, int C) {
  if (T <= 0 || C <= 0) return 0;
  // partial sums [T][S][2][C] + backward coefficients [T][3][C]
  return ((int64_t)T * soda::BN_MAX_SLICES * 2 * C + (int64_t)T * 3 * C) * 4;
}

extern "C" int soda_bn_train_fwd(const float* x, int T, int64_t n_pix, int C, const float* gamma, const float* beta,
                                 float eps, float momentum, float* running_mean, float* running_var, float* y,
                                 float* save_mean, float* save_invstd, void* workspace, void* stream) {
  using namespace soda;
  if (!x || !y || !save_mean || !save_invstd || !workspace) return fail(SODA_ERR_BAD_ARG, "bn_train_fwd: null pointer");
  if ((running_mean == nullptr) != (running_var == nullptr))
    return fail(SODA_ERR_BAD_ARG, "bn_train_fwd: running_mean / running_var must come together");
  int rc = bn_check(T, n_pix, C, x, y, workspace);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  float* partial = reinterpret_cast<float*>(workspace);
  float* scale = partial + (int64_t)T * BN_MAX_SLICES * 2 * C;   // [T][C] scale, [T][C] shift (reuses the coefficient area)
  float* shift = scale + (int64_t)T * C;
  const int S = bn_slices(n_pix, T);
  bn_partial_kernel<0><<<dim3(S, T), BN_THREADS, 0, st>>>(x, nullptr, nullptr, nullptr, n_pix, C, partial);
  rc = after_launch("bn_partial_kernel");
  if (rc) return rc;
  bn_finalize_fwd_kernel<<<(C + 31) / 32, 32 * BN_FT, 0, st>>>(partial, nullptr, T, S, C, n_pix, eps, momentum, gamma, beta, save_mean,
                                                          save_invstd, scale, shift, running_mean, running_var);
  rc = after_launch("bn_finalize_fwd_kernel");
  if (rc) return rc;
  const int64_t n_vec = n_pix * C / 4, total = n_vec * T;
  const int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)kNumSMs * 16);
  bn_apply_kernel<false><<<grid, 256, 0, st>>>(x, nullptr, scale, shift, n_vec, C, total, y);
  return after_launch("bn_apply_kernel");
}

extern "C" int soda_bn_train_bwd(const float* grad_y, const float* x, int T, int64_t n_pix, int C, const float* gamma,
                                 const float* save_mean, const float* save_invstd, float* grad_x, float* grad_gamma,
                                 float* grad_beta, void* workspace, void* stream) {
  using namespace soda;
  if (!grad_y || !x || !save_mean || !save_invstd || !grad_x || !workspace)
    return fail(SODA_ERR_BAD_ARG, "bn_train_bwd: null pointer");
  int rc = bn_check(T, n_pix, C, x, grad_y, grad_x);
  if (rc) return rc;
  if (reinterpret_cast<uintptr_t>(workspace) & 15) return fail(SODA_ERR_BAD_ARG, "bn_train: workspace must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  float* partial = reinterpret_cast<float*>(workspace);
  float* coef = partial + (int64_t)T * BN_MAX_SLICES * 2 * C;
  const int S = bn_slices(n_pix, T);
  bn_partial_kernel<1><<<dim3(S, T), BN_THREADS, 0, st>>>(x, grad_y, save_mean, save_invstd, n_pix, C, partial);
  rc = after_launch("bn_partial_kernel");
  if (rc) return rc;
  bn_finalize_bwd_kernel<<<(C + 31) / 32, 32 * BN_FT, 0, st>>>(partial, nullptr, nullptr, T, S, C, n_pix, gamma, save_mean, save_invstd, coef,
                                                          grad_gamma, grad_beta);
  rc = after_launch("bn_finalize_bwd_kernel");
  if (rc) return rc;
  const int64_t n_vec = n_pix * C / 4, total = n_vec * T;
  const int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)kNumSMs * 16);
  bn_apply_kernel<true><<<grid, 256, 0, st>>>(x, grad_y, coef, nullptr, n_vec, C, total, grad_x);
  return after_launch("bn_apply_kernel");
}

// ---- cross-rank (SyncBatchNorm) variant: statistics -> [caller all-reduces the sums] -> normalise ------------------
// The reference trains with sync_batchnorm (config.yaml:76): every time step's statistics are taken over the GLOBAL
// batch.  soda_bn_train_stats leaves the per-(t, c) sums of this rank in a small fp64 buffer [T][2][C]; the caller
// all-reduces it (one NCCL call per layer and direction, <= 128 KB) and passes the global sums and the global pixel
// count to the *_from_sums entry points.  dgamma / dbeta come from the LOCAL sums (the flat gradient all-reduce
// averages them over ranks like every other parameter gradient).
extern "C" int soda_bn_train_stats(const float* x, const float* grad_y, const float* save_mean, const float* save_invstd,
                                   int T, int64_t n_pix, int C, double* sums, void* workspace, void* stream) {
  using namespace soda;
  if (!x || !sums || !workspace) return fail(SODA_ERR_BAD_ARG, "bn_train_stats: null pointer");
  if (grad_y && (!save_mean || !save_invstd)) return fail(SODA_ERR_BAD_ARG, "bn_train_stats: backward sums need mean / invstd");
  int rc = bn_check(T, n_pix, C, x, grad_y, workspace);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  float* partial = reinterpret_cast<float*>(workspace);
  const int S = bn_slices(n_pix, T);
  if (grad_y) bn_partial_kernel<1><<<dim3(S, T), BN_THREADS, 0, st>>>(x, grad_y, save_mean, save_invstd, n_pix, C, partial);
  else bn_partial_kernel<0><<<dim3(S, T), BN_THREADS, 0, st>>>(x, nullptr, nullptr, nullptr, n_pix, C, partial);
  rc = after_launch("bn_partial_kernel");
  if (rc) return rc;
  const int64_t n = (int64_t)T * 2 * C;
  bn_reduce_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(partial, T, S, C, sums);
  return after_launch("bn_reduce_kernel");
}

extern "C" int soda_bn_train_fwd_from_sums(const float* x, const double* sums, int64_t n_total, int T, int64_t n_pix, int C,
                                           const float* gamma, const float* beta, float eps, float momentum,
                                           float* running_mean, float* running_var, float* y, float* save_mean,
                                           float* save_invstd, void* workspace, void* stream) {
  using namespace soda;
  if (!x || !sums || !y || !save_mean || !save_invstd || !workspace)
    return fail(SODA_ERR_BAD_ARG, "bn_train_fwd_from_sums: null pointer");
  if ((running_mean == nullptr) != (running_var == nullptr) || n_total < n_pix)
    return fail(SODA_ERR_BAD_ARG, "bn_train_fwd_from_sums: bad running statistics or global count");
  int rc = bn_check(T, n_pix, C, x, y, workspace);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  float* scale = reinterpret_cast<float*>(workspace) + (int64_t)T * BN_MAX_SLICES * 2 * C;
  float* shift = scale + (int64_t)T * C;
  bn_finalize_fwd_kernel<<<(C + 31) / 32, 32 * BN_FT, 0, st>>>(nullptr, sums, T, 0, C, n_total, eps, momentum, gamma, beta,
                                                              save_mean, save_invstd, scale, shift, running_mean, running_var);
  rc = after_launch("bn_finalize_fwd_kernel");
  if (rc) return rc;
  const int64_t n_vec = n_pix * C / 4, total = n_vec * T;
  const int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)kNumSMs * 16);
  bn_apply_kernel<false><<<grid, 256, 0, st>>>(x, nullptr, scale, shift, n_vec, C, total, y);
  return after_launch("bn_apply_kernel");
}

extern "C" int soda_bn_train_bwd_from_sums(const float* grad_y, const float* x, const double* sums_global, int64_t n_total,
                                           const double* sums_local, int T, int64_t n_pix, int C, const float* gamma,
                                           const float* save_mean, const float* save_invstd, float* grad_x,
                                           float* grad_gamma, float* grad_beta, void* workspace, void* stream) {
  using namespace soda;
  if (!grad_y || !x || !sums_global || !save_mean || !save_invstd || !grad_x || !workspace)
    return fail(SODA_ERR_BAD_ARG, "bn_train_bwd_from_sums: null pointer");
  if (n_total < n_pix) return fail(SODA_ERR_BAD_ARG, "bn_train_bwd_from_sums: global count below the local one");
  int rc = bn_check(T, n_pix, C, x, grad_y, grad_x);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  float* coef = reinterpret_cast<float*>(workspace) + (int64_t)T * BN_MAX_SLICES * 2 * C;
  bn_finalize_bwd_kernel<<<(C + 31) / 32, 32 * BN_FT, 0, st>>>(nullptr, sums_global, sums_local, T, 0, C, n_total, gamma,
                                                              save_mean, save_invstd, coef, grad_gamma, grad_beta);
  rc = after_launch("bn_finalize_bwd_kernel");
  if (rc) return rc;
  const int64_t n_vec = n_pix * C / 4, total = n_vec * T;
  const int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)kNumSMs * 16);
  bn_apply_kernel<true><<<grid, 256, 0, st>>>(x, grad_y, coef, nullptr, n_vec, C, total, grad_x);
  return after_launch("bn_apply_kernel");
}
