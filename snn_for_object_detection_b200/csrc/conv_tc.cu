// tcgen05 / TMEM / TMA implicit-GEMM convolution (soda_conv2d_fwd kind SODA_CONV_TC).
//
// Reference arithmetic: nn.Conv2d(bias=False, padding=k/2) models/modules/layer_gen.py:129-136 with
// the eval BatchNorm affine (:211-214), Residual sum (models/generator.py:146) and concat-slice
// store (:158) in the epilogue.
//
//   D[m, co] = sum_{tap, ci} A_tap[m, ci] * Wp[tap][co][ci]      m = output pixel, 128 per tile
//
// * A tiles (128 pixels x 32 channels fp32 = 128 rows x 128 B, SWIZZLE_128B) are fetched by TMA
//   straight from the pixel-major activation tensor through a 4-D tiled tensor map (C, W, H, N):
//   the box is 32 channels x BW x BH pixels of one image, shifted by the filter tap; out-of-bounds
//   coordinates are zero-filled by the TMA unit, which IS the convolution's zero padding; stride-2
//   convolutions use the map's element strides.  1x1 convolutions view the tensor as a flat list
//   of pixels (box 32 x 128).
// * B tiles (BN output channels x 32 input channels) come from the packed weights.
// * Precision: fp32-level results from the TF32 tensor pipe by operand splitting.  The packed
//   weights hold hi = tf32(w) and lo = w - hi; each A tile is split the same way; D += A_hi*B_hi + A_lo*B_hi + A_hi*B_lo with fp32 accumulation in TMEM
//   ("3xTF32").  When the caller declares the input exactly representable in TF32 (spikes, pooled
//   spike counts) the A split and the A_lo pass are skipped (2 passes).
// * The A operand of the MMAs lives in TENSOR MEMORY (tcgen05.mma "TS" form): a transform
//   warpgroup reads each fp32 A tile from shared memory once, splits it into tf32 hi / lo and
//   writes both to TMEM with tcgen05.st.  The 12 MMAs of a K block then read A from TMEM and only
//   B from shared memory, which keeps small-N layers (Cout 32/64) off the shared-memory roof.
// * Persistent, warp-specialised: warp 0 = TMA producer, warp 1 = MMA issuer (single thread) and
//   TMEM owner, warps 2-5 = A split -> TMEM, warps 6-9 = epilogue (tcgen05.ld -> +shift
//   -> +residual -> global store).  For BN <= 64 two TMEM accumulator stages let the epilogue of
//   tile i overlap the main loop of tile i+1.
#include "common.cuh"
#include "tc_common.cuh"

namespace soda {

using namespace tc;

constexpr int TC_BM = 128;
constexpr int TC_BK = 32;
constexpr int TC_A_BYTES = TC_BM * TC_BK * 4;  // 16 KB
constexpr int TC_THREADS = 320;

struct TcParams {
  int N, H, W, Cin, Cout, k, stride, pad, Ho, Wo;
  int BH, BW;                      // spatial tile, BH*BW == 128 (flat mode: 1 x 128)
  int tiles_x, tiles_y;            // tiles per image (flat mode: tiles_x = all tiles, tiles_y = 1)
  int num_m_tiles, num_n_tiles;
  int flat;                        // 1x1 stride-1 conv over the flattened pixel list
  int taps, cblocks;               // k*k, Cin/32
  int a_exact;
  int64_t total_pix;               // N*Ho*Wo
  int64_t out_bs, out_ps, res_bs, res_ps;
  const float* shift;
  const float* residual;
  float* y;
};

template <int BN>
struct TcSmem {
  static constexpr int B_BYTES = BN * TC_BK * 4;
  static constexpr int STAGE_BYTES = TC_A_BYTES + 2 * B_BYTES;  // A (fp32), B_hi, B_lo
  static constexpr int STAGES = 4;                              // = A-operand stages in TMEM
  static constexpr int BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
};

// TMEM map (512 columns): [0,256) accumulators, [256,512) A operand, 4 stages x (hi 32 | lo 32).
// Accumulator stage a: main at a*2*BN, correction at a*2*BN + BN.  BN=128 -> one stage.
constexpr uint32_t TC_TMEM_COLS = 512;
constexpr uint32_t TC_TMEM_A0 = 256;

template <int BN>
__global__ void __launch_bounds__(TC_THREADS, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, TcParams p) {
  using S = TcSmem<BN>;
  constexpr int STAGES = S::STAGES;
  constexpr int ACC_STAGES = BN == 128 ? 1 : 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * S::STAGE_BYTES);
  uint64_t* full_bar = bars;                 // TMA bytes landed                  [STAGES]
  uint64_t* ready_bar = bars + STAGES;       // A split written to TMEM           [STAGES]
  uint64_t* empty_bar = bars + 2 * STAGES;   // MMAs reading the stage completed  [STAGES]
  uint64_t* acc_full = bars + 3 * STAGES;    // accumulator complete              [2]
  uint64_t* acc_empty = bars + 3 * STAGES + 2;  // accumulator drained            [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * STAGES + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.num_m_tiles * p.num_n_tiles;
  const int kblocks = p.taps * p.cblocks;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&ready_bar[s], 128);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&acc_full[a], 1);
      mbar_init(&acc_empty[a], 128);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TC_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  auto stage_a = [&](int s) { return smem + s * S::STAGE_BYTES; };
  auto stage_bhi = [&](int s) { return smem + s * S::STAGE_BYTES + TC_A_BYTES; };
  auto stage_blo = [&](int s) { return smem + s * S::STAGE_BYTES + TC_A_BYTES + S::B_BYTES; };

  if (warp == 0) {
    // ================================ TMA producer ============================================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int nt = tile % p.num_n_tiles;
        const int mt = tile / p.num_n_tiles;
        int img = 0, iy0 = 0, ix0 = 0;
        if (p.flat) {
          ix0 = mt * TC_BM;
        } else {
          const int per_img = p.tiles_x * p.tiles_y;
          img = mt / per_img;
          const int r = mt - img * per_img;
          iy0 = (r / p.tiles_x) * p.BH * p.stride - p.pad;
          ix0 = (r % p.tiles_x) * p.BW * p.stride - p.pad;
        }
        for (int kb = 0; kb < kblocks; ++kb) {
          const int tap = kb / p.cblocks;
          const int cb = kb - tap * p.cblocks;
          const int ky = tap / p.k, kx = tap - ky * p.k;
          mbar_wait(&empty_bar[s], ph ^ 1);
          mbar_expect_tx(&full_bar[s], TC_A_BYTES + 2 * S::B_BYTES);
          tma_load_4d(stage_a(s), &tmA, &full_bar[s], cb * TC_BK, ix0 + kx, iy0 + ky, img);
          tma_load_3d(stage_bhi(s), &tmB, &full_bar[s], cb * TC_BK, nt * BN, tap);
          tma_load_3d(stage_blo(s), &tmB, &full_bar[s], cb * TC_BK, nt * BN, p.taps + tap);
          if (++s == STAGES) s = 0, ph ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ==============================================
    if (lane == 0) {
      constexpr uint32_t idesc = idesc_tf32(TC_BM, BN);
      int s = 0;
      uint32_t ph = 0;
      int local = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
        const int acc = local % ACC_STAGES;
        const uint32_t acc_ph = (local / ACC_STAGES) & 1;
        mbar_wait(&acc_empty[acc], acc_ph ^ 1);
        tc_fence_after();
        const uint32_t d_main = tmem_base + acc * 2 * BN;  // sum of A_hi*B_hi
        const uint32_t d_corr = d_main + BN;               // sum of the 2^-11-sized split terms
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&full_bar[s], ph);    // B tiles landed
          mbar_wait(&ready_bar[s], ph);   // A_hi / A_lo written to TMEM
          tc_fence_after();
          const uint32_t a_hi = tmem_base + TC_TMEM_A0 + s * 64;
          const uint32_t a_lo = a_hi + 32;
          const uint64_t b_hi = smem_desc_sw128(smem_u32(stage_bhi(s)));
          const uint64_t b_lo = smem_desc_sw128(smem_u32(stage_blo(s)));
#pragma unroll
          for (int kk = 0; kk < TC_BK / 8; ++kk) {
            const uint64_t adv = (uint64_t)((kk * 32) >> 4);  // 8 tf32 = 32 B along K inside the 128 B row
            // The tensor pipe truncates the fp32 accumulator on every accumulation, so the error
            // grows with (#MMAs into an accumulator) x ulp(|accumulator|): the small split terms
            // get their own accumulator and are added to the main one once, in the epilogue.
            mma_tf32_ts(d_corr, a_hi + kk * 8, b_lo + adv, idesc, (kb | kk) != 0);
            if (!p.a_exact) mma_tf32_ts(d_corr, a_lo + kk * 8, b_hi + adv, idesc, 1);
            mma_tf32_ts(d_main, a_hi + kk * 8, b_hi + adv, idesc, (kb | kk) != 0);
          }
          mma_commit(&empty_bar[s]);
          if (kb == kblocks - 1) mma_commit(&acc_full[acc]);
          if (++s == STAGES) s = 0, ph ^= 1;
        }
      }
    }
  } else if (warp < 6) {
    // ============== A operand: shared memory (fp32) -> tf32 hi / lo split -> tensor memory =======
    const int q = warp & 3;            // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;     // tile row = TMEM lane
    int s = 0;
    uint32_t ph = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      for (int kb = 0; kb < kblocks; ++kb) {
        mbar_wait(&full_bar[s], ph);
        const uint8_t* arow = stage_a(s) + row * 128;
        uint32_t hi[32], lo[32];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          // SWIZZLE_128B: 16-byte chunk c of row r lives at chunk position c ^ (r & 7)
          const float4 v = *reinterpret_cast<const float4*>(arow + ((c ^ (row & 7)) << 4));
          const float f[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const uint32_t h = __float_as_uint(f[e]) & 0xFFFFE000u;
            hi[c * 4 + e] = h;
            lo[c * 4 + e] = __float_as_uint(f[e] - __uint_as_float(h));
          }
        }
        // stage s of the previous round has been consumed: the producer only refilled the shared
        // memory stage after the MMAs reading TMEM stage s committed (empty_bar)
        const uint32_t t_a = tmem_base + ((uint32_t)(q * 32) << 16) + TC_TMEM_A0 + s * 64;
        tmem_st_32x32(t_a, hi);
        if (!p.a_exact) tmem_st_32x32(t_a + 32, lo);
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&ready_bar[s]);
        if (++s == STAGES) s = 0, ph ^= 1;
      }
    }
  } else {
    // ================================ epilogue ================================================
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;
    int local = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++local) {
      const int acc = local % ACC_STAGES;
      const uint32_t acc_ph = (local / ACC_STAGES) & 1;
      const int nt = tile % p.num_n_tiles;
      const int mt = tile / p.num_n_tiles;
      // output pixel of this thread
      bool valid;
      int64_t out_off, res_off;
      if (p.flat) {
        const int64_t pix = (int64_t)mt * TC_BM + row;
        valid = pix < p.total_pix;
        const int64_t hw = (int64_t)p.Ho * p.Wo;
        const int64_t n = pix / hw, r = pix - n * hw;
        out_off = n * p.out_bs + r * p.out_ps;
        res_off = n * p.res_bs + r * p.res_ps;
      } else {
        const int per_img = p.tiles_x * p.tiles_y;
        const int img = mt / per_img;
        const int r = mt - img * per_img;
        const int oy = (r / p.tiles_x) * p.BH + row / p.BW;
        const int ox = (r % p.tiles_x) * p.BW + row % p.BW;
        valid = oy < p.Ho && ox < p.Wo;
        const int64_t rr = (int64_t)oy * p.Wo + ox;
        out_off = (int64_t)img * p.out_bs + rr * p.out_ps;
        res_off = (int64_t)img * p.res_bs + rr * p.res_ps;
      }
      mbar_wait(&acc_full[acc], acc_ph);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + acc * 2 * BN;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t r[32], rc[32];
        tmem_ld_32x32(t_row + c0, r);
        tmem_ld_32x32(t_row + BN + c0, rc);
        tmem_ld_wait();
        if (c0 + 32 >= BN) {
          // last TMEM read of this tile: hand the accumulator back before the global stores
          tc_fence_before();
          mbar_arrive(&acc_empty[acc]);
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(rc[j]));
        const int co0 = nt * BN + c0;
        if (valid && co0 < p.Cout) {
          float* dst = p.y + out_off + co0;
          const float* rs = p.residual ? p.residual + res_off + co0 : nullptr;
          const int nvalid = min(32, p.Cout - co0);
          const bool vec = nvalid == 32 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) &&
                           (rs == nullptr || (reinterpret_cast<uintptr_t>(rs) & 15) == 0);
          if (vec) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              float4 o;
              o.x = __uint_as_float(r[j + 0]), o.y = __uint_as_float(r[j + 1]);
              o.z = __uint_as_float(r[j + 2]), o.w = __uint_as_float(r[j + 3]);
              if (p.shift) {
                const float4 sh = __ldg(reinterpret_cast<const float4*>(p.shift + co0 + j));
                o.x += sh.x, o.y += sh.y, o.z += sh.z, o.w += sh.w;
              }
              if (rs) {
                const float4 rv = __ldg(reinterpret_cast<const float4*>(rs + j));
                o.x += rv.x, o.y += rv.y, o.z += rv.z, o.w += rv.w;
              }
              *reinterpret_cast<float4*>(dst + j) = o;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (j < nvalid) {
                float o = __uint_as_float(r[j]);
                if (p.shift) o += __ldg(p.shift + co0 + j);
                if (rs) o += __ldg(rs + j);
                dst[j] = o;
              }
            }
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TC_TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------
// host side: weight packing, tensor maps, dispatch
// ------------------------------------------------------------------------------------------------
static inline int tc_bn_for(int Cout) { return Cout <= 32 ? 32 : (Cout <= 64 ? 64 : 128); }
static inline int tc_cout_pad(int Cout) {
  const int bn = tc_bn_for(Cout);
  return (Cout + bn - 1) / bn * bn;
}

int64_t conv_tc_packed_elems(int Cout, int Cin, int k) {
  if (Cin % TC_BK != 0) return 0;
  return 2LL * k * k * tc_cout_pad(Cout) * Cin;
}

// [Cout, Cin, k, k] -> [plane][tap][Cout_pad][Cin]; plane 0 = tf32(w * scale), plane 1 = remainder
__global__ void pack_weights_tc_kernel(const float* __restrict__ w, const float* __restrict__ scale, int Cout,
                                       int Cin, int taps, int Cout_pad, float* __restrict__ wp) {
  const int64_t plane = (int64_t)taps * Cout_pad * Cin;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < plane;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int ci = (int)(i % Cin);
    const int64_t r = i / Cin;
    const int co = (int)(r % Cout_pad);
    const int tap = (int)(r / Cout_pad);
    float v = 0.f;
    if (co < Cout) {
      v = w[((int64_t)co * Cin + ci) * taps + tap];
      if (scale != nullptr) v = __fmul_rn(v, scale[co]);
    }
    const float hi = __uint_as_float(__float_as_uint(v) & 0xFFFFE000u);
    wp[i] = hi;
    wp[plane + i] = v - hi;
  }
}

int conv_tc_pack(const float* w, const float* scale, int Cout, int Cin, int k, float* packed, cudaStream_t st) {
  if (Cin % TC_BK != 0) return fail(SODA_ERR_UNSUPPORTED, "conv_tc: Cin must be a multiple of 32");
  const int64_t plane = (int64_t)k * k * tc_cout_pad(Cout) * Cin;
  pack_weights_tc_kernel<<<(unsigned)std::min<int64_t>(ceil_div(plane, 256), 4096), 256, 0, st>>>(
      w, scale, Cout, Cin, k * k, tc_cout_pad(Cout), packed);
  return after_launch("pack_weights_tc_kernel");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

static int encode_map(CUtensorMap* m, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides_b,
                      const cuuint32_t* box, const cuuint32_t* estr) {
  EncodeTiledFn fn = get_encode_fn();
  if (fn == nullptr) return fail(SODA_ERR_CUDA, "conv_tc: cuTensorMapEncodeTiled entry point not available");
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<void*>(base), dims, strides_b, box,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("conv_tc: cuTensorMapEncodeTiled failed (%d)", (int)r);
    return (int)SODA_ERR_CUDA;
  }
  return 0;
}

template <int BN>
static int launch_tc(const CUtensorMap& tmA, const CUtensorMap& tmB, const TcParams& p, cudaStream_t st) {
  using S = TcSmem<BN>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(conv_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::BYTES);
    if (e != cudaSuccess) {
      set_error("conv_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return (int)SODA_ERR_CUDA;
    }
    attr_set = true;
  }
  const int tiles = p.num_m_tiles * p.num_n_tiles;
  const int grid = tiles < kNumSMs ? tiles : kNumSMs;
  conv_tc_kernel<BN><<<grid, TC_THREADS, S::BYTES, st>>>(tmA, tmB, p);
  return after_launch("conv_tc_kernel");
}

bool conv_tc_supported(const soda_conv_desc_t* d, const float* x) {
  if (d->in_nchw || d->Cin % TC_BK != 0) return false;
  if (d->in_pix_stride % 4 != 0 || d->in_batch_stride % 4 != 0) return false;
  if ((reinterpret_cast<uintptr_t>(x) & 15) != 0) return false;
  if (d->k != 1 && d->k != 3) return false;
  if (d->stride != 1 && d->stride != 2) return false;
  return true;
}

int conv_tc_dispatch(const soda_conv_desc_t* d, const float* x, const float* packed, const float* shift,
                     const float* residual, float* y, void* stream) {
  if (!conv_tc_supported(d, x))
    return fail(SODA_ERR_UNSUPPORTED,
                "conv_tc: needs pixel-major input, Cin % 32 == 0, 16-byte aligned strides, k in {1,3}, stride in {1,2}");
  TcParams p;
  p.N = d->N, p.H = d->H, p.W = d->W, p.Cin = d->Cin, p.Cout = d->Cout, p.k = d->k, p.stride = d->stride;
  p.pad = d->k / 2;
  p.Ho = (d->H + 2 * p.pad - d->k) / d->stride + 1;
  p.Wo = (d->W + 2 * p.pad - d->k) / d->stride + 1;
  if (p.Ho <= 0 || p.Wo <= 0) return fail(SODA_ERR_BAD_ARG, "conv: empty output");
  p.taps = d->k * d->k;
  p.cblocks = d->Cin / TC_BK;
  p.a_exact = (d->flags & SODA_CONV_FLAG_INPUT_EXACT_TF32) ? 1 : 0;
  p.total_pix = (int64_t)p.N * p.Ho * p.Wo;
  p.out_bs = d->out_batch_stride, p.out_ps = d->out_pix_stride;
  p.res_bs = d->res_batch_stride, p.res_ps = d->res_pix_stride;
  p.shift = shift, p.residual = residual, p.y = y;
  const int BN = tc_bn_for(d->Cout);
  const int Cout_pad = tc_cout_pad(d->Cout);
  p.num_n_tiles = Cout_pad / BN;

  CUtensorMap tmA, tmB;
  const int64_t HW = (int64_t)d->H * d->W;
  p.flat = (d->k == 1 && d->stride == 1 && (d->N == 1 || d->in_batch_stride == HW * d->in_pix_stride)) ? 1 : 0;
  if (p.flat) {
    p.BH = 1, p.BW = TC_BM;
    p.num_m_tiles = (int)ceil_div(p.total_pix, TC_BM);
    p.tiles_x = p.num_m_tiles, p.tiles_y = 1;
    const cuuint64_t dims[4] = {(cuuint64_t)d->Cin, (cuuint64_t)p.total_pix, 1, 1};
    const cuuint64_t str[3] = {(cuuint64_t)d->in_pix_stride * 4, (cuuint64_t)p.total_pix * d->in_pix_stride * 4,
                               (cuuint64_t)p.total_pix * d->in_pix_stride * 4};
    const cuuint32_t box[4] = {TC_BK, TC_BM, 1, 1};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    int rc = encode_map(&tmA, x, 4, dims, str, box, es);
    if (rc) return rc;
  } else {
    static const int shapes[6][2] = {{8, 16}, {16, 8}, {4, 32}, {32, 4}, {2, 64}, {1, 128}};
    int64_t best = -1;
    for (int i = 0; i < 6; ++i) {
      const int64_t t = ceil_div(p.Ho, shapes[i][0]) * ceil_div(p.Wo, shapes[i][1]);
      if (best < 0 || t < best) best = t, p.BH = shapes[i][0], p.BW = shapes[i][1];
    }
    p.tiles_y = (int)ceil_div(p.Ho, p.BH), p.tiles_x = (int)ceil_div(p.Wo, p.BW);
    p.num_m_tiles = p.N * p.tiles_x * p.tiles_y;
    const cuuint64_t dims[4] = {(cuuint64_t)d->Cin, (cuuint64_t)d->W, (cuuint64_t)d->H, (cuuint64_t)d->N};
    const cuuint64_t str[3] = {(cuuint64_t)d->in_pix_stride * 4, (cuuint64_t)d->W * d->in_pix_stride * 4,
                               (cuuint64_t)(d->N == 1 ? HW * d->in_pix_stride : d->in_batch_stride) * 4};
    const cuuint32_t box[4] = {TC_BK, (cuuint32_t)(p.BW * d->stride), (cuuint32_t)(p.BH * d->stride), 1};
    const cuuint32_t es[4] = {1, (cuuint32_t)d->stride, (cuuint32_t)d->stride, 1};
    int rc = encode_map(&tmA, x, 4, dims, str, box, es);
    if (rc) return rc;
  }
  {
    const cuuint64_t dims[3] = {(cuuint64_t)d->Cin, (cuuint64_t)Cout_pad, (cuuint64_t)(2 * p.taps)};
    const cuuint64_t str[2] = {(cuuint64_t)d->Cin * 4, (cuuint64_t)Cout_pad * d->Cin * 4};
    const cuuint32_t box[3] = {TC_BK, (cuuint32_t)BN, 1};
    const cuuint32_t es[3] = {1, 1, 1};
    int rc = encode_map(&tmB, packed, 3, dims, str, box, es);
    if (rc) return rc;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (BN == 32) return launch_tc<32>(tmA, tmB, p, st);
  if (BN == 64) return launch_tc<64>(tmA, tmB, p, st);
  return launch_tc<128>(tmA, tmB, p, st);
}

}  // namespace soda
