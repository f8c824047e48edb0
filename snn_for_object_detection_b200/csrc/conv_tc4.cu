// tcgen05 / TMEM / TMA implicit-GEMM convolution over PRE-SPLIT activations (soda_pconv_fwd).
//
// Reference arithmetic: nn.Conv2d(bias=False, padding=k/2) models/modules/layer_gen.py:129-136, the eval
// BatchNorm affine (:211-214), norse LIFCell (:233) iterated over the T steps of models/soda.py:142, the Residual
// sum (models/generator.py:146) and the concat-slice store (:158) in the epilogue.
//
//   D[m, co] = sum_{tap, ci} A_tap[m, ci] * W[tap][co][ci]          m = output pixel, 128 per tile
//
// What is different from conv_tc3.cu (fp32 activations, split in shared memory by 8 warps, tf32 + bf16 passes):
//
// * Activations live in HBM already split for the tensor core, written that way by the PRODUCING kernel's epilogue:
//     SODA_ACT_PAIR : per pixel and 32-channel group 128 B = fp16 hi[32] | fp16 lo[32], x = hi + lo (22 significant
//                     bits; the same 4 bytes per value as fp32, so a PAIR tensor has an fp32 tensor's strides)
//     SODA_ACT_HALF : fp16, for values that are exact in it (spikes {0,1}, spike counts <= 2048): 2 bytes per value
//   so TMA's bytes ARE the MMA operands: no split warps, no second shared-memory plane, no tf32 stream.
// * Products, all kind::f16 with fp32 accumulation in TMEM:
//     PAIR : a_hi*[w_hi | w_lo] + a_lo*w_hi         (4 MMAs per 32 channels and tap: the weight k-block is N-stacked,
//                                                    hi rows then lo rows, so one N = 2*BN MMA feeds the adjacent main |
//                                                    correction accumulators; a_lo*w_lo ~ 2^-22 is dropped)
//     HALF : a*[w_hi | w_lo]                         (4 MMAs per 64 channels and tap)
//   Weights are split the same way (fp16 hi | fp16 lo) after multiplication by the folded BatchNorm scale and by a
//   per-tensor power of two that puts them in fp16's normal range; the epilogue multiplies by its inverse (exact).
// * One weight k-block is BN x 128 B per 32 channels (was 2 x BN x 128 B): twice as many layers keep their weights
//   resident in shared memory, streamed layers move half the bytes from L2.
// * The epilogue emits F32, PAIR or HALF as the consumer wants (fp32 membrane state never leaves the registers in
//   the LIF variant) through a staged, fully coalesced store: 4 rows x 128 B per instruction.
//
// "im2col by descriptor" (one 18 x 10 halo tile per 32-channel slab, the nine taps are nine descriptors into it),
// the persistent item loop, the weight-stationary / streamed plans and programmatic dependent launch are those of
// conv_tc3.cu.  12 warps: w0-7 epilogue (two per TMEM lane quarter), w8 A producer (TMA), w9 B producer (TMA),
// w10 MMA issuer + TMEM owner, w11 idle.  The fused Conv -> BN -> LIF instantiations run 20 warps: w0-15 epilogue (four
// per lane quarter, 16 channels each - their neuron arithmetic is instruction-bound and needs the schedulers full),
// w16 / w17 producers, w18 issuer.
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include <cuda_fp16.h>

#include "common.cuh"
#include "neuron.cuh"
#include "tc_common.cuh"

namespace soda {

using namespace tc;

constexpr int T4_BM = 128;
constexpr int T4_BH = 16, T4_BW = 8;
constexpr int T4_MAX_RING = 8;
#ifdef SODA_TC4_EPI_DBG
constexpr bool T4_EPI_DBG = true;
#else
constexpr bool T4_EPI_DBG = false;
#endif
// accumulator stages in TMEM; a stage = main accumulator (hi * hi products) | correction accumulator (hi * lo, lo * hi):
// the tensor core truncates on every accumulate, an error proportional to the accumulator's magnitude and to the number
// of MMAs into it - the 2^-11-small corrections, two thirds of the MMAs, must not pay it at the main sum's magnitude
// (measured: one accumulator 4.1e-6 of the output scale at K = 1152, split 1.2e-6)
__host__ __device__ constexpr int t4_nacc(int BN) { return BN == 128 ? 2 : 4; }
constexpr int T4_EPI_BYTES = 4096;              // per epilogue warp: 32 rows x 128 B staging for the TMA store
constexpr int T4_SMEM_LIMIT = 232448 - 1024 /*alignment slack*/ - 1024 /*barriers*/ - 8 * T4_EPI_BYTES;

struct Tc4Params {
  int H, W, Cin, Cout, k, stride, pad, Ho, Wo;
  int tiles_x, tiles_y, n_spatial;
  int mt;                        // M-tiles per work item sharing every streamed weight k-block (1 or 2)
  int n_sp_items;
  int it_step_sp, it_step_outer;
  int taps, cblocks, kblocks;
  int mode;                      // 0 = HALO (one A stage per channel slab, `taps` k-blocks), 1 = TAP (one stage per k-block group)
  int kg;                        // k-blocks per A / B ring stage outside the 3x3 halo mode (1 or 2): every stage hand-off
                                 // (two barrier round trips + commits) then covers twice the MMAs
  uint32_t a_tile_bytes;         // one k-block's A tile inside a stage
  int flat;                      // 1x1 stride-1 over the flattened pixel list
  int in_fmt;                    // SODA_ACT_PAIR / SODA_ACT_HALF
  int a_exact;                   // PAIR input whose lo halves are all zero: the a_lo*w_hi product is skipped
  int rows, WW;
  uint32_t sbo, plane_bytes, a_tx_bytes;
  uint32_t b_kb_bytes;           // one weight k-block in shared memory: BN*128 (PAIR) or 2*BN*128 (HALF)
  int nch, n_outer, T_inner, Bsz, n_items;
  int b_stationary, b_stages, a_stages;
  uint32_t b_region_bytes;
  int64_t total_pix;
  int out_fmt, res_fmt;
  int64_t out_bs, out_ps, res_bs, res_ps;   // BYTES
  const float* shift;
  const float* descale;          // device scalar: inverse of the weights' power-of-two scale
  const uint8_t* residual;
  uint8_t* y;
  unsigned int* status;          // [0] counts values that left fp16's range when encoded (NULL = not counted)
  int dbg;                       // SODA_TC4_DBG timing experiments (results are garbage when set): 2 = no stores,
                                 // 4 = no epilogue arithmetic / stores, 8 = no A loads, 16 = no (streamed) B loads
  int lif;
  NP np;
  const float* v0;
  const float* i0;
  float* v_out;
  float* i_out;
};

struct Tc4Iter {
  int sp, outer;
  __device__ __forceinline__ void init(const Tc4Params& p) {
    const int r0 = (int)blockIdx.x / p.nch;
    sp = r0 % p.n_sp_items;
    outer = r0 / p.n_sp_items;
  }
  __device__ __forceinline__ void next(const Tc4Params& p) {
    sp += p.it_step_sp;
    outer += p.it_step_outer;
    if (sp >= p.n_sp_items) sp -= p.n_sp_items, ++outer;
  }
};

// kind::f16 with fp16 inputs, fp32 accumulate, both operands K-major
__host__ __device__ constexpr uint32_t idesc_f16(int M, int N) {
  return (1u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_h2(float lo, float hi) {   // lo -> bits [0,16)
  uint32_t d;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
  return d;
}
__device__ __forceinline__ float2 unpack_h2(uint32_t v) {
  const __half2 h = *reinterpret_cast<const __half2*>(&v);
  return __half22float2(h);
}
__device__ __forceinline__ uint4 ldg_u4(const uint8_t* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

// TMA store of a staged tile (shared -> global), bulk-group completion
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* m, uint32_t smem_src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_src), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all but the newest N bulk groups of this thread have finished READING their shared-memory source
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ================================ MMA issuer ===================================================
// The issuer is ONE warp and every tcgen05.mma needs its two 64-bit descriptors in uniform registers: measured
// (SODA_TC4_DBG=32), a k-block of 6 MMAs cost ~600 clk of issue time when the descriptors were assembled with 64-bit
// adds inside the elected-lane region - 2.5 x the 240 clk the tensor core needs for them at N = 32.  So: the upper
// descriptor words are kernel constants, the lower words (start address | leading byte offset) are 32-bit values
// that only ever get small immediates added (an address field never carries out of its 14 bits for offsets inside
// the shared-memory window), the tap loop is unrolled with compile-time tap offsets, and the operand format is a
// template parameter.
__device__ __forceinline__ void mma_f16_lo(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                           uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "mov.b64 da, {%1, %2};\n\t"
      "mov.b64 db, {%3, %4};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
      ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}

// One k-block (one 128-byte A row slab x one filter tap) of one tile.  The weight k-block in shared memory is
// [w_hi rows (BN) ; w_lo rows (BN)], so ONE MMA of N = 2*BN multiplies an A slice by both halves and lands in the main |
// correction accumulator pair (adjacent TMEM columns): the A slice is fetched from shared memory once for two
// products.  SS-mode MMAs are paced by operand fetch (128 B/clk of shared memory: measured 85 clk for an M128 N128 K16
// MMA that needs 64 clk of math), so this is what the products cost:
//   PAIR : a_hi x [w_hi ; w_lo] (N = 2BN) + a_lo x w_hi (N = BN)   2 + 2 MMAs per 32 channels (was 6)
//   HALF : a    x [w_hi ; w_lo] (N = 2BN)                          4 MMAs per 64 channels (was 8)
template <int BN, bool PAIR>
__device__ __forceinline__ void tc4_issue_kblock(uint32_t d_main, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo,
                                                 uint32_t b_hi, uint32_t accum, bool lo_pass) {
#ifdef SODA_TC4_NO_MMA   // timing experiment (build with -DSODA_TC4_NO_MMA): the issue loop without tensor work
  return;
#endif
  constexpr uint32_t idesc1 = idesc_f16(T4_BM, BN);
  constexpr uint32_t idesc2 = idesc_f16(T4_BM, 2 * BN);
  const uint32_t d_corr = d_main + BN;
  if constexpr (PAIR) {
    // A row = hi[0..15] hi[16..31] lo[0..15] lo[16..31] (128 B); B rows of 64 B = 32 channels (SWIZZLE_64B)
    mma_f16_lo(d_main, a_lo, a_hi, b_lo, b_hi, idesc2, accum);             // a_hi x [w_hi ; w_lo]
    mma_f16_lo(d_main, a_lo + 2, a_hi, b_lo + 2, b_hi, idesc2, 1);
    if (lo_pass) {
      mma_f16_lo(d_corr, a_lo + 4, a_hi, b_lo, b_hi, idesc1, 1);           // a_lo x w_hi
      mma_f16_lo(d_corr, a_lo + 6, a_hi, b_lo + 2, b_hi, idesc1, 1);
    }
  } else {
    // A row = a[0..63] (128 B); B rows of 128 B = 64 channels (SWIZZLE_128B)
    mma_f16_lo(d_main, a_lo, a_hi, b_lo, b_hi, idesc2, accum);
    mma_f16_lo(d_main, a_lo + 2, a_hi, b_lo + 2, b_hi, idesc2, 1);
    mma_f16_lo(d_main, a_lo + 4, a_hi, b_lo + 4, b_hi, idesc2, 1);
    mma_f16_lo(d_main, a_lo + 6, a_hi, b_lo + 6, b_hi, idesc2, 1);
  }
}

// HALO3: 3x3 stride-1 layers (one A stage = one channel slab with its halo, nine taps = nine descriptor offsets);
// otherwise one A stage = one k-block (1x1 layers, stride-2 tap tiles).
template <int BN, int MT, bool PAIR, int MODE>
__device__ __forceinline__ void tc4_mma_role(const Tc4Params& p, uint32_t tmem_base, uint32_t smem_a_u32,
                                             uint32_t smem_b_u32, uint64_t* a_full, uint64_t* a_empty,
                                             uint64_t* b_full, uint64_t* b_empty, uint64_t* acc_full,
                                             uint64_t* acc_empty) {
  constexpr uint32_t NACC = t4_nacc(BN);
  constexpr uint32_t B_KB16 = ((BN * 128) >> 4) * (PAIR ? 1 : 2);   // one weight k-block, in descriptor units
  constexpr bool HALO3 = MODE == 0;
  constexpr int WW = T4_BW + 2;                                      // halo row pitch in pixels
  constexpr int PW = T4_BW + 1;                                      // phase-plane row pitch in pixels (stride-2 layers)
  // BN = 32: the two epilogue warps of a lane quarter alternate WORK UNITS (item x m-tile), and a warp must see every
  // phase of a barrier it waits on - so the units of warpgroup e cycle through their own two stages {2e, 2e + 1}
  constexpr bool UNIT_SPLIT = BN == 32;
  const uint64_t a_desc0 = smem_desc_sw128_sbo(smem_a_u32, p.sbo);
  // weight rows: 64 B (32 channels, SWIZZLE_64B, 8-row groups 512 B apart) for PAIR input, 128 B (64 channels,
  // SWIZZLE_128B) for HALF input
  const uint64_t b_desc0 = PAIR ? smem_desc_swz(smem_b_u32, 512, 4) : smem_desc_sw128(smem_b_u32);
  const uint32_t a_hi = (uint32_t)(a_desc0 >> 32), b_hi = (uint32_t)(b_desc0 >> 32);
  const uint32_t a_lo0 = (uint32_t)a_desc0, b_lo0 = (uint32_t)b_desc0;
  const bool stat = p.b_stationary != 0;
  const bool lo_pass = PAIR && !p.a_exact;
  const int stages_per_tile = HALO3 ? p.cblocks : (MODE == 2 ? p.cblocks * 4 : p.kblocks / p.kg);
  const uint32_t a_stage16 = p.plane_bytes >> 4;
  const uint32_t b_tap16 = (uint32_t)p.cblocks * B_KB16;   // resident weights are stored (tap, cb)
  uint32_t sa = 0, pa = 0, sb = 0, pb = 0, cnt = 0;
  uint32_t unit = 0, ce0 = 0, ce1 = 0;
  if (stat) mbar_wait_uniform(&b_full[0], 0);
  const long long dbg_t0 = clock64();
  long long dbg_wa = 0, dbg_wb = 0, dbg_wc = 0;
  const bool dbg_on = (p.dbg & 32) != 0;
  for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, unit += MT) {
    for (int t = 0; t < p.T_inner; ++t) {
      uint32_t d_acc[2], acc_i[2];
      _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < MT) {
        uint32_t a, ph;
        if (UNIT_SPLIT) {
          const uint32_t e = (unit + m) & 1u;
          const uint32_t c = e ? ce1 : ce0;
          a = 2u * e + (c & 1u), ph = (c >> 1) & 1u;
          if (e) ++ce1; else ++ce0;
        } else {
          a = (cnt + m) % NACC, ph = ((cnt + m) / NACC) & 1u;
        }
        { const long long w0 = dbg_on ? clock64() : 0; mbar_wait_uniform(&acc_empty[a], ph ^ 1); if (dbg_on) dbg_wc += clock64() - w0; }
        acc_i[m] = a;
        d_acc[m] = tmem_base + a * (2 * BN);   // main; correction accumulator BN columns further
      }
      tc_fence_after();
      for (int st = 0; st < stages_per_tile; ++st) {
        uint32_t s_m[2], a_lo[2];
        _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < MT) {
          { const long long w0 = dbg_on ? clock64() : 0; mbar_wait_uniform(&a_full[sa], pa); if (dbg_on) dbg_wa += clock64() - w0; }
          a_lo[m] = a_lo0 + sa * a_stage16;
          s_m[m] = sa;
          if (++sa == (uint32_t)p.a_stages) sa = 0, pa ^= 1;
        }
        tc_fence_after();
        const uint32_t acc0 = st != 0;
        if constexpr (HALO3) {
          uint32_t b_lo = b_lo0 + (uint32_t)st * B_KB16;   // resident: k-block (tap 0, slab st)
          _Pragma("unroll") for (int tp = 0; tp < 9; ++tp) {
            if (!stat) {
              { const long long w0 = dbg_on ? clock64() : 0; mbar_wait_uniform(&b_full[sb], pb); if (dbg_on) dbg_wb += clock64() - w0; }
              tc_fence_after();
              b_lo = b_lo0 + sb * B_KB16;
            }
            constexpr int dummy = 0;
            (void)dummy;
            const uint32_t tap_off = (uint32_t)((tp / 3) * WW + (tp % 3)) * 8u;
            if (elect_one()) {
              _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < MT)
                tc4_issue_kblock<BN, PAIR>(d_acc[m], a_lo[m] + tap_off, a_hi, b_lo, b_hi, tp == 0 ? acc0 : 1u, lo_pass);
              if (!stat) mma_commit(&b_empty[sb]);
            }
            if (stat) b_lo += b_tap16;
            else if (++sb == (uint32_t)p.b_stages) sb = 0, pb ^= 1;
          }
        } else if constexpr (MODE == 2) {
          // stride-2 3x3 layer: stage st = parity plane (st & 3) of channel slab (st >> 2).  A plane holds every
          // second input row / column of the tile's halo (17 x 9 pixels); the taps that read it are descriptor shifts:
          //   plane 0 (even rows, even cols): tap (1,1)           plane 1 (even rows, odd cols): (1,0) (1,2)
          //   plane 2 (odd rows, even cols) : (0,1) (2,1)         plane 3 (odd rows, odd cols) : (0,0) (0,2) (2,0) (2,2)
          const int pl = st & 3, cb = st >> 2;
          const int ntap = pl == 0 ? 1 : (pl == 3 ? 4 : 2);
          for (int j = 0; j < ntap; ++j) {
            int ky, kx;
            if (pl == 0) ky = 1, kx = 1;
            else if (pl == 1) ky = 1, kx = 2 * j;
            else if (pl == 2) ky = 2 * j, kx = 1;
            else ky = 2 * (j >> 1), kx = 2 * (j & 1);
            const uint32_t tap_off = (uint32_t)((ky >> 1) * PW + (kx >> 1)) * 8u;
            uint32_t b_lo;
            if (stat) {
              b_lo = b_lo0 + (uint32_t)((ky * 3 + kx) * p.cblocks + cb) * B_KB16;
            } else {
              { const long long w0 = dbg_on ? clock64() : 0; mbar_wait_uniform(&b_full[sb], pb); if (dbg_on) dbg_wb += clock64() - w0; }
              tc_fence_after();
              b_lo = b_lo0 + sb * B_KB16;
            }
            if (elect_one()) {
              _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < MT)
                tc4_issue_kblock<BN, PAIR>(d_acc[m], a_lo[m] + tap_off, a_hi, b_lo, b_hi, (st == 0 && j == 0) ? 0u : 1u, lo_pass);
              if (!stat) mma_commit(&b_empty[sb]);
            }
            if (!stat && ++sb == (uint32_t)p.b_stages) sb = 0, pb ^= 1;
          }
        } else {
          uint32_t b_lo;
          if (stat) {
            b_lo = b_lo0 + (uint32_t)(st * p.kg) * B_KB16;
          } else {
            { const long long w0 = dbg_on ? clock64() : 0; mbar_wait_uniform(&b_full[sb], pb); if (dbg_on) dbg_wb += clock64() - w0; }
            tc_fence_after();
            b_lo = b_lo0 + sb * (B_KB16 * (uint32_t)p.kg);
          }
          const uint32_t a_tile16 = p.a_tile_bytes >> 4;
          if (elect_one()) {
            _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < MT)
              tc4_issue_kblock<BN, PAIR>(d_acc[m], a_lo[m], a_hi, b_lo, b_hi, acc0, lo_pass);
            if (p.kg == 2) {
              _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < MT)
                tc4_issue_kblock<BN, PAIR>(d_acc[m], a_lo[m] + a_tile16, a_hi, b_lo + B_KB16, b_hi, 1u, lo_pass);
            }
            if (!stat) mma_commit(&b_empty[sb]);
          }
          if (!stat && ++sb == (uint32_t)p.b_stages) sb = 0, pb ^= 1;
        }
        if (elect_one()) {
          _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < MT) mma_commit(&a_empty[s_m[m]]);
        }
      }
      if (elect_one()) {
        _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < MT) mma_commit(&acc_full[acc_i[m]]);
      }
      cnt += MT;
    }
  }
  if ((p.dbg & 32) && blockIdx.x == 0 && (threadIdx.x & 31) == 0)
    printf("tc4 mma warp: %lld clk total, waits: A %lld B %lld acc %lld; %u tiles, BN=%d MT=%d MODE=%d\n", clock64() - dbg_t0,
           dbg_wa, dbg_wb, dbg_wc, cnt, BN, MT, MODE);
}

// ---- the 16-warp epilogue of the fused Conv -> BN -> LIF kernel (NE = 16) ---------------------------------------
// Warp w: TMEM lane quarter q = w & 3 (32 pixels), column half h = (w >> 2) & 1, group g = w >> 3.
//   BN = 64: all warps work on the same tile; g = its 32-channel group, the warp owns channels g*32 + h*16 .. + 16.
//   BN = 32: g = the work unit's owner (units alternate, each with its own accumulator stages {2g, 2g+1}), the warp owns
//            channels h*16 .. + 16 of the unit's single group.
// The two warps (h = 0, 1) of a (q, g) pair fill ONE 32-pixel x 32-channel staging tile and the h = 0 warp hands it to
// TMA; two named barriers per step (tile free -> both write -> tile complete) keep the pair in step.
__device__ __forceinline__ void pair_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }

template <int BN>
__device__ __forceinline__ void tc4_lif16_epilogue(const Tc4Params& p, const CUtensorMap& tmY, uint32_t tmem_base,
                                                   uint64_t* acc_full, uint64_t* acc_empty, uint8_t* epi_base,
                                                   int my_chunk, int warp, int lane) {
  constexpr uint32_t NACC = t4_nacc(BN);
  constexpr bool UNIT_SPLIT = BN == 32;
  static_assert(BN == 32 || BN == 64, "fused LIF: 32 or 64 output channels per CTA");
  const int q = warp & 3, h = (warp >> 2) & 1, g = warp >> 3;
  const int pair = q + 4 * g;
  const uint32_t stage_u32 = smem_u32(epi_base) + (uint32_t)pair * T4_EPI_BYTES;
  const int row = q * 32 + lane;
  const int64_t hw = (int64_t)p.Ho * p.Wo;
  const float descale = __ldg(p.descale);
  const f32x2_t descale2 = pk2(descale, descale);
  const NP2 np2 = np_pairs(p.np);
  const bool half_out = p.out_fmt == SODA_ACT_HALF;
  const uint32_t st_row = stage_u32 + (uint32_t)lane * (half_out ? 64u : 128u);
  const uint32_t st_xor = half_out ? (((uint32_t)lane >> 1) & 3u) : ((uint32_t)lane & 7u);
  const int col_base = UNIT_SPLIT ? h * 16 : g * 32 + h * 16;
  const int co_grp = my_chunk * BN + (UNIT_SPLIT ? 0 : g * 32);   // first channel of the pair's 32-channel group
  const int co_base = my_chunk * BN + col_base;                    // first of this warp's 16 channels
  const bool group_live = co_grp < p.Cout;                          // Cout % 32 == 0: a group is whole or absent
  f32x2_t sh2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) sh2[j] = pk2(0.f, 0.f);
  if (p.shift != nullptr && group_live) {
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(p.shift + co_base + j));
      sh2[j / 2] = pk2(v.x, v.y), sh2[j / 2 + 1] = pk2(v.z, v.w);
    }
  }
  unsigned int n_over = 0;
  uint32_t local = 0, unit = 0, mine_cnt = 0;
  Tc4Iter it;
  it.init(p);
  for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, it.next(p)) {
    for (int m = 0; m < p.mt; ++m) {
      const bool mine = !UNIT_SPLIT || ((unit & 1u) == (uint32_t)g);
      ++unit;
      if (!mine) {
        local += (uint32_t)p.T_inner;
        continue;
      }
      const int spt = it.sp * p.mt + m;
      const int ty = spt / p.tiles_x, tx = spt - ty * p.tiles_x;
      const int oy = ty * T4_BH + (row >> 3), ox = tx * T4_BW + (row & 7);
      const bool valid = oy < p.Ho && ox < p.Wo;
      const int64_t rr = (int64_t)oy * p.Wo + ox;
      const int64_t frame0 = it.outer;
      const int sx = tx * T4_BW, sy = ty * T4_BH + q * 4;
      const int64_t soff = ((int64_t)it.outer * hw + rr) * p.Cout + co_base;
      f32x2_t sv[8], si[8];
#pragma unroll
      for (int c = 0; c < 16; c += 4) {
        float4 a = make_float4(p.np.v_leak, p.np.v_leak, p.np.v_leak, p.np.v_leak);
        float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
        if (valid && p.v0 != nullptr && group_live) {
          a = __ldg(reinterpret_cast<const float4*>(p.v0 + soff + c));
          b = __ldg(reinterpret_cast<const float4*>(p.i0 + soff + c));
        }
        sv[c / 2] = pk2(a.x, a.y), sv[c / 2 + 1] = pk2(a.z, a.w);
        si[c / 2] = pk2(b.x, b.y), si[c / 2 + 1] = pk2(b.z, b.w);
      }
      for (int t = 0; t < p.T_inner; ++t, ++local, ++mine_cnt) {
        const uint32_t acc = UNIT_SPLIT ? 2u * (uint32_t)g + (mine_cnt & 1u) : local % NACC;
        const uint32_t acc_ph = UNIT_SPLIT ? (mine_cnt >> 1) & 1u : (local / NACC) & 1u;
        const int64_t frame = frame0 + (int64_t)t * p.Bsz;
        const bool has_res = p.residual != nullptr && valid && group_live;
        // this warp's 16 residual channels, fetched ahead of the accumulator wait: F32 64 B, HALF 32 B, PAIR 32 B of the
        // group's hi half + 32 B of its lo half (64 B further)
        uint4 rp[4];
        if (has_res) {
          const uint8_t* rg = p.residual + frame * p.res_bs + rr * p.res_ps;
          if (p.res_fmt == SODA_ACT_F32) {
            const uint8_t* r0 = rg + (int64_t)co_base * 4;
#pragma unroll
            for (int j = 0; j < 4; ++j) rp[j] = ldg_u4(r0 + j * 16);
          } else if (p.res_fmt == SODA_ACT_PAIR) {
            const uint8_t* r0 = rg + (int64_t)co_grp * 4 + h * 32;
            rp[0] = ldg_u4(r0), rp[1] = ldg_u4(r0 + 16), rp[2] = ldg_u4(r0 + 64), rp[3] = ldg_u4(r0 + 80);
          } else {
            const uint8_t* r0 = rg + (int64_t)co_base * 2;
            rp[0] = ldg_u4(r0), rp[1] = ldg_u4(r0 + 16);
          }
        }
        mbar_wait(&acc_full[acc], acc_ph);
        tc_fence_after();
        const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + acc * (2 * BN) + (uint32_t)col_base;
        float o[16];
        {
          uint32_t r[16], rc[16];
          tmem_ld_32x16(t_row, r);
          tmem_ld_32x16(t_row + BN, rc);
          tmem_ld_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&acc_empty[acc]);
          if (p.shift != nullptr) {
#pragma unroll
            for (int j = 0; j < 16; j += 2) {
              const f32x2_t sum = add2(pk2(__uint_as_float(r[j]), __uint_as_float(r[j + 1])),
                                       pk2(__uint_as_float(rc[j]), __uint_as_float(rc[j + 1])));
              upk2(fma2(sum, descale2, sh2[j / 2]), o[j], o[j + 1]);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) o[j] = __fmul_rn(__uint_as_float(r[j]) + __uint_as_float(rc[j]), descale);
          }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) lif_step2(np2, pk2(o[2 * j], o[2 * j + 1]), sv[j], si[j], o[2 * j], o[2 * j + 1]);
        if (has_res) {
          if (p.res_fmt == SODA_ACT_F32) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              o[4 * j] += __uint_as_float(rp[j].x), o[4 * j + 1] += __uint_as_float(rp[j].y);
              o[4 * j + 2] += __uint_as_float(rp[j].z), o[4 * j + 3] += __uint_as_float(rp[j].w);
            }
          } else if (p.res_fmt == SODA_ACT_PAIR) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const uint32_t hw4[4] = {rp[j].x, rp[j].y, rp[j].z, rp[j].w};
              const uint32_t lw4[4] = {rp[j + 2].x, rp[j + 2].y, rp[j + 2].z, rp[j + 2].w};
#pragma unroll
              for (int w = 0; w < 4; ++w) {
                const float2 a = unpack_h2(hw4[w]), b = unpack_h2(lw4[w]);
                o[8 * j + 2 * w] += a.x + b.x;
                o[8 * j + 2 * w + 1] += a.y + b.y;
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const uint32_t hw4[4] = {rp[j].x, rp[j].y, rp[j].z, rp[j].w};
#pragma unroll
              for (int w = 0; w < 4; ++w) {
                const float2 a = unpack_h2(hw4[w]);
                o[8 * j + 2 * w] += a.x, o[8 * j + 2 * w + 1] += a.y;
              }
            }
          }
        }
        // ---- the pair's staging tile: free (the leader's previous store has read it) -> both write -> complete ----
        if (h == 0 && lane == 0) bulk_wait_read0();
        pair_sync(1 + pair);
        if (p.out_fmt == SODA_ACT_F32) {
#pragma unroll
          for (uint32_t j = 0; j < 4; ++j)
            sts128(st_row + (((uint32_t)h * 4u + j) ^ st_xor) * 16u, o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
        } else {
          uint32_t whi[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) whi[j] = pack_h2(o[2 * j], o[2 * j + 1]);
          if (p.residual != nullptr) {   // spikes {0, 1} alone cannot leave the fp16 range; spike + residual is checked
            float amax = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fmaxf(fabsf(o[2 * j]), fabsf(o[2 * j + 1])));
            n_over += amax > 65504.f;
          }
#pragma unroll
          for (uint32_t j = 0; j < 2; ++j)
            sts128(st_row + (((uint32_t)h * 2u + j) ^ st_xor) * 16u, __uint_as_float(whi[4 * j]), __uint_as_float(whi[4 * j + 1]),
                   __uint_as_float(whi[4 * j + 2]), __uint_as_float(whi[4 * j + 3]));
          if (p.out_fmt == SODA_ACT_PAIR) {
            uint32_t wlo[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float2 hf = unpack_h2(whi[j]);
              wlo[j] = pack_h2(o[2 * j] - hf.x, o[2 * j + 1] - hf.y);
            }
#pragma unroll
            for (uint32_t j = 0; j < 2; ++j)
              sts128(st_row + ((4u + (uint32_t)h * 2u + j) ^ st_xor) * 16u, __uint_as_float(wlo[4 * j]),
                     __uint_as_float(wlo[4 * j + 1]), __uint_as_float(wlo[4 * j + 2]), __uint_as_float(wlo[4 * j + 3]));
          }
        }
        fence_proxy_async_smem();
        pair_sync(1 + pair);
        if (h == 0 && lane == 0 && group_live) {
          tma_store_4d(&tmY, stage_u32, co_grp, sx, sy, (int)frame);
          bulk_commit();
        }
      }
      if (valid && group_live) {
#pragma unroll
        for (int c = 0; c < 16; c += 4) {
          float4 a, b;
          upk2(sv[c / 2], a.x, a.y), upk2(sv[c / 2 + 1], a.z, a.w);
          upk2(si[c / 2], b.x, b.y), upk2(si[c / 2 + 1], b.z, b.w);
          *reinterpret_cast<float4*>(p.v_out + soff + c) = a;
          *reinterpret_cast<float4*>(p.i_out + soff + c) = b;
        }
      }
    }
  }
  if (h == 0 && lane == 0) bulk_wait_all();
  if (p.status != nullptr && n_over != 0) atomicAdd(p.status, n_over);
}

// NE = epilogue warps.  8: two per TMEM lane quarter, 32 output channels each.  16 (fused LIF only, SODA_TC4_LIF16=1):
// four per quarter, 16 channels each - the LIF epilogue is instruction-bound (~700 warp-instructions per warp and time
// step), and two warps per scheduler cannot hide its fixed latencies; four can.  Role warps follow the epilogue warps.
template <int BN, bool LIF, int NE = 8>
__global__ void __launch_bounds__((NE + 4) * 32, 1)
conv_tc4_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmY, Tc4Params p) {
  static_assert(NE == 8 || (NE == 16 && LIF), "16 epilogue warps: fused LIF variants only");
  constexpr int T4_W_PROD_A = NE, T4_W_PROD_B = NE + 1, T4_W_MMA = NE + 2;
  constexpr uint32_t NACC = t4_nacc(BN);
  constexpr uint32_t TMEM_COLS = NACC * 2 * BN;
  constexpr bool UNIT_SPLIT = BN == 32;          // the two warps of a lane quarter alternate work units
  constexpr int NG = BN >= 64 ? BN / 64 : 1;     // 32-channel groups per epilogue warp
  static_assert(!LIF || NG == 1, "fused LIF keeps 2 x 32 state registers per group");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_b = smem;
  uint8_t* smem_a = smem + p.b_region_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_a + (size_t)p.a_stages * p.plane_bytes);
  uint64_t* a_full = bars;             // [8]
  uint64_t* a_empty = bars + 8;        // [8]
  uint64_t* b_full = bars + 16;        // [8]
  uint64_t* b_empty = bars + 24;       // [8]
  uint64_t* acc_full = bars + 32;      // [4]
  uint64_t* acc_empty = bars + 36;     // [4]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 40);
  uint8_t* epi_base = reinterpret_cast<uint8_t*>(bars) + 1024;   // 1024-aligned: the store's swizzle is address-based

  // the shuffle tells ptxas that the warp index is warp-uniform: the role branches below are then uniform branches
  // and each role's scalar state (ring positions, descriptor words, barrier addresses) can live in uniform registers
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;

  if (warp == T4_W_PROD_A && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    prefetch_tmap(&tmY);
    for (int s = 0; s < T4_MAX_RING; ++s) {
      mbar_init(&a_full[s], 1);
      mbar_init(&a_empty[s], 1);
      mbar_init(&b_full[s], 1);
      mbar_init(&b_empty[s], 1);
    }
    for (int a = 0; a < 4; ++a) {
      mbar_init(&acc_full[a], 1);
      mbar_init(&acc_empty[a], (UNIT_SPLIT ? 4 : 8) * (NE / 8));   // one arrival per epilogue warp that reads the stage
    }
    fence_barrier_init();
  }
  if (warp == T4_W_MMA) tmem_alloc(tmem_slot, TMEM_COLS);
  pdl_trigger();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();
  const int my_chunk = blockIdx.x % p.nch;   // gridDim.x is a multiple of nch: fixed per CTA
  // (setmaxnreg re-balancing between the warpgroups was tried: this ptxas keeps the kernel-wide cap of 168 for the
  // epilogue and only adds spills to the de-allocated roles - measured with -Xptxas -v, not kept)

  if (warp == T4_W_PROD_A) {
    // ================================ A producer (TMA) =========================================
    uint32_t s = 0, ph = 0;
    const uint32_t smem_a_u32 = smem_u32(smem_a);
    const int rowch = p.in_fmt == SODA_ACT_PAIR ? 32 : 64;   // tensor-map elements per 128-byte row
    Tc4Iter it;
    it.init(p);
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, it.next(p)) {
      int x0[2] = {0, 0}, y0[2] = {0, 0};
      _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < p.mt) {
        // a tile index past the last one (odd tile count) loads zeros (out of bounds) and is never stored
        const int spt = it.sp * p.mt + m;
        if (p.flat) {
          x0[m] = spt * T4_BM;
        } else {
          const int ty = spt / p.tiles_x, tx = spt - ty * p.tiles_x;
          y0[m] = ty * T4_BH * p.stride - p.pad;
          x0[m] = tx * T4_BW * p.stride - p.pad;
        }
      }
      for (int t = 0; t < p.T_inner; ++t) {
        const int img = p.flat ? 0 : t * p.Bsz + it.outer;
        if (p.mode == 2) {
          // four parity planes per channel slab; x0 / y0 = 2 * (tile origin) - 1 is the ODD start, + 1 the even one
          for (int cb = 0; cb < p.cblocks; ++cb)
            for (int pl = 0; pl < 4; ++pl) {
              _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < p.mt) {
                mbar_wait_uniform(&a_empty[s], ph ^ 1);
                if (elect_one()) {
                  mbar_expect_tx(&a_full[s], p.a_tx_bytes);
                  tma_load_4d_s(smem_a_u32 + s * p.plane_bytes, &tmA, &a_full[s], cb * rowch, x0[m] + ((pl & 1) ? 0 : 1),
                                y0[m] + ((pl & 2) ? 0 : 1), img);
                }
                __syncwarp();
                if (++s == (uint32_t)p.a_stages) s = 0, ph ^= 1;
              }
            }
          continue;
        }
        const int n_tap = p.mode == 0 ? 1 : p.taps;
        int ky = 0, kx = 0;
        for (int tap = 0; tap < n_tap; ++tap) {
          for (int cb = 0; cb < p.cblocks; cb += p.kg) {
            _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < p.mt) {
              mbar_wait_uniform(&a_empty[s], ph ^ 1);
              if (elect_one()) {
                if (p.dbg & 8) {
                  mbar_arrive(&a_full[s]);
                } else {
                  mbar_expect_tx(&a_full[s], p.a_tx_bytes);
                  tma_load_4d_s(smem_a_u32 + s * p.plane_bytes, &tmA, &a_full[s], cb * rowch, x0[m] + kx, y0[m] + ky,
                                img);
                  if (p.kg == 2)
                    tma_load_4d_s(smem_a_u32 + s * p.plane_bytes + p.a_tile_bytes, &tmA, &a_full[s], (cb + 1) * rowch,
                                  x0[m] + kx, y0[m] + ky, img);
                }
              }
              __syncwarp();
              if (++s == (uint32_t)p.a_stages) s = 0, ph ^= 1;
            }
          }
          if (++kx == p.k) kx = 0, ++ky;
        }
      }
    }
  } else if (warp == T4_W_PROD_B) {
    // ================================ B producer (TMA) =========================================
    // packed weights: [tap][Cout_pad][2 * Cin] fp16; PAIR: per 32 channels hi[32] | lo[32] (one 128-byte row);
    // HALF: per 64 channels hi[64] | lo[64] (two rows -> two BN x 128 B tiles)
    const uint32_t smem_b_u32 = smem_u32(smem_b);
    const bool pair = p.in_fmt == SODA_ACT_PAIR;
    if (p.b_stationary) {
      if (elect_one()) mbar_expect_tx(&b_full[0], (uint32_t)p.kblocks * p.b_kb_bytes);
      __syncwarp();
      for (int tap = 0; tap < p.taps; ++tap)
        for (int cb = 0; cb < p.cblocks; ++cb) {
          const uint32_t dst = smem_b_u32 + (uint32_t)(tap * p.cblocks + cb) * p.b_kb_bytes;
          if (elect_one()) {
            tma_load_4d_s(dst, &tmB, &b_full[0], cb * (pair ? 32 : 64), my_chunk * BN, 0, tap);
          }
          __syncwarp();
        }
    } else {
      uint32_t s = 0, ph = 0;
      const bool halo3 = p.mode == 0 && p.k == 3;          // halo mode walks (cb, tap); everything else (tap, cb)
      const int n_o = halo3 ? p.cblocks : p.taps;
      const int n_i = halo3 ? p.taps : p.cblocks;
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
        for (int t = 0; t < p.T_inner; ++t) {
          if (p.mode == 2) {
            // the issuer's order: per channel slab the taps of plane 0, 1, 2, 3
            for (int cb = 0; cb < p.cblocks; ++cb)
              for (int j = 0; j < 9; ++j) {
                const int tap = j == 0 ? 4 : j == 1 ? 3 : j == 2 ? 5 : j == 3 ? 1 : j == 4 ? 7 : j == 5 ? 0 : j == 6 ? 2 : j == 7 ? 6 : 8;
                mbar_wait_uniform(&b_empty[s], ph ^ 1);
                if (elect_one()) {
                  mbar_expect_tx(&b_full[s], p.b_kb_bytes);
                  const uint32_t dst = smem_b_u32 + s * p.b_kb_bytes;
                  tma_load_4d_s(dst, &tmB, &b_full[s], cb * (pair ? 32 : 64), my_chunk * BN, 0, tap);
                }
                __syncwarp();
                if (++s == (uint32_t)p.b_stages) s = 0, ph ^= 1;
              }
            continue;
          }
          for (int o = 0; o < n_o; ++o)
            for (int i = 0; i < n_i; i += p.kg) {   // (kg == 2 only outside the halo mode: i walks the channel slabs)
              const int tap = halo3 ? i : o, cb = halo3 ? o : i;
              mbar_wait_uniform(&b_empty[s], ph ^ 1);
              if (elect_one() && (p.dbg & 16)) mbar_arrive(&b_full[s]);
              if (elect_one() && !(p.dbg & 16)) {
                mbar_expect_tx(&b_full[s], p.b_kb_bytes * (uint32_t)p.kg);
                for (int j = 0; j < p.kg; ++j) {
                  const uint32_t dst = smem_b_u32 + (s * (uint32_t)p.kg + (uint32_t)j) * p.b_kb_bytes;
                  tma_load_4d_s(dst, &tmB, &b_full[s], (cb + j) * (pair ? 32 : 64), my_chunk * BN, 0, tap);
                }
              }
              __syncwarp();
              if (++s == (uint32_t)p.b_stages) s = 0, ph ^= 1;
            }
        }
      }
    }
  } else if (warp == T4_W_MMA) {
    const uint32_t sa_u32 = smem_u32(smem_a), sb_u32 = smem_u32(smem_b);
    const bool pair = p.in_fmt == SODA_ACT_PAIR;
    const int rmode = p.mode == 2 ? 2 : ((p.mode == 0 && p.k == 3) ? 0 : 1);
#define TC4_ROLE(MT_, PAIR_, MODE_) \
    tc4_mma_role<BN, MT_, PAIR_, MODE_>(p, tmem_base, sa_u32, sb_u32, a_full, a_empty, b_full, b_empty, acc_full, acc_empty)
#define TC4_ROLE_M(MT_, PAIR_) \
    do { if (rmode == 0) TC4_ROLE(MT_, PAIR_, 0); else if (rmode == 1) TC4_ROLE(MT_, PAIR_, 1); else TC4_ROLE(MT_, PAIR_, 2); } while (0)
    if (p.mt == 2) {
      if (pair) TC4_ROLE_M(2, true); else TC4_ROLE_M(2, false);
    } else {
      if (pair) TC4_ROLE_M(1, true); else TC4_ROLE_M(1, false);
    }
#undef TC4_ROLE_M
#undef TC4_ROLE
  } else if (warp < NE) {
    if constexpr (NE == 16) {
      tc4_lif16_epilogue<BN>(p, tmY, tmem_base, acc_full, acc_empty, epi_base, my_chunk, warp, lane);
    } else {
    // ================================ epilogue ================================================
    // Two warps per TMEM lane quarter.  BN >= 64: warpgroup e owns columns [e*BN/2, (e+1)*BN/2) of every tile;
    // BN = 32: the warpgroups alternate work units.  Per 32-channel group a warp reads main + correction accumulator,
    // finishes the values (descale, shift, LIF step, residual), encodes its 32 rows x 128 (64) bytes into a swizzled
    // 4 KB staging tile and hands it to TMA: the store engine clips the tile at the frame border and un-swizzles, no
    // per-row addressing, predicates or global store instructions in the warp.
    const int e = warp >> 2;
    const int q = warp & 3;
    const uint32_t stage_u32 = smem_u32(epi_base) + (uint32_t)warp * T4_EPI_BYTES;
    const int row = q * 32 + lane;
    const int64_t hw = (int64_t)p.Ho * p.Wo;
    const float descale = __ldg(p.descale);
    const f32x2_t descale2 = pk2(descale, descale);
    [[maybe_unused]] const NP2 np2 = np_pairs(p.np);
    const bool half_out = p.out_fmt == SODA_ACT_HALF;
    // byte address of 16-byte chunk j of this lane's row inside the staging tile (TMA SWIZZLE_128B / SWIZZLE_64B)
    const uint32_t st_row = stage_u32 + (uint32_t)lane * (half_out ? 64u : 128u);
    const uint32_t st_xor = half_out ? (((uint32_t)lane >> 1) & 3u) : ((uint32_t)lane & 7u);
    unsigned int n_over = 0;
    uint32_t local = 0, unit = 0, mine_cnt = 0;
    const bool edbg = T4_EPI_DBG && (p.dbg & 32) != 0;   // (make CFLAGS=-DSODA_TC4_EPI_DBG: per-warp clock64 counters)
    const long long e_t0 = clock64();
    long long e_wacc = 0, e_wst = 0, e_ld = 0;
    uint32_t e_steps = 0;
    Tc4Iter it;
    it.init(p);
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, it.next(p)) {
      for (int m = 0; m < p.mt; ++m) {   // (mt == 2 with LIF only at BN = 32, where tile m belongs to warpgroup m)
        const bool mine = !UNIT_SPLIT || ((unit & 1u) == (uint32_t)e);
        ++unit;
        if (!mine) {
          local += (uint32_t)p.T_inner;
          continue;
        }
        const int spt = it.sp * p.mt + m;
        bool valid;
        int64_t rr;        // pixel index inside its frame
        int64_t frame0;    // frame of step t = 0
        int sx, sy;        // store coordinates of this warp's 32 rows: (x0, y0) of its 8 x 4 patch, or (first pixel, 0)
        if (p.flat) {
          const int64_t pix = (int64_t)spt * T4_BM + row;
          valid = pix < p.total_pix;
          frame0 = 0;
          rr = pix;
          sx = spt * T4_BM + q * 32, sy = 0;
        } else {
          const int ty = spt / p.tiles_x, tx = spt - ty * p.tiles_x;
          const int oy = ty * T4_BH + (row >> 3);
          const int ox = tx * T4_BW + (row & 7);
          valid = oy < p.Ho && ox < p.Wo;
          rr = (int64_t)oy * p.Wo + ox;
          frame0 = it.outer;
          sx = tx * T4_BW, sy = ty * T4_BH + q * 4;
        }
        const int col_base = UNIT_SPLIT ? 0 : e * (BN / 2);
        const int co_base = my_chunk * BN + col_base;
        f32x2_t sv[LIF ? 16 : 1], si[LIF ? 16 : 1];   // membrane / current of this lane's 32 neurons, two per register pair
        if constexpr (LIF) {
          const int64_t soff = ((int64_t)it.outer * hw + rr) * p.Cout + co_base;
#pragma unroll
          for (int c = 0; c < 32; c += 4) {
            float4 a = make_float4(p.np.v_leak, p.np.v_leak, p.np.v_leak, p.np.v_leak);
            float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
            if (valid && p.v0 != nullptr && co_base < p.Cout) {
              a = __ldg(reinterpret_cast<const float4*>(p.v0 + soff + c));
              b = __ldg(reinterpret_cast<const float4*>(p.i0 + soff + c));
            }
            sv[c / 2] = pk2(a.x, a.y), sv[c / 2 + 1] = pk2(a.z, a.w);
            si[c / 2] = pk2(b.x, b.y), si[c / 2 + 1] = pk2(b.z, b.w);
          }
        }
        for (int t = 0; t < p.T_inner; ++t, ++local, ++mine_cnt) {
          // (see tc4_mma_role: with BN = 32 this warpgroup's units use its own two stages)
          const uint32_t acc = UNIT_SPLIT ? 2u * (uint32_t)e + (mine_cnt & 1u) : local % NACC;
          const uint32_t acc_ph = UNIT_SPLIT ? (mine_cnt >> 1) & 1u : (local / NACC) & 1u;
          const int64_t frame = frame0 + (int64_t)t * p.Bsz;
          const bool has_res = p.residual != nullptr && valid;
          const uint8_t* res_row = p.residual + frame * p.res_bs + rr * p.res_ps;
          // the residual row of the first group is fetched BEFORE the wait for the accumulator: its latency then
          // overlaps the MMAs of this tile instead of sitting between the TMEM read and the store
          uint4 rp[8];
          if (has_res && co_base < p.Cout) {
            const uint8_t* rp0 = res_row + (int64_t)co_base * (p.res_fmt == SODA_ACT_HALF ? 2 : 4);
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (j < 4 || p.res_fmt != SODA_ACT_HALF) rp[j] = ldg_u4(rp0 + j * 16);
          }
          { const long long w0 = edbg ? clock64() : 0; mbar_wait(&acc_full[acc], acc_ph); if (edbg) e_wacc += clock64() - w0, ++e_steps; }
          tc_fence_after();
          const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + acc * (2 * BN) + (uint32_t)col_base;
#pragma unroll
          for (int g = 0; g < NG; ++g) {
            const int co0 = co_base + g * 32;
            const bool group_live = co0 < p.Cout;   // Cout % 32 == 0: a group is whole or absent
            float o[32];
#pragma unroll
            for (int h = 0; h < 2; ++h) {   // 16 columns of main + correction at a time: 32 transient registers
              uint32_t r[16], rc[16];
              const long long l0 = edbg ? clock64() : 0;
              tmem_ld_32x16(t_row + g * 32 + h * 16, r);
              tmem_ld_32x16(t_row + BN + g * 32 + h * 16, rc);
              tmem_ld_wait();
              if (edbg) e_ld += clock64() - l0;
              if (g == NG - 1 && h == 1) {
                // last TMEM read of this tile: hand the accumulator back before the rest
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&acc_empty[acc]);
              }
              if (p.shift != nullptr && group_live) {
#pragma unroll
                for (int j = 0; j < 16; j += 4) {
                  // fma(main + correction, descale, shift), two columns per FADD2 / FFMA2
                  const float4 sh = __ldg(reinterpret_cast<const float4*>(p.shift + co0 + h * 16 + j));
                  const f32x2_t s0 = add2(pk2(__uint_as_float(r[j]), __uint_as_float(r[j + 1])),
                                          pk2(__uint_as_float(rc[j]), __uint_as_float(rc[j + 1])));
                  const f32x2_t s1 = add2(pk2(__uint_as_float(r[j + 2]), __uint_as_float(r[j + 3])),
                                          pk2(__uint_as_float(rc[j + 2]), __uint_as_float(rc[j + 3])));
                  upk2(fma2(s0, descale2, pk2(sh.x, sh.y)), o[h * 16 + j], o[h * 16 + j + 1]);
                  upk2(fma2(s1, descale2, pk2(sh.z, sh.w)), o[h * 16 + j + 2], o[h * 16 + j + 3]);
                }
              } else {
#pragma unroll
                for (int j = 0; j < 16; ++j) o[h * 16 + j] = (__uint_as_float(r[j]) + __uint_as_float(rc[j])) * descale;
              }
            }
            if (p.dbg & 4) continue;
            if constexpr (LIF) {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                lif_step2(np2, pk2(o[2 * j], o[2 * j + 1]), sv[j], si[j], o[2 * j], o[2 * j + 1]);
            }
            if (has_res && group_live) {
              if (g > 0) {
                const uint8_t* rpg = res_row + (int64_t)co0 * (p.res_fmt == SODA_ACT_HALF ? 2 : 4);
#pragma unroll
                for (int j = 0; j < 8; ++j)
                  if (j < 4 || p.res_fmt != SODA_ACT_HALF) rp[j] = ldg_u4(rpg + j * 16);
              }
              if (p.res_fmt == SODA_ACT_F32) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  o[4 * j] += __uint_as_float(rp[j].x), o[4 * j + 1] += __uint_as_float(rp[j].y);
                  o[4 * j + 2] += __uint_as_float(rp[j].z), o[4 * j + 3] += __uint_as_float(rp[j].w);
                }
              } else if (p.res_fmt == SODA_ACT_PAIR) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const uint32_t hw4[4] = {rp[j].x, rp[j].y, rp[j].z, rp[j].w};
                  const uint32_t lw4[4] = {rp[j + 4].x, rp[j + 4].y, rp[j + 4].z, rp[j + 4].w};
#pragma unroll
                  for (int w = 0; w < 4; ++w) {
                    const float2 a = unpack_h2(hw4[w]), b = unpack_h2(lw4[w]);
                    o[8 * j + 2 * w] += a.x + b.x;
                    o[8 * j + 2 * w + 1] += a.y + b.y;
                  }
                }
              } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const uint32_t hw4[4] = {rp[j].x, rp[j].y, rp[j].z, rp[j].w};
#pragma unroll
                  for (int w = 0; w < 4; ++w) {
                    const float2 a = unpack_h2(hw4[w]);
                    o[8 * j + 2 * w] += a.x, o[8 * j + 2 * w + 1] += a.y;
                  }
                }
              }
            }
            // ---- encode this row's 32 channels: 8 (F32 / PAIR) or 4 (HALF) chunks of 16 bytes ----
            uint32_t w[32];
            // (one uniform branch per format around whole loops: a format test inside the loop gets if-converted and
            // the PAIR arithmetic is then issued, predicated off, for HALF output as well)
            if (p.out_fmt == SODA_ACT_F32) {
#pragma unroll
              for (int j = 0; j < 32; ++j) w[j] = __float_as_uint(o[j]);
            } else if (p.out_fmt == SODA_ACT_PAIR) {
              float amax = 0.f;
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const uint32_t hi = pack_h2(o[2 * j], o[2 * j + 1]);
                const float2 hf = unpack_h2(hi);
                w[j] = hi;
                w[16 + j] = pack_h2(o[2 * j] - hf.x, o[2 * j + 1] - hf.y);
                amax = fmaxf(amax, fmaxf(fabsf(o[2 * j]), fabsf(o[2 * j + 1])));
              }
              n_over += amax > 65504.f;
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) w[j] = pack_h2(o[2 * j], o[2 * j + 1]);
              if (!(LIF && !has_res)) {   // spikes {0, 1} cannot leave the fp16 range; anything else is checked
                float amax = 0.f;
#pragma unroll
                for (int j = 0; j < 16; ++j) amax = fmaxf(amax, fmaxf(fabsf(o[2 * j]), fabsf(o[2 * j + 1])));
                n_over += amax > 65504.f;
              }
            }
            // ---- stage (swizzled) and let TMA store it ----
            { const long long w0 = edbg ? clock64() : 0;
              if (lane == 0) bulk_wait_read0();   // this warp's previous store has read the staging tile
              __syncwarp();
              if (edbg) e_wst += clock64() - w0; }
#pragma unroll
            for (uint32_t j = 0; j < 8; ++j)
              if (j < 4 || !half_out)
                sts128(st_row + ((j ^ st_xor) << 4), __uint_as_float(w[4 * j]), __uint_as_float(w[4 * j + 1]),
                       __uint_as_float(w[4 * j + 2]), __uint_as_float(w[4 * j + 3]));
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && group_live && !(p.dbg & 2)) {
              tma_store_4d(&tmY, stage_u32, co0, sx, sy, p.flat ? 0 : (int)frame);
              bulk_commit();
            }
          }
        }
        if constexpr (LIF) {
          if (valid && co_base < p.Cout) {
            const int64_t soff = ((int64_t)it.outer * hw + rr) * p.Cout + co_base;
#pragma unroll
            for (int c = 0; c < 32; c += 4) {
              float4 a, b;
              upk2(sv[c / 2], a.x, a.y), upk2(sv[c / 2 + 1], a.z, a.w);
              upk2(si[c / 2], b.x, b.y), upk2(si[c / 2 + 1], b.z, b.w);
              *reinterpret_cast<float4*>(p.v_out + soff + c) = a;
              *reinterpret_cast<float4*>(p.i_out + soff + c) = b;
            }
          }
        }
      }  // m
    }
    if (edbg && blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == 4))
      printf("tc4 epilogue warp %d: %lld clk total, %u tile-steps, waits: acc_full %lld, tmem ld %lld, staging %lld\n", warp,
             clock64() - e_t0, e_steps, e_wacc, e_ld, e_wst);
    if (lane == 0) bulk_wait_all();
    if (p.status != nullptr && n_over != 0) atomicAdd(p.status, n_over);
    }   // NE == 8
  }

  tc_fence_before();
  __syncthreads();
  if (warp == T4_W_MMA) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------
// host side: weight packing, tensor maps, configuration, dispatch
// ------------------------------------------------------------------------------------------------
static inline int tc4_bn_for(int Cout, bool lif) {
  static const int env_bn = []() { const char* e = getenv("SODA_TC4_BN"); return e ? atoi(e) : 0; }();
  if (Cout <= 32) return 32;
  if (env_bn == 32 || env_bn == 64 || (env_bn == 128 && !lif)) return Cout <= 64 ? std::min(env_bn, 64) : env_bn;
  if (Cout <= 64 || lif) return 64;
  return 128;
}
static inline int tc4_cout_pad(int Cout) {
  if (Cout <= 32) return 32;
  if (Cout <= 64) return 64;
  return (Cout + 127) / 128 * 128;
}

int64_t conv_tc4_packed_bytes(int Cout, int Cin, int k) {
  if (Cin % 32 != 0) return 0;
  return 4LL * k * k * tc4_cout_pad(Cout) * Cin;
}

// [Cout, Cin, k, k] * scale[co] * mult -> fp16 planes [tap][hi | lo][Cout_pad][Cin]: one layout for both input formats
// (a TMA box {32 or 64 channels, BN rows, 2 planes} is one weight k-block: the w_hi rows, then the w_lo rows)
__global__ void pack_weights_tc4_kernel(const float* __restrict__ w, const float* __restrict__ scale,
                                        const float* __restrict__ mult, int Cout, int Cin, int taps, int Cout_pad,
                                        __half* __restrict__ wp) {
  const int64_t n = (int64_t)taps * Cout_pad * Cin;
  const float mu = mult != nullptr ? __ldg(mult) : 1.0f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int ci = (int)(i % Cin);
    const int64_t r = i / Cin;   // tap * Cout_pad + co
    const int co = (int)(r % Cout_pad);
    const int tap = (int)(r / Cout_pad);
    float v = 0.f;
    if (co < Cout) {
      v = w[((int64_t)co * Cin + ci) * taps + tap];
      if (scale != nullptr) v = __fmul_rn(v, scale[co]);
      v = __fmul_rn(v, mu);
    }
    const __half hi = __float2half_rn(v);
    const __half lo = __float2half_rn(v - __half2float(hi));
    const int64_t base = ((int64_t)(tap * 2) * Cout_pad + co) * Cin + ci;
    wp[base] = hi;
    wp[base + (int64_t)Cout_pad * Cin] = lo;
  }
}

int conv_tc4_pack(const float* w, const float* scale, const float* mult, int Cout, int Cin, int k, int in_fmt,
                  void* packed, cudaStream_t st) {
  if (in_fmt != SODA_ACT_PAIR && in_fmt != SODA_ACT_HALF)
    return fail(SODA_ERR_BAD_ARG, "pconv_pack: input format must be SODA_ACT_PAIR or SODA_ACT_HALF");
  const int grp = in_fmt == SODA_ACT_PAIR ? 32 : 64;
  if (Cin % grp != 0) return fail(SODA_ERR_UNSUPPORTED, "pconv_pack: Cin must be a multiple of 32 (PAIR) / 64 (HALF)");
  const int64_t n = (int64_t)k * k * tc4_cout_pad(Cout) * Cin;
  pack_weights_tc4_kernel<<<(unsigned)std::min<int64_t>(ceil_div(n, 256), 4096), 256, 0, st>>>(
      w, scale, mult, Cout, Cin, k * k, tc4_cout_pad(Cout), reinterpret_cast<__half*>(packed));
  return after_launch("pack_weights_tc4_kernel");
}

typedef CUresult (*EncodeTiledFn4)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn4 get_encode_fn4() {
  static EncodeTiledFn4 fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn4>(ptr);
  }
  return fn;
}

static int encode_map4(CUtensorMap* m, CUtensorMapDataType dt, const void* base, int rank, const cuuint64_t* dims,
                       const cuuint64_t* strides_b, const cuuint32_t* box, const cuuint32_t* estr,
                       CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn4 fn = get_encode_fn4();
  if (fn == nullptr) return fail(SODA_ERR_CUDA, "pconv: cuTensorMapEncodeTiled entry point not available");
  CUresult r = fn(m, dt, (cuuint32_t)rank, const_cast<void*>(base), dims, strides_b, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("pconv: cuTensorMapEncodeTiled failed (%d)", (int)r);
    return (int)SODA_ERR_CUDA;
  }
  return 0;
}

template <int BN, bool LIF, int NE = 8>
static int launch_tc4(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmY, const Tc4Params& p, int grid,
                      size_t smem, cudaStream_t st) {
  constexpr int T4_THREADS = (NE + 4) * 32;
  // the opt-in is per device: set it on every launch (cheap), not once per process
  cudaError_t e = cudaFuncSetAttribute(conv_tc4_kernel<BN, LIF, NE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
  if (e != cudaSuccess) {
    set_error("pconv: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return (int)SODA_ERR_CUDA;
  }
  static const bool pdl = []() { const char* e = getenv("SODA_PDL"); return !(e && atoi(e) == 0); }();
  if (pdl) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)grid), cfg.blockDim = dim3(T4_THREADS);
    cfg.dynamicSmemBytes = smem, cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at, cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, conv_tc4_kernel<BN, LIF, NE>, tmA, tmB, tmY, p);
  } else {
    conv_tc4_kernel<BN, LIF, NE><<<grid, T4_THREADS, smem, st>>>(tmA, tmB, tmY, p);
  }
  return after_launch("conv_tc4_kernel");
}

static inline int act_bytes(int fmt) { return fmt == SODA_ACT_HALF ? 2 : 4; }

bool conv_tc4_supported(const soda_pconv_desc_t* d, const void* x, const void* y, const void* residual) {
  if (d->in_fmt != SODA_ACT_PAIR && d->in_fmt != SODA_ACT_HALF) return false;
  if (d->out_fmt < SODA_ACT_F32 || d->out_fmt > SODA_ACT_HALF) return false;
  const int grp = d->in_fmt == SODA_ACT_PAIR ? 32 : 64;
  if (d->Cin % grp != 0 || d->Cout % 32 != 0) return false;
  if (d->k != 1 && d->k != 3) return false;
  if (d->stride != 1 && d->stride != 2) return false;
  const int ib = act_bytes(d->in_fmt), ob = act_bytes(d->out_fmt);
  // every 32-channel group of a pixel is one aligned 128-byte (64-byte for HALF output) piece
  if ((reinterpret_cast<uintptr_t>(x) & 127) || (d->in_pix_stride * ib) % 128 || (d->in_batch_stride * ib) % 128) return false;
  const int oal = d->out_fmt == SODA_ACT_HALF ? 64 : 128;
  if ((reinterpret_cast<uintptr_t>(y) & (oal - 1)) || (d->out_pix_stride * ob) % oal || (d->out_batch_stride * ob) % oal)
    return false;
  if (residual != nullptr) {
    if (d->res_fmt < SODA_ACT_F32 || d->res_fmt > SODA_ACT_HALF) return false;
    const int rb = act_bytes(d->res_fmt);
    const int ral = d->res_fmt == SODA_ACT_HALF ? 64 : 128;
    if ((reinterpret_cast<uintptr_t>(residual) & (ral - 1)) || (d->res_pix_stride * rb) % ral ||
        (d->res_batch_stride * rb) % ral)
      return false;
  }
  return true;
}

struct Tc4Lif {
  int T;
  NP np;
  const float* v0;
  const float* i0;
  float* v_out;
  float* i_out;
};

int conv_tc4_dispatch(const soda_pconv_desc_t* d, const void* x, const void* packed, const float* descale,
                      const float* shift, const void* residual, void* y, unsigned int* status, const Tc4Lif* lif,
                      void* stream) {
  if (!conv_tc4_supported(d, x, y, residual))
    return fail(SODA_ERR_UNSUPPORTED,
                "pconv: needs PAIR / HALF input with Cin % 32 (64) == 0, Cout % 32 == 0, 128-byte aligned channel groups, "
                "k in {1,3}, stride in {1,2}");
  Tc4Params p;
  memset(&p, 0, sizeof(p));
  {
    static const int env_dbg = []() { const char* e = getenv("SODA_TC4_DBG"); return e ? atoi(e) : 0; }();
    p.dbg = env_dbg;
  }
  p.H = d->H, p.W = d->W, p.Cin = d->Cin, p.Cout = d->Cout, p.k = d->k, p.stride = d->stride;
  p.pad = d->k / 2;
  p.Ho = (d->H + 2 * p.pad - d->k) / d->stride + 1;
  p.Wo = (d->W + 2 * p.pad - d->k) / d->stride + 1;
  if (p.Ho <= 0 || p.Wo <= 0) return fail(SODA_ERR_BAD_ARG, "pconv: empty output");
  p.in_fmt = d->in_fmt, p.out_fmt = d->out_fmt, p.res_fmt = d->res_fmt;
  p.a_exact = (d->flags & SODA_PCONV_FLAG_INPUT_EXACT) ? 1 : 0;
  const int rowch = d->in_fmt == SODA_ACT_PAIR ? 32 : 64;
  const int ib = act_bytes(d->in_fmt), ob = act_bytes(d->out_fmt), rb = act_bytes(d->res_fmt);
  p.taps = d->k * d->k;
  p.cblocks = d->Cin / rowch;
  p.kblocks = p.taps * p.cblocks;
  p.total_pix = (int64_t)d->N * p.Ho * p.Wo;
  p.out_bs = d->out_batch_stride * ob, p.out_ps = d->out_pix_stride * ob;
  p.res_bs = d->res_batch_stride * rb, p.res_ps = d->res_pix_stride * rb;
  p.shift = shift, p.descale = descale;
  p.residual = reinterpret_cast<const uint8_t*>(residual), p.y = reinterpret_cast<uint8_t*>(y);
  p.status = status;
  const bool is_lif = lif != nullptr;
  if (is_lif) {
    if (lif->T <= 0 || d->N % lif->T != 0) return fail(SODA_ERR_BAD_ARG, "pconv_lif: N must be T * batch");
    if (!lif->v_out || !lif->i_out) return fail(SODA_ERR_BAD_ARG, "pconv_lif: null state output");
    if ((lif->v0 == nullptr) != (lif->i0 == nullptr)) return fail(SODA_ERR_BAD_ARG, "pconv_lif: v0 / i0 must come together");
    p.lif = 1, p.np = lif->np, p.v0 = lif->v0, p.i0 = lif->i0, p.v_out = lif->v_out, p.i_out = lif->i_out;
    p.T_inner = lif->T, p.Bsz = d->N / lif->T, p.n_outer = p.Bsz;
  } else {
    p.T_inner = 1, p.Bsz = d->N, p.n_outer = d->N;
  }
  const int BN = tc4_bn_for(d->Cout, is_lif);
  const int Cout_pad = tc4_cout_pad(d->Cout);
  p.nch = (d->Cout + BN - 1) / BN;
  const uint32_t b_kb = (uint32_t)BN * 128u * (d->in_fmt == SODA_ACT_PAIR ? 1u : 2u);
  p.b_kb_bytes = b_kb;

  CUtensorMap tmA, tmB;
  const CUtensorMapDataType adt = d->in_fmt == SODA_ACT_PAIR ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  const int64_t HW = (int64_t)d->H * d->W;
  const bool frames_dense = d->N == 1 || d->in_batch_stride == HW * d->in_pix_stride;
  {
    static const int env_phase = []() { const char* e = getenv("SODA_TC4_PHASE"); return e ? atoi(e) : 1; }();
    // stride 1: halo mode; 3x3 stride 2: parity planes (each input pixel loaded ~1.1 x instead of 2.25 x); else one tap
    // tile per k-block
    p.mode = d->stride == 1 ? 0 : ((d->k == 3 && d->stride == 2 && env_phase) ? 2 : 1);
  }
  const int64_t HWo = (int64_t)p.Ho * p.Wo;
  const bool out_dense = (d->N == 1 || d->out_batch_stride == HWo * d->out_pix_stride) &&
                         (residual == nullptr || d->N == 1 || d->res_batch_stride == HWo * d->res_pix_stride);
  p.flat = (d->k == 1 && d->stride == 1 && frames_dense && out_dense && !is_lif) ? 1 : 0;
  p.rows = T4_BM, p.WW = T4_BW, p.sbo = 1024;
  if (p.flat) {
    p.n_spatial = (int)ceil_div(p.total_pix, T4_BM);
    p.tiles_x = p.n_spatial, p.tiles_y = 1;
    p.n_outer = 1;
    const cuuint64_t dims[4] = {(cuuint64_t)d->Cin, (cuuint64_t)p.total_pix, 1, 1};
    const cuuint64_t str[3] = {(cuuint64_t)d->in_pix_stride * ib, (cuuint64_t)p.total_pix * d->in_pix_stride * ib,
                               (cuuint64_t)p.total_pix * d->in_pix_stride * ib};
    const cuuint32_t box[4] = {(cuuint32_t)rowch, T4_BM, 1, 1};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    int rc = encode_map4(&tmA, adt, x, 4, dims, str, box, es);
    if (rc) return rc;
  } else {
    p.tiles_y = (int)ceil_div(p.Ho, T4_BH), p.tiles_x = (int)ceil_div(p.Wo, T4_BW);
    p.n_spatial = p.tiles_x * p.tiles_y;
    const cuuint64_t dims[4] = {(cuuint64_t)d->Cin, (cuuint64_t)d->W, (cuuint64_t)d->H, (cuuint64_t)d->N};
    const cuuint64_t str[3] = {(cuuint64_t)d->in_pix_stride * ib, (cuuint64_t)d->W * d->in_pix_stride * ib,
                               (cuuint64_t)(d->N == 1 ? HW * d->in_pix_stride : d->in_batch_stride) * ib};
    if (p.mode == 0) {
      p.WW = T4_BW + 2 * p.pad;
      const int HH = T4_BH + 2 * p.pad;
      p.rows = p.WW * HH;
      p.sbo = (uint32_t)p.WW * 128u;
      const cuuint32_t box[4] = {(cuuint32_t)rowch, (cuuint32_t)p.WW, (cuuint32_t)HH, 1};
      const cuuint32_t es[4] = {1, 1, 1, 1};
      int rc = encode_map4(&tmA, adt, x, 4, dims, str, box, es);
      if (rc) return rc;
    } else if (p.mode == 2) {
      p.WW = T4_BW + 1;
      p.rows = p.WW * (T4_BH + 1);          // 9 x 17 pixels of one parity plane
      p.sbo = (uint32_t)p.WW * 128u;
      const cuuint32_t box[4] = {(cuuint32_t)rowch, (cuuint32_t)(2 * p.WW), (cuuint32_t)(2 * (T4_BH + 1)), 1};
      const cuuint32_t es[4] = {1, 2, 2, 1};
      int rc = encode_map4(&tmA, adt, x, 4, dims, str, box, es);
      if (rc) return rc;
    } else {
      const cuuint32_t box[4] = {(cuuint32_t)rowch, (cuuint32_t)(T4_BW * d->stride), (cuuint32_t)(T4_BH * d->stride), 1};
      const cuuint32_t es[4] = {1, (cuuint32_t)d->stride, (cuuint32_t)d->stride, 1};
      int rc = encode_map4(&tmA, adt, x, 4, dims, str, box, es);
      if (rc) return rc;
    }
  }
  {
    static const int env_kg = []() { const char* e = getenv("SODA_TC4_KG"); return e ? atoi(e) : 1; }();   // measured: 2 halves the ring depth where shared memory is tight and loses 15-50 % on the streamed layers; only the fused stride-2 + LIF case gained
    const bool halo3_ = p.mode == 0 && p.k == 3;
    p.kg = (!halo3_ && p.mode != 2 && env_kg == 2 && p.cblocks % 2 == 0) ? 2 : 1;
  }
  p.a_tile_bytes = ((uint32_t)p.rows * 128u + 1023u) & ~1023u;
  p.a_tx_bytes = (uint32_t)p.rows * 128u * (uint32_t)p.kg;
  p.plane_bytes = p.a_tile_bytes * (uint32_t)p.kg;
  {
    // packed weights [tap][hi | lo][Cout_pad][Cin]: one k-block = box {32 | 64 channels, BN rows, both planes}
    const cuuint64_t dims[4] = {(cuuint64_t)d->Cin, (cuuint64_t)Cout_pad, 2, (cuuint64_t)p.taps};
    const cuuint64_t str[3] = {(cuuint64_t)d->Cin * 2, (cuuint64_t)Cout_pad * d->Cin * 2, (cuuint64_t)Cout_pad * d->Cin * 4};
    const cuuint32_t box[4] = {(cuuint32_t)rowch, (cuuint32_t)BN, 2, 1};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    int rc = encode_map4(&tmB, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, packed, 4, dims, str, box, es,
                         d->in_fmt == SODA_ACT_PAIR ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  CUtensorMap tmY;
  {
    // the epilogue's TMA store: 32 channels x (8 x 4 pixels | 32 pixels) per instruction, clipped at the frame border
    const CUtensorMapDataType ydt = d->out_fmt == SODA_ACT_HALF ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    const CUtensorMapSwizzle yswz = d->out_fmt == SODA_ACT_HALF ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B;
    const cuuint32_t es[4] = {1, 1, 1, 1};
    int rc;
    if (p.flat) {
      const cuuint64_t dims[4] = {(cuuint64_t)d->Cout, (cuuint64_t)p.total_pix, 1, 1};
      const cuuint64_t str[3] = {(cuuint64_t)d->out_pix_stride * ob, (cuuint64_t)p.total_pix * d->out_pix_stride * ob,
                                 (cuuint64_t)p.total_pix * d->out_pix_stride * ob};
      const cuuint32_t box[4] = {32, 32, 1, 1};
      rc = encode_map4(&tmY, ydt, y, 4, dims, str, box, es, yswz);
    } else {
      const cuuint64_t dims[4] = {(cuuint64_t)d->Cout, (cuuint64_t)p.Wo, (cuuint64_t)p.Ho, (cuuint64_t)d->N};
      const cuuint64_t str[3] = {(cuuint64_t)d->out_pix_stride * ob, (cuuint64_t)p.Wo * d->out_pix_stride * ob,
                                 (cuuint64_t)(d->N == 1 ? HWo * d->out_pix_stride : d->out_batch_stride) * ob};
      const cuuint32_t box[4] = {32, T4_BW, 4, 1};
      rc = encode_map4(&tmY, ydt, y, 4, dims, str, box, es, yswz);
    }
    if (rc) return rc;
  }
  // shared-memory plan: weights resident if one output-channel chunk fits beside an A ring of >= min_a planes,
  // else a weight ring as well
  const int64_t b_total = (int64_t)p.kblocks * b_kb;
  const bool halo3 = p.mode == 0 && p.k == 3;
  const int64_t plane = p.plane_bytes;
  static const int env_min_a = []() { const char* e = getenv("SODA_TC4_MIN_A"); return e ? atoi(e) : 0; }();
  const int min_a = env_min_a > 0 ? env_min_a : (halo3 || p.kg == 2 ? 2 : (p.mode == 2 ? 3 : 4));
  static const int env_mt = []() { const char* e = getenv("SODA_TC4_MT"); return e ? atoi(e) : 0; }();
  if (b_total + min_a * plane <= T4_SMEM_LIMIT) {
    p.b_stationary = 1, p.b_stages = 1, p.mt = 1;
    p.b_region_bytes = (uint32_t)b_total;
    p.a_stages = (int)std::min<int64_t>(T4_MAX_RING, (T4_SMEM_LIMIT - b_total) / plane);
    // LIF at BN = 32: a unit (tile x all T steps) belongs to ONE epilogue warpgroup, which walks its T steps serially
    // (the membrane lives in its registers).  Two tiles per item, issued step-interleaved into the two warpgroups' own
    // accumulator stages, keep both warpgroups busy; with one tile per item they took turns, one always idle.
    if (is_lif && BN == 32 && p.n_spatial > 1 && env_mt != 1 && env_mt != 3 && p.a_stages >= 4) p.mt = 2;
  } else {
    // streamed weights: an item takes TWO M-tiles that share each weight k-block (two accumulators of the four)
    p.b_stationary = 0;
    // Measured per layer (bench.py --per-layer, SODA_TC4_MT=2 against 4): sharing the k-blocks pays only on the
    // stride-2 3x3 layers (~7 % per tile) and only when pairing does not cost more in wave quantisation / an odd tile
    // count than that (256 -> 256 on 15 x 19 output: 3 tiles per frame, a quarter of the paired slots empty); 3x3
    // stride-1 and 1x1 layers run 4 .. 30 % faster unpaired.  (SODA_TC4_MT: 1 never, 2 LIF only, 3 streamed only by
    // this rule, 4 every streamed layer.)
    bool want_pair = false;
    if (!is_lif && p.n_spatial > 1 && env_mt != 1 && env_mt != 2) {
      const int64_t slots = std::max(1, kNumSMs / p.nch);
      const int64_t per_chunk_1 = (int64_t)p.n_spatial * p.n_outer, per_chunk_2 = (int64_t)((p.n_spatial + 1) / 2) * p.n_outer;
      const int64_t rounds_1 = (per_chunk_1 + slots - 1) / slots, rounds_2 = 2 * ((per_chunk_2 + slots - 1) / slots);
      want_pair = env_mt == 4 || (p.mode == 2 && rounds_2 * 93 < rounds_1 * 100);
    }
    p.mt = want_pair ? 2 : 1;
    const int64_t b_st = (int64_t)b_kb * p.kg;   // one B ring stage
    int bs = 4;
    while (bs > 2 && (int64_t)bs * b_st + 2 * plane > T4_SMEM_LIMIT) --bs;
    int64_t a = (T4_SMEM_LIMIT - (int64_t)bs * b_st) / plane;
    if (a < 2) return fail(SODA_ERR_UNSUPPORTED, "pconv: shared-memory plan does not fit");
    const int a_pref = (halo3 || p.mode == 2) ? 4 : (p.kg == 2 ? 4 : 6);
    p.a_stages = (int)std::min<int64_t>(std::min<int64_t>(a, T4_MAX_RING), a_pref);
    p.b_stages = (int)std::min<int64_t>(T4_MAX_RING, (T4_SMEM_LIMIT - (int64_t)p.a_stages * plane) / b_st);
    if (p.b_stages < 2) return fail(SODA_ERR_UNSUPPORTED, "pconv: shared-memory plan does not fit");
    p.b_region_bytes = (uint32_t)(p.b_stages * b_st);
  }
  p.n_sp_items = (p.n_spatial + p.mt - 1) / p.mt;
  p.n_items = p.nch * p.n_sp_items * p.n_outer;
  int dev = 0, n_sm = kNumSMs;
  if (cudaGetDevice(&dev) == cudaSuccess) {
    static int sm_cache[64] = {0};
    if (dev >= 0 && dev < 64) {
      if (sm_cache[dev] == 0) cudaDeviceGetAttribute(&sm_cache[dev], cudaDevAttrMultiProcessorCount, dev);
      if (sm_cache[dev] > 0) n_sm = sm_cache[dev];
    }
  }
  int grid = p.n_items < n_sm ? p.n_items : n_sm;
  grid = grid / p.nch * p.nch;
  if (grid < p.nch) grid = p.nch;
  p.it_step_sp = (grid / p.nch) % p.n_sp_items;
  p.it_step_outer = (grid / p.nch) / p.n_sp_items;
  const size_t smem = (size_t)p.b_region_bytes + (size_t)p.a_stages * p.plane_bytes + 1024 + 8 * T4_EPI_BYTES + 1024;
  cudaStream_t st = (cudaStream_t)stream;
  if (is_lif) {
    // 16 epilogue warps by default (measured on the step: fused-LIF kernels 1.02 -> 0.79 ms); SODA_TC4_LIF16=0: 8
    static const int lif16 = []() { const char* e = getenv("SODA_TC4_LIF16"); return e ? atoi(e) : 1; }();
    if (lif16) {
      if (BN == 32) return launch_tc4<32, true, 16>(tmA, tmB, tmY, p, grid, smem, st);
      return launch_tc4<64, true, 16>(tmA, tmB, tmY, p, grid, smem, st);
    }
    if (BN == 32) return launch_tc4<32, true>(tmA, tmB, tmY, p, grid, smem, st);
    return launch_tc4<64, true>(tmA, tmB, tmY, p, grid, smem, st);
  }
  if (BN == 32) return launch_tc4<32, false>(tmA, tmB, tmY, p, grid, smem, st);
  if (BN == 64) return launch_tc4<64, false>(tmA, tmB, tmY, p, grid, smem, st);
  return launch_tc4<128, false>(tmA, tmB, tmY, p, grid, smem, st);
}

}  // namespace soda

extern "C" int64_t soda_pconv_packed_bytes(int Cout, int Cin, int k) {
  if (Cout <= 0 || Cin <= 0 || k <= 0) return 0;
  return soda::conv_tc4_packed_bytes(Cout, Cin, k);
}

extern "C" int soda_pconv_pack_weights(const float* weight, const float* scale, const float* mult, int Cout, int Cin,
                                       int k, int in_fmt, void* packed, void* stream) {
  using namespace soda;
  if (!weight || !packed) return fail(SODA_ERR_BAD_ARG, "pconv_pack: null pointer");
  if (Cout <= 0 || Cin <= 0 || k <= 0) return fail(SODA_ERR_BAD_ARG, "pconv_pack: non-positive size");
  return conv_tc4_pack(weight, scale, mult, Cout, Cin, k, in_fmt, packed, (cudaStream_t)stream);
}

extern "C" int soda_pconv_supported(const soda_pconv_desc_t* d, const void* x, const void* y, const void* residual) {
  if (!d || !x || !y) return 0;
  return soda::conv_tc4_supported(d, x, y, residual) ? 1 : 0;
}

extern "C" int soda_pconv_fwd(const soda_pconv_desc_t* d, const void* x, const void* packed, const float* descale,
                              const float* shift, const void* residual, void* y, unsigned int* status, void* stream) {
  using namespace soda;
  if (!d || !x || !packed || !y || !descale) return fail(SODA_ERR_BAD_ARG, "pconv: null pointer");
  if (d->N <= 0 || d->H <= 0 || d->W <= 0 || d->Cin <= 0 || d->Cout <= 0 || d->k <= 0 || d->stride <= 0)
    return fail(SODA_ERR_BAD_ARG, "pconv: non-positive size");
  return conv_tc4_dispatch(d, x, packed, descale, shift, residual, y, status, nullptr, stream);
}

extern "C" int soda_pconv_lif_fwd(const soda_pconv_desc_t* d, const void* x, const void* packed, const float* descale,
                                  const float* shift, const void* residual, int T, const soda_neuron_params_t* np,
                                  const float* v0, const float* i0, void* y, float* v_out, float* i_out,
                                  unsigned int* status, void* stream) {
  using namespace soda;
  if (!d || !x || !packed || !y || !descale || !np || !v_out || !i_out)
    return fail(SODA_ERR_BAD_ARG, "pconv_lif: null pointer");
  if (d->N <= 0 || d->H <= 0 || d->W <= 0 || d->Cin <= 0 || d->Cout <= 0 || d->k <= 0 || d->stride <= 0)
    return fail(SODA_ERR_BAD_ARG, "pconv_lif: non-positive size");
  Tc4Lif l;
  l.T = T, l.np = to_np(np), l.v0 = v0, l.i0 = i0, l.v_out = v_out, l.i_out = i_out;
  return conv_tc4_dispatch(d, x, packed, descale, shift, residual, y, status, &l, stream);
}
