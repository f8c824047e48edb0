// tcgen05 / TMEM / TMA implicit-GEMM convolution (soda_conv2d_fwd kind SODA_CONV_TC and the fused
// soda_conv_lif_fwd).
//
// Reference arithmetic: nn.Conv2d(bias=False, padding=k/2) models/modules/layer_gen.py:129-136, the eval
// BatchNorm affine (:211-214), norse LIFCell (:233), Residual sum (models/generator.py:146) and the
// concat-slice store (:158) in the epilogue.
//
//   D[m, co] = sum_{tap, ci} A_tap[m, ci] * W[tap][co][ci]          m = output pixel, 128 per tile
//
// Design (each point is a measured fact about B200, see DESIGN.md and scripts/microbench/):
//
// * "im2col by descriptor".  The swizzle of a SWIZZLE_128B shared-memory matrix descriptor is a function of
//   the absolute shared-memory address, so a descriptor may start at any 128-byte row and its 8-row groups
//   may be any multiple of 16 B apart.  A tile is 16 x 8 output pixels; TMA loads its 18 x 10 input halo
//   ONCE per 32-channel slab (out-of-bounds = the convolution's zero padding) as 180 rows of 128 B, and the
//   A operand of filter tap (ky, kx) is simply the descriptor {start = slab + (ky*10 + kx)*128 B, group
//   stride = 10*128 B}: the nine taps read the same slab, nothing is gathered or copied.  1x1 layers load
//   the tile itself; stride-2 layers load one dense tap tile per filter tap through the tensor map's element
//   strides (2.25x the input instead of 9x).
// * Precision: "tf32 + bf16 correction", 2 tensor passes instead of 3xTF32's 3:
//       a*b ~= hi(a)*hi(b)            kind::tf32, exact products, fp32 accumulate  -> main accumulator
//            + bf16(hi a)*bf16(lo b)
//            + bf16(lo a)*bf16(hi b)  kind::f16, K-stacked in one 128-byte row     -> correction accumulator
//   The correction terms are 2^-11 of the main term, so 8 mantissa bits on them keep ~2^-20 overall.
//   Each slab is split ONCE in shared memory by 8 split warps into a second plane of [bf16 hi | bf16 lo] rows
//   with the same row addressing, not once per tap; the fp32 plane itself is the tf32 operand as it is
//   (kind::tf32 truncates), so the main pass starts as soon as the TMA bytes land.  Inputs declared exactly
//   representable (spikes, pooled spike counts) skip the lo(a) half.
// * All synchronisation is per slab (9 k-blocks = 72 MMAs for a 3x3 layer), not per k-block: an mbarrier
//   round trip costs several hundred clocks, a k-block of a 32-channel tile only 128.
// * B (weights: tf32 hi plane + bf16 [lo|hi] plane, BatchNorm scale folded) is SHARED-MEMORY RESIDENT for
//   the whole kernel when one output-channel chunk of it fits; otherwise it streams through its own ring.
// * Two MMA issuer warps, one per accumulator (main / correction): independent instruction streams.
// * Epilogue: 8 warps (two per TMEM lane quarter, splitting the columns): main + correction + shift
//   (+ residual) -> global / concat slice.  LIF variant: the work item is (sample, tile, channel chunk)
//   and the kernel walks t = 0..T-1 INSIDE the item, so membrane potential and synaptic current stay in
//   the epilogue warps' registers for all T steps and never touch HBM.
//
// Persistent, warp-specialised, 20 warps: w0-7 split, w8-15 epilogue, w16 A producer (TMA), w17 main-pass MMA
// issuer + TMEM owner, w18 B producer (TMA), w19 correction-pass MMA issuer.
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "common.cuh"
#include "neuron.cuh"
#include "tc_common.cuh"

namespace soda {

using namespace tc;

constexpr int T3_BM = 128;
constexpr int T3_BK = 32;
constexpr int T3_BH = 16, T3_BW = 8;     // spatial tile (one 8-row descriptor group per output row)
constexpr int T3_THREADS = 640;
constexpr int T3_W_PROD_A = 16, T3_W_MMA_MAIN = 17, T3_W_PROD_B = 18, T3_W_MMA_CORR = 19;
constexpr int T3_MAX_RING = 8;           // max stages of the A / B shared-memory rings
constexpr int T3_STAGE_BYTES = 8 * 1536;  // per epilogue warp: 1 KB transpose staging (16 rows x 16 cols) + 512 B row-offset table
constexpr int T3_SMEM_LIMIT = 232448 - 1024 /*alignment slack*/ - 512 /*barriers*/ - T3_STAGE_BYTES;

struct Tc3Params {
  int H, W, Cin, Cout, k, stride, pad, Ho, Wo;
  int tiles_x, tiles_y, n_spatial;
  int mt;                        // M-tiles per work item sharing every streamed weight k-block (1 or 2)
  int n_sp_items;                // ceil(n_spatial / mt)
  int it_step_sp, it_step_outer; // (gridDim.x / nch) decomposed over (sp, outer): per-CTA item iterator increments
  int taps, cblocks, kblocks;
  int mode;                      // 0 = HALO (one A stage per 32-channel slab, `taps` k-blocks), 1 = TAP
  int flat;                      // 1x1 stride-1 over the flattened pixel list
  int a_exact;
  int spk;                       // input exact in bf16 (spikes, pooled spike counts): bf16-only products against the
                                 // [bf16 hi | bf16 mid] weight plane - no tf32 pass, no main weight plane
  uint32_t b_stage_bytes;        // one k-block of weights in shared memory: BN*128 (spk) or 2*BN*128
  int rows;                      // 128-byte rows in one A plane (180 halo pixels, or 128)
  int WW;                        // halo row pitch in pixels
  uint32_t sbo;                  // bytes between the 8-row groups of the A operand
  uint32_t plane_bytes;          // hi plane size = offset of the correction plane inside a stage
  int nch;                       // output-channel chunks of BN
  int n_outer, T_inner, Bsz;     // work items = nch * n_spatial * n_outer, each walks T_inner frames
  int n_items;
  int b_stationary, b_stages;
  int a_stages;                  // fp32 planes in the TMA ring
  int c_stages;                  // correction planes (written by the split warps, read by the correction MMAs)
  uint32_t a_stage_bytes, a_tx_bytes, b_region_bytes;
  int64_t total_pix;
  int64_t out_bs, out_ps, res_bs, res_ps;
  const float* shift;
  const float* residual;
  float* y;
  int prefetch;                  // A producer issues L2 prefetches one item ahead
  int vec_store;                 // Cout % 4 == 0 and 16-byte aligned output / residual: coalesced staged stores
  int dbg;                       // SODA_TC_DBG timing experiments (results are garbage when set)
  unsigned long long* tl;        // dbg & 64: per-role clock64 timeline of CTA 0's first tiles (TL_TILES x 16 events)
  int lif;
  NP np;
  const float* v0;
  const float* i0;
  float* v_out;
  float* i_out;
};

constexpr int TL_TILES = 40;
// events: 0 producer saw a_empty / issues TMA, 1 split saw a_full, 2 split arrived a_ready, 3 main-MMA saw a_full,
// 4 main-MMA committed the tile, 5 corr-MMA saw a_ready, 6 corr-MMA committed, 7 epilogue saw acc_full,
// 8 epilogue released the accumulator, 9 main-MMA saw acc_empty (tile start)
// Compiled in only with -DSODA_TC_TIMELINE (make TIMELINE=1) and then enabled by SODA_TC_DBG=64.
__device__ __forceinline__ void tl_mark(const Tc3Params& p, int tile, int ev) {
#ifdef SODA_TC_TIMELINE
  if ((p.dbg & 64) && blockIdx.x == 0 && tile < TL_TILES && (threadIdx.x & 31) == 0)
    p.tl[tile * 16 + ev] = (unsigned long long)clock64();
#endif
}

struct Tc3Item {
  int chunk, sp, outer;
};
__device__ __forceinline__ Tc3Item tc3_item(const Tc3Params& p, int item) {
  Tc3Item it;
  it.chunk = item % p.nch;
  const int r = item / p.nch;
  it.sp = r % p.n_sp_items;      // tiles sp * mt .. sp * mt + mt - 1
  it.outer = r / p.n_sp_items;
  return it;
}

// A CTA's items are blockIdx.x + k * gridDim.x.  gridDim.x is a multiple of nch, so the channel chunk is fixed and
// (sp, outer) advance by a constant step: no integer divisions in the per-tile paths (they cost the epilogue warps
// ~1200 clk per tile when done with / and % - measured with the SODA_TC_DBG=64 timeline).
struct Tc3Iter {
  int sp, outer;
  __device__ __forceinline__ void init(const Tc3Params& p) {
    const int r0 = (int)blockIdx.x / p.nch;
    sp = r0 % p.n_sp_items;
    outer = r0 / p.n_sp_items;
  }
  __device__ __forceinline__ void next(const Tc3Params& p) {
    sp += p.it_step_sp;
    outer += p.it_step_outer;
    if (sp >= p.n_sp_items) sp -= p.n_sp_items, ++outer;
  }
};

// ================================ MMA issuers ===================================================
// One warp per accumulator.  CORR = false: tf32 main pass over the hi planes; CORR = true: bf16 correction
// pass over the [bf16 hi | bf16 lo] x [bf16 lo | bf16 hi] planes.  Warp-uniform control flow; one elected
// lane issues, so descriptors live in uniform registers.
template <int BN, bool CORR, int MT, bool SPK>
__device__ __forceinline__ void tc3_mma_role(const Tc3Params& p, uint32_t tmem_base, uint32_t smem_a_u32,
                                             uint32_t smem_b_u32,
                                             uint64_t* a_ready /* CORR: correction plane written (c_full); main: TMA landed */,
                                             uint64_t* a_empty, uint64_t* c_empty,
                                             uint64_t* b_full, uint64_t* b_empty, uint64_t* acc_full,
                                             uint64_t* acc_empty) {
  static_assert(CORR || !SPK, "the bf16-only scheme has no tf32 stream");
  constexpr uint32_t B_TILE = BN * 128, B_STAGE = SPK ? B_TILE : 2 * B_TILE;
  constexpr uint32_t idesc = CORR ? idesc_bf16(T3_BM, BN) : idesc_tf32(T3_BM, BN);
  // + (bytes >> 4) on a descriptor moves its start address
  // fp32 planes [a_stages] then correction planes [c_stages]: two rings, so that the TMA ring can be deeper than
  // the short-lived correction planes (measured: with 32 KB coupled stages the ring of a 1x1 layer held ONE tile and
  // the tile period was TMA latency + one stage of MMAs)
  const uint64_t a_desc0 = smem_desc_sw128_sbo(smem_a_u32 + (CORR ? (uint32_t)p.a_stages * p.plane_bytes : 0u), p.sbo);
  const uint64_t b_desc0 = smem_desc_sw128(smem_b_u32 + ((CORR && !SPK) ? B_TILE : 0u));
  const bool half = CORR && p.a_exact;   // exact inputs have no lo(a) half
  const bool stat = p.b_stationary != 0;
  const bool halo3 = p.mode == 0 && p.k == 3;
  const int stages_per_tile = p.mode == 0 ? p.cblocks : p.kblocks;
  const int kb_per_stage = p.mode == 0 ? p.taps : 1;
  const uint32_t a_stage16 = p.plane_bytes >> 4;
  uint32_t sc = 0, pc = 0;    // correction ring slot / parity (CORR only)
  constexpr uint32_t NACC = BN <= 64 ? 4 : 2;   // accumulator stages of (main | correction) in TMEM
  uint32_t sa = 0, pa = 0;    // A ring stage / parity
  uint32_t sb = 0, pb = 0;    // B ring stage / parity (streamed B)
  uint32_t cnt = 0;           // running accumulator-stage counter
  constexpr int mt = MT;
  if (stat) mbar_wait(&b_full[0], 0);
  for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
    for (int t = 0; t < p.T_inner; ++t) {
      uint32_t d_acc[2];
      _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < mt) {
        const uint32_t a = (cnt + m) % NACC, ph = ((cnt + m) / NACC) & 1u;
        mbar_wait(&acc_empty[a], ph ^ 1);
        d_acc[m] = tmem_base + a * (2 * BN) + ((CORR && !SPK) ? (uint32_t)BN : 0u);
      }
      if (!CORR) tl_mark(p, (int)cnt, 9);
      tc_fence_after();
      uint32_t accum = 0;
      for (int st = 0; st < stages_per_tile; ++st) {
        // the mt tiles of the item own consecutive ring stages for this slab / k-block
        uint32_t s_m[2], c_m[2];
        uint64_t a_st[2];
        _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < mt) {
          if (CORR) {
            mbar_wait(&a_ready[sc], pc);
            c_m[m] = sc;
            a_st[m] = a_desc0 + (uint64_t)(sc * a_stage16);
            if (++sc == (uint32_t)p.c_stages) sc = 0, pc ^= 1;
          } else {
            mbar_wait(&a_ready[sa], pa);
            a_st[m] = a_desc0 + (uint64_t)(sa * a_stage16);
          }
          s_m[m] = sa;
          if (++sa == (uint32_t)p.a_stages) sa = 0, pa ^= 1;
        }
        if (st == 0) tl_mark(p, (int)cnt, CORR ? 5 : 3);
        tc_fence_after();
        int ky = 0, kx = 0;
        for (int tp = 0; tp < kb_per_stage; ++tp) {
          // resident weights are stored (tap, cb); HALO stage = channel slab st, k-block = tap tp
          uint64_t b_desc;
          if (stat) {
            const int bi = p.mode == 0 ? tp * p.cblocks + st : st;
            b_desc = b_desc0 + (uint64_t)((uint32_t)bi * (B_STAGE >> 4));
          } else {
            mbar_wait(&b_full[sb], pb);
            tc_fence_after();
            b_desc = b_desc0 + (uint64_t)(sb * (B_STAGE >> 4));
          }
          // tap (ky, kx) of the halo = the same slab, started (ky * WW + kx) rows further
          const uint64_t tap_off = (uint64_t)(halo3 ? (uint32_t)(ky * p.WW + kx) * 8u : 0u);
          if (elect_one()) {
            // 32 bytes along the 128-byte row = 8 tf32 / 16 bf16 K elements; descriptor unit is 16 B
            if (!(p.dbg & 4)) {
              _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < mt) {
                const uint64_t a_desc = a_st[m] + tap_off;
                if (SPK) {
                  // a(32 ch, bf16, first half of the row) x [W_hi | W_mid]: the A descriptor revisits the same 64 bytes
                  mma_bf16_ss(d_acc[m], a_desc, b_desc, idesc, accum);
                  mma_bf16_ss(d_acc[m], a_desc + 2, b_desc + 2, idesc, 1);
                  mma_bf16_ss(d_acc[m], a_desc, b_desc + 4, idesc, 1);
                  mma_bf16_ss(d_acc[m], a_desc + 2, b_desc + 6, idesc, 1);
                } else if (CORR) {
                  mma_bf16_ss(d_acc[m], a_desc, b_desc, idesc, accum);
                  mma_bf16_ss(d_acc[m], a_desc + 2, b_desc + 2, idesc, 1);
                  if (!half) {
                    mma_bf16_ss(d_acc[m], a_desc + 4, b_desc + 4, idesc, 1);
                    mma_bf16_ss(d_acc[m], a_desc + 6, b_desc + 6, idesc, 1);
                  }
                } else {
                  mma_tf32_ss(d_acc[m], a_desc, b_desc, idesc, accum);
                  mma_tf32_ss(d_acc[m], a_desc + 2, b_desc + 2, idesc, 1);
                  mma_tf32_ss(d_acc[m], a_desc + 4, b_desc + 4, idesc, 1);
                  mma_tf32_ss(d_acc[m], a_desc + 6, b_desc + 6, idesc, 1);
                }
              }
            }
            if (!stat) mma_commit(&b_empty[sb]);
          }
          accum = 1;
          if (++kx == 3) kx = 0, ++ky;
          if (!stat && ++sb == (uint32_t)p.b_stages) sb = 0, pb ^= 1;
        }
        if (elect_one()) {
          _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < mt) {
            if (!SPK) mma_commit(&a_empty[s_m[m]]);   // SPK: the split warps release the fp32 plane themselves
            if (CORR) mma_commit(&c_empty[c_m[m]]);
          }
        }
      }
      if (elect_one()) {
        _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < mt) mma_commit(&acc_full[(cnt + m) % NACC]);
      }
      tl_mark(p, (int)cnt, CORR ? 6 : 4);
      cnt += mt;
    }
  }
}

template <int BN, bool LIF, bool SPK>
__global__ void __launch_bounds__(T3_THREADS, 1)
conv_tc3_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, Tc3Params p) {
  constexpr uint32_t B_TILE = BN * 128;        // one plane of one k-block: BN rows x 128 B
  constexpr uint32_t B_STAGE = SPK ? B_TILE : 2 * B_TILE;   // tf32 hi plane + bf16 [lo|hi] plane, or the [hi|mid] plane alone
  constexpr uint32_t NACC = BN <= 64 ? 4 : 2;  // accumulator stages of (main | correction)
  constexpr uint32_t TMEM_COLS = NACC * 2 * BN;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_b = smem;
  uint8_t* smem_a = smem + p.b_region_bytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_a + (size_t)(p.a_stages + p.c_stages) * p.plane_bytes);
  uint64_t* a_full = bars;                        // TMA bytes of an A stage landed            [8]
  uint64_t* a_ready = bars + 8;                   // stage split (hi in place + correction plane) [8]
  uint64_t* a_empty = bars + 16;                  // both MMA streams done with the stage     [8]
  uint64_t* b_full = bars + 24;                   // [8]
  uint64_t* b_empty = bars + 32;                  // [8]
  uint64_t* acc_full = bars + 40;                 // [4]
  uint64_t* acc_empty = bars + 44;                // [4]
  uint64_t* c_empty = bars + 48;                  // correction MMAs done with a correction plane  [8]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 56);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == T3_W_PROD_A && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
    for (int s = 0; s < T3_MAX_RING; ++s) {
      mbar_init(&a_full[s], 1);
      mbar_init(&a_ready[s], 8);      // one arrival per split warp
      mbar_init(&a_empty[s], SPK ? 8 : 2);   // one commit per MMA stream; spk: one arrival per split warp
      mbar_init(&b_full[s], 1);
      mbar_init(&b_empty[s], SPK ? 1 : 2);
      mbar_init(&c_empty[s], 1);      // the correction MMA stream's commit
    }
    for (int a = 0; a < 4; ++a) {
      mbar_init(&acc_full[a], SPK ? 1 : 2);
      mbar_init(&acc_empty[a], 8);    // one arrival per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == T3_W_MMA_MAIN) tmem_alloc(tmem_slot, TMEM_COLS);
  pdl_trigger();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Launched as a programmatic dependent: everything above (barrier init, TMEM allocation, descriptor prefetch)
  // overlapped the tail of the previous kernel of the stream; its results are visible from here on.
  pdl_wait();
  const int my_chunk = blockIdx.x % p.nch;   // gridDim.x is a multiple of nch: fixed per CTA

  // Roles 16-19 run warp-uniform control flow and elect one lane only around the asynchronous instructions
  // themselves: descriptors, coordinates and barrier addresses then live in uniform registers.
  if (warp == T3_W_PROD_A) {
    // ================================ A producer (TMA) =========================================
    uint32_t s = 0, ph = 0;  // ring position / parity
    int tl_tile = 0;
    const uint32_t smem_a_u32 = smem_u32(smem_a);
    Tc3Iter it;
    it.init(p);
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, it.next(p)) {
      int x0[2] = {0, 0}, y0[2] = {0, 0};
      _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < p.mt) {
        // a tile index past the last one (odd tile count) loads zeros (out of bounds) and is never stored
        const int spt = it.sp * p.mt + m;
        if (p.flat) {
          x0[m] = spt * T3_BM;
        } else {
          const int ty = spt / p.tiles_x, tx = spt - ty * p.tiles_x;
          y0[m] = ty * T3_BH * p.stride - p.pad;
          x0[m] = tx * T3_BW * p.stride - p.pad;
        }
      }
      for (int t = 0; t < p.T_inner; ++t) {
        const int img = p.flat ? 0 : t * p.Bsz + it.outer;
        const int n_tap = p.mode == 0 ? 1 : p.taps;   // HALO: one stage per channel slab
        if (p.prefetch) {
          // L2 prefetch of what this CTA loads next (its next item, or the next time step of this one): the
          // shared-memory ring holds only 2-6 stages, far less than HBM latency x bandwidth per SM
          int pimg = img, px0[2] = {x0[0], x0[1]}, py0[2] = {y0[0], y0[1]};
          bool have = true;
          if (t + 1 < p.T_inner) {
            pimg = img + p.Bsz;
          } else if (item + (int)gridDim.x < p.n_items) {
            const Tc3Item nx = tc3_item(p, item + (int)gridDim.x);
            pimg = p.flat ? 0 : nx.outer;
            _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < p.mt) {
              const int spt = nx.sp * p.mt + m;
              if (p.flat) {
                px0[m] = spt * T3_BM, py0[m] = 0;
              } else {
                const int ty = spt / p.tiles_x, tx = spt - ty * p.tiles_x;
                py0[m] = ty * T3_BH * p.stride - p.pad;
                px0[m] = tx * T3_BW * p.stride - p.pad;
              }
            }
          } else {
            have = false;
          }
          if (have && elect_one()) {
            int qy = 0, qx = 0;
            for (int tap = 0; tap < n_tap; ++tap) {
              for (int cb = 0; cb < p.cblocks; ++cb)
                _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < p.mt)
                  tma_prefetch_4d(&tmA, cb * T3_BK, px0[m] + qx, py0[m] + qy, pimg);
              if (++qx == p.k) qx = 0, ++qy;
            }
          }
          __syncwarp();
        }
        int ky = 0, kx = 0;
        for (int tap = 0; tap < n_tap; ++tap) {
          for (int cb = 0; cb < p.cblocks; ++cb) {
            _Pragma("unroll") for (int m = 0; m < 2; ++m) if (m < p.mt) {
              mbar_wait(&a_empty[s], ph ^ 1);
              if (tap == 0 && cb == 0 && m == 0) tl_mark(p, tl_tile, 0);
              if (elect_one()) {
                mbar_expect_tx(&a_full[s], p.a_tx_bytes);
                tma_load_4d_s(smem_a_u32 + s * p.plane_bytes, &tmA, &a_full[s], cb * T3_BK, x0[m] + kx, y0[m] + ky,
                              img);
              }
              __syncwarp();
              if (++s == (uint32_t)p.a_stages) s = 0, ph ^= 1;
            }
          }
          if (++kx == p.k) kx = 0, ++ky;
        }
        tl_tile += p.mt;
      }
    }
  } else if (warp == T3_W_PROD_B) {
    // ================================ B producer (TMA) =========================================
    const uint32_t smem_b_u32 = smem_u32(smem_b);
    if (p.b_stationary) {
      if (elect_one()) mbar_expect_tx(&b_full[0], (uint32_t)p.kblocks * B_STAGE);
      __syncwarp();
      for (int tap = 0; tap < p.taps; ++tap)
        for (int cb = 0; cb < p.cblocks; ++cb) {
          const uint32_t dst = smem_b_u32 + (uint32_t)(tap * p.cblocks + cb) * B_STAGE;
          if (elect_one()) {
            if (SPK) {
              tma_load_3d_s(dst, &tmB, &b_full[0], cb * T3_BK, my_chunk * BN, 2 * p.taps + tap);
            } else {
              tma_load_3d_s(dst, &tmB, &b_full[0], cb * T3_BK, my_chunk * BN, tap);
              tma_load_3d_s(dst + B_TILE, &tmB, &b_full[0], cb * T3_BK, my_chunk * BN, p.taps + tap);
            }
          }
          __syncwarp();
        }
    } else {
      uint32_t s = 0, ph = 0;
      const int n_o = p.mode == 0 ? p.cblocks : p.taps;   // HALO walks (cb, tap), TAP walks (tap, cb)
      const int n_i = p.mode == 0 ? p.taps : p.cblocks;
      for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
        for (int t = 0; t < p.T_inner; ++t) {
          for (int o = 0; o < n_o; ++o)
            for (int i = 0; i < n_i; ++i) {
              const int tap = p.mode == 0 ? i : o, cb = p.mode == 0 ? o : i;
              mbar_wait(&b_empty[s], ph ^ 1);
              if (elect_one()) {
                mbar_expect_tx(&b_full[s], B_STAGE);
                const uint32_t dst = smem_b_u32 + s * B_STAGE;
                if (SPK) {
                  tma_load_3d_s(dst, &tmB, &b_full[s], cb * T3_BK, my_chunk * BN, 2 * p.taps + tap);
                } else {
                  tma_load_3d_s(dst, &tmB, &b_full[s], cb * T3_BK, my_chunk * BN, tap);
                  tma_load_3d_s(dst + B_TILE, &tmB, &b_full[s], cb * T3_BK, my_chunk * BN, p.taps + tap);
                }
              }
              __syncwarp();
              if (++s == (uint32_t)p.b_stages) s = 0, ph ^= 1;
            }
        }
      }
    }
  } else if (warp == T3_W_MMA_MAIN) {
    // bf16-only scheme (SPK): no tf32 stream, this warp only owns the TMEM allocation
    if constexpr (!SPK) {
      if (p.mt == 2)
        tc3_mma_role<BN, false, 2, false>(p, tmem_base, smem_u32(smem_a), smem_u32(smem_b), a_full, a_empty, c_empty, b_full,
                                          b_empty, acc_full, acc_empty);
      else
        tc3_mma_role<BN, false, 1, false>(p, tmem_base, smem_u32(smem_a), smem_u32(smem_b), a_full, a_empty, c_empty, b_full,
                                          b_empty, acc_full, acc_empty);
    }
  } else if (warp == T3_W_MMA_CORR) {
    if (p.mt == 2)
      tc3_mma_role<BN, true, 2, SPK>(p, tmem_base, smem_u32(smem_a), smem_u32(smem_b), a_ready, a_empty, c_empty, b_full,
                                     b_empty, acc_full, acc_empty);
    else
      tc3_mma_role<BN, true, 1, SPK>(p, tmem_base, smem_u32(smem_a), smem_u32(smem_b), a_ready, a_empty, c_empty, b_full,
                                     b_empty, acc_full, acc_empty);
  } else if (warp < 8) {
    // ============ split: fp32 slab -> tf32 hi (in place) + [bf16 hi | bf16 lo] correction plane =========
    // Task = (pixel row P, 8 channels q).  SWIZZLE_128B: logical 16-byte chunk c of row P sits at chunk
    // position c ^ (P & 7) in both planes (TMA wrote the hi plane that way; the correction plane copies the
    // convention so that one descriptor arithmetic serves both).
    const uint32_t smem_a_u32 = smem_u32(smem_a);
    const int n_task = p.rows * 4;
    const int stages_per_tile = (p.mode == 0 ? p.cblocks : p.kblocks) * p.mt;
    uint32_t s = 0, ph = 0;
    uint32_t c = 0, cph = 0;   // correction ring slot / parity
    int tl_tile = 0;
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
      for (int t = 0; t < p.T_inner; ++t, tl_tile += p.mt) {
        for (int st = 0; st < stages_per_tile; ++st) {
          mbar_wait(&a_full[s], ph);
          if (st == 0 && warp == 0) tl_mark(p, tl_tile, 1);
          mbar_wait(&c_empty[c], cph ^ 1);
          const uint32_t hi_pl = smem_a_u32 + s * p.plane_bytes;
          const uint32_t cr_pl = smem_a_u32 + ((uint32_t)p.a_stages + c) * p.plane_bytes;
          if (!(p.dbg & 1)) {
            for (int task = (int)threadIdx.x; task < n_task; task += 256) {
              const uint32_t P = (uint32_t)task >> 2, q = (uint32_t)task & 3u, sw = P & 7u;
              const uint32_t a0 = hi_pl + P * 128u + (((2u * q) ^ sw) << 4);
              const uint32_t a1 = hi_pl + P * 128u + (((2u * q + 1u) ^ sw) << 4);
              const float4 v0 = lds128(a0), v1 = lds128(a1);
              const float f[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
              if constexpr (SPK) {
                // values are exact in bf16: one conversion, first half of the row only
                sts128(cr_pl + P * 128u + ((q ^ sw) << 4), __uint_as_float(pack_bf16x2(f[0], f[1])),
                       __uint_as_float(pack_bf16x2(f[2], f[3])), __uint_as_float(pack_bf16x2(f[4], f[5])),
                       __uint_as_float(pack_bf16x2(f[6], f[7])));
                continue;
              }
              float fh[8], fl[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                fh[e] = __uint_as_float(__float_as_uint(f[e]) & 0xFFFFE000u);
                fl[e] = f[e] - fh[e];
              }
              const uint32_t cr = cr_pl + P * 128u;
              sts128(cr + ((q ^ sw) << 4), __uint_as_float(pack_bf16x2(fh[0], fh[1])),
                     __uint_as_float(pack_bf16x2(fh[2], fh[3])), __uint_as_float(pack_bf16x2(fh[4], fh[5])),
                     __uint_as_float(pack_bf16x2(fh[6], fh[7])));
              // The fp32 plane is left as TMA wrote it: kind::tf32 reads the upper 19 bits of each element and
              // ignores the rest (measured: bit-identical results with and without rewriting hi in place),
              // so hi(a) = truncation is exactly what the main pass sees.
              if (!p.a_exact) {
                sts128(cr + (((4u + q) ^ sw) << 4), __uint_as_float(pack_bf16x2(fl[0], fl[1])),
                       __uint_as_float(pack_bf16x2(fl[2], fl[3])), __uint_as_float(pack_bf16x2(fl[4], fl[5])),
                       __uint_as_float(pack_bf16x2(fl[6], fl[7])));
              }
            }
          }
          // generic-proxy writes must be visible to the tensor core's (async proxy) operand reads
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            mbar_arrive(&a_ready[c]);
            if (SPK) mbar_arrive(&a_empty[s]);   // the fp32 plane has been read; nobody else needs it
          }
          if (++c == (uint32_t)p.c_stages) c = 0, cph ^= 1;
          if (st == 0 && warp == 0) tl_mark(p, tl_tile, 2);
          if (++s == (uint32_t)p.a_stages) s = 0, ph ^= 1;
        }
      }
    }
  } else if (warp < 16) {
    // ================================ epilogue ================================================
    // Two warps per TMEM lane quarter; warpgroup e owns columns [e*BN/2, (e+1)*BN/2) of the tile.
    constexpr int NCOL = BN / 2;
    static_assert(!LIF || NCOL <= 32, "fused LIF keeps 2*NCOL state registers per thread");
    const int e = (warp - 8) >> 2;
    const int q = warp & 3;
    const uint32_t stage_u32 = smem_u32(reinterpret_cast<uint8_t*>(bars) + 512) + (uint32_t)(warp - 8) * 1536u;
    const uint32_t table_u32 = stage_u32 + 1024u;
    const int row = q * 32 + lane;
    const int64_t hw = (int64_t)p.Ho * p.Wo;
    uint32_t local = 0;
    Tc3Iter it;
    it.init(p);
    for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, it.next(p)) {
      for (int m = 0; m < p.mt; ++m) {   // (mt == 2 only without LIF: T_inner == 1)
      const int spt = it.sp * p.mt + m;
      bool valid;
      int64_t rr;      // pixel index inside its frame
      int64_t frame0;  // frame of step t = 0
      if (p.flat) {
        // flat mode is only planned when output (and residual) frames are dense in pixels, so the flattened
        // pixel index addresses them directly: frame 0, "row" = global pixel
        const int64_t pix = (int64_t)spt * T3_BM + row;
        valid = pix < p.total_pix;
        frame0 = 0;
        rr = pix;
      } else {
        const int ty = spt / p.tiles_x, tx = spt - ty * p.tiles_x;
        const int oy = ty * T3_BH + (row >> 3);
        const int ox = tx * T3_BW + (row & 7);
        valid = oy < p.Ho && ox < p.Wo;
        rr = (int64_t)oy * p.Wo + ox;
        frame0 = it.outer;
      }
      const int co_base = my_chunk * BN + e * NCOL;
      float sv[LIF ? NCOL : 1], si[LIF ? NCOL : 1];
      if constexpr (LIF) {
        const int64_t soff = ((int64_t)it.outer * hw + rr) * p.Cout + co_base;
#pragma unroll
        for (int c = 0; c < NCOL; c += 4) {
          float4 a = make_float4(p.np.v_leak, p.np.v_leak, p.np.v_leak, p.np.v_leak);
          float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
          if (valid && p.v0 != nullptr && co_base + c < p.Cout) {
            a = __ldg(reinterpret_cast<const float4*>(p.v0 + soff + c));
            b = __ldg(reinterpret_cast<const float4*>(p.i0 + soff + c));
          }
          sv[c] = a.x, sv[c + 1] = a.y, sv[c + 2] = a.z, sv[c + 3] = a.w;
          si[c] = b.x, si[c + 1] = b.y, si[c + 2] = b.z, si[c + 3] = b.w;
        }
      }
      for (int t = 0; t < p.T_inner; ++t, ++local) {
        const uint32_t acc = local % NACC;
        const uint32_t acc_ph = (local / NACC) & 1u;
        const int64_t frame = frame0 + (int64_t)t * p.Bsz;
        const int64_t out_off = frame * p.out_bs + rr * p.out_ps;
        const int64_t res_off = frame * p.res_bs + rr * p.res_ps;
        if (warp == 8) tl_mark(p, (int)local, 10);
        if (p.vec_store) {
          // row -> (output offset | -1, residual offset) table of this warp's 32 rows for the staged stores
          __syncwarp();
          const int64_t my_out = valid ? out_off : -1;
          asm volatile("st.shared.v2.b64 [%0], {%1, %2};" ::"r"(table_u32 + (uint32_t)lane * 16u), "l"(my_out), "l"(res_off)
                       : "memory");
          __syncwarp();
        }
        mbar_wait(&acc_full[acc], acc_ph);
        if (warp == 8) tl_mark(p, (int)local, 7);
        tc_fence_after();
        const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + acc * (2 * BN) + e * NCOL;
#pragma unroll
        for (int c0 = 0; c0 < NCOL; c0 += 16) {
          uint32_t r[16], rc[16];
          tmem_ld_32x16(t_row + c0, r);
          if constexpr (!SPK) tmem_ld_32x16(t_row + BN + c0, rc);
          tmem_ld_wait();
          if (c0 == 0 && warp == 8) tl_mark(p, (int)local, 11);
          if (c0 + 16 >= NCOL) {
            // last TMEM read of this tile: hand the accumulator back before the global traffic
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[acc]);
            if (warp == 8) tl_mark(p, (int)local, 8);
          }
          const int co0 = co_base + c0;
          const bool live = valid && co0 < p.Cout;
          float o[16];
#pragma unroll
          for (int jj = 0; jj < 16; ++jj)
            o[jj] = SPK ? __uint_as_float(r[jj]) : __uint_as_float(r[jj]) + __uint_as_float(rc[jj]);
          if (p.shift != nullptr && co0 < p.Cout) {
            if (co0 + 16 <= p.Cout) {
#pragma unroll
              for (int jj = 0; jj < 16; jj += 4) {
                const float4 sh = __ldg(reinterpret_cast<const float4*>(p.shift + co0 + jj));
                o[jj] += sh.x, o[jj + 1] += sh.y, o[jj + 2] += sh.z, o[jj + 3] += sh.w;
              }
            } else {
#pragma unroll
              for (int jj = 0; jj < 16; ++jj)
                if (co0 + jj < p.Cout) o[jj] += __ldg(p.shift + co0 + jj);
            }
          }
          if constexpr (LIF) {
#pragma unroll
            for (int jj = 0; jj < 16; ++jj) {
              float z, vd;
              lif_step(p.np, o[jj], sv[c0 + jj], si[c0 + jj], z, vd);
              o[jj] = z;
            }
          }
          if (p.vec_store) {
            // Coalesced store: the warp transposes its 32 rows x 16 columns through 1 KB of shared memory,
            // 16 rows at a time, so that one store instruction writes 8 rows x 64 contiguous bytes (8 sectors
            // pairs) instead of 32 scattered 16-byte pieces; the residual is read with the same mapping.
            if (!(p.dbg & 8)) {
#pragma unroll
              for (int half = 0; half < 2; ++half) {
                __syncwarp();
                if ((lane >> 4) == half) {
                  const uint32_t r16 = lane & 15u;
                  const uint32_t rowa = stage_u32 + r16 * 64u;
                  const uint32_t f = (r16 >> 1) & 3u;
#pragma unroll
                  for (uint32_t j = 0; j < 4; ++j)
                    sts128(rowa + ((j ^ f) << 4), o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
                }
                __syncwarp();
#pragma unroll
                for (int it = 0; it < 2; ++it) {
                  const uint32_t rrow = (uint32_t)it * 8u + ((uint32_t)lane >> 2);
                  const uint32_t j = lane & 3u;
                  const float4 val = lds128(stage_u32 + rrow * 64u + ((j ^ ((rrow >> 1) & 3u)) << 4));
                  int64_t ooff, roff;
                  asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];"
                               : "=l"(ooff), "=l"(roff)
                               : "r"(table_u32 + ((uint32_t)half * 16u + rrow) * 16u));
                  const int col = co0 + (int)j * 4;
                  if (ooff >= 0 && col < p.Cout) {
                    float4 ov = val;
                    if (p.residual != nullptr) {
                      const float4 rv = __ldg(reinterpret_cast<const float4*>(p.residual + roff + col));
                      ov.x += rv.x, ov.y += rv.y, ov.z += rv.z, ov.w += rv.w;
                    }
                    *reinterpret_cast<float4*>(p.y + ooff + col) = ov;
                  }
                }
              }
            }
          } else if (live && !(p.dbg & 8)) {
            float* dst = p.y + out_off + co0;
            const float* rs = p.residual ? p.residual + res_off + co0 : nullptr;
            const int nvalid = min(16, p.Cout - co0);
#pragma unroll
            for (int jj = 0; jj < 16; ++jj) {
              if (jj < nvalid) {
                float ov = o[jj];
                if (rs) ov += __ldg(rs + jj);
                dst[jj] = ov;
              }
            }
          }
        }
        if (warp == 8) tl_mark(p, (int)local, 12);
      }
      if constexpr (LIF) {
        if (valid && co_base < p.Cout) {
          const int64_t soff = ((int64_t)it.outer * hw + rr) * p.Cout + co_base;
#pragma unroll
          for (int c = 0; c < NCOL; c += 4) {
            *reinterpret_cast<float4*>(p.v_out + soff + c) = make_float4(sv[c], sv[c + 1], sv[c + 2], sv[c + 3]);
            *reinterpret_cast<float4*>(p.i_out + soff + c) = make_float4(si[c], si[c + 1], si[c + 2], si[c + 3]);
          }
        }
      }
      }  // m
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == T3_W_MMA_MAIN) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------
// host side: weight packing, tensor maps, configuration, dispatch
// ------------------------------------------------------------------------------------------------
static inline int tc3_bn_for(int Cout, bool lif) {
  if (Cout <= 32) return 32;
  if (Cout <= 64 || lif) return 64;
  return 128;
}
// The packed layout pads Cout to a multiple of 128 (or to 32 / 64 below that) so that every BN the
// dispatcher may pick (plain and LIF launches share one packed buffer) addresses whole tiles.
static inline int tc3_cout_pad(int Cout) {
  if (Cout <= 32) return 32;
  if (Cout <= 64) return 64;
  return (Cout + 127) / 128 * 128;
}

int64_t conv_tc3_packed_elems(int Cout, int Cin, int k) {
  if (Cin % T3_BK != 0) return 0;
  return 3LL * k * k * tc3_cout_pad(Cout) * Cin;
}

__device__ __forceinline__ uint16_t tc3_bf16_bits(float f) {
  return (uint16_t)(pack_bf16x2(f, 0.0f) & 0xFFFFu);
}

// [Cout, Cin, k, k] -> plane 0: [tap][Cout_pad][Cin] tf32 hi of (w * scale)
//                      plane 1: [tap][Cout_pad][Cin/32] rows of 128 B = bf16 lo[32] | bf16 hi[32]
//                      plane 2: same rows = bf16 rn(w)[32] | bf16 rn(w - rn(w))[32]   (bf16-only scheme for
//                               inputs exact in bf16: 16+ significant weight bits, no tf32 pass)
__global__ void pack_weights_tc3_kernel(const float* __restrict__ w, const float* __restrict__ scale, int Cout,
                                        int Cin, int taps, int Cout_pad, float* __restrict__ wp) {
  const int64_t plane = (int64_t)taps * Cout_pad * Cin;
  uint16_t* corr = reinterpret_cast<uint16_t*>(wp + plane);
  uint16_t* spk = reinterpret_cast<uint16_t*>(wp + 2 * plane);
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
    const int64_t row16 = (r * Cin + (ci & ~31)) * 2;  // start of this 128-byte row, in uint16
    corr[row16 + (ci & 31)] = tc3_bf16_bits(v - hi);
    corr[row16 + 32 + (ci & 31)] = tc3_bf16_bits(hi);
    const uint16_t wh = tc3_bf16_bits(v);
    spk[row16 + (ci & 31)] = wh;
    spk[row16 + 32 + (ci & 31)] = tc3_bf16_bits(v - __uint_as_float((uint32_t)wh << 16));
  }
}

int conv_tc3_pack(const float* w, const float* scale, int Cout, int Cin, int k, float* packed, cudaStream_t st) {
  if (Cin % T3_BK != 0) return fail(SODA_ERR_UNSUPPORTED, "conv_tc: Cin must be a multiple of 32");
  const int64_t plane = (int64_t)k * k * tc3_cout_pad(Cout) * Cin;
  pack_weights_tc3_kernel<<<(unsigned)std::min<int64_t>(ceil_div(plane, 256), 4096), 256, 0, st>>>(
      w, scale, Cout, Cin, k * k, tc3_cout_pad(Cout), packed);
  return after_launch("pack_weights_tc3_kernel");
}

typedef CUresult (*EncodeTiledFn3)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn3 get_encode_fn3() {
  static EncodeTiledFn3 fn = nullptr;
  if (fn == nullptr) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn3>(ptr);
  }
  return fn;
}

static int encode_map3(CUtensorMap* m, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides_b,
                       const cuuint32_t* box, const cuuint32_t* estr) {
  EncodeTiledFn3 fn = get_encode_fn3();
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

template <int BN, bool LIF, bool SPK>
static int launch_tc3_(const CUtensorMap& tmA, const CUtensorMap& tmB, const Tc3Params& p, int grid, size_t smem,
                      cudaStream_t st) {
  {   // the opt-in is per device: set on every launch (cheap), not once per process
    cudaError_t e = cudaFuncSetAttribute(conv_tc3_kernel<BN, LIF, SPK>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         232448);
    if (e != cudaSuccess) {
      set_error("conv_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return (int)SODA_ERR_CUDA;
    }
  }
  static const bool pdl = []() { const char* e = getenv("SODA_PDL"); return !(e && atoi(e) == 0); }();
  if (pdl) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)grid), cfg.blockDim = dim3(T3_THREADS);
    cfg.dynamicSmemBytes = smem, cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at, cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, conv_tc3_kernel<BN, LIF, SPK>, tmA, tmB, p);
  } else {
    conv_tc3_kernel<BN, LIF, SPK><<<grid, T3_THREADS, smem, st>>>(tmA, tmB, p);
  }
  return after_launch("conv_tc3_kernel");
}

template <int BN, bool LIF>
static int launch_tc3(const CUtensorMap& tmA, const CUtensorMap& tmB, const Tc3Params& p, int grid, size_t smem,
                      cudaStream_t st) {
  return p.spk ? launch_tc3_<BN, LIF, true>(tmA, tmB, p, grid, smem, st)
               : launch_tc3_<BN, LIF, false>(tmA, tmB, p, grid, smem, st);
}

bool conv_tc3_supported(const soda_conv_desc_t* d, const float* x) {
  if (d->in_nchw || d->Cin % T3_BK != 0) return false;
  if (d->in_pix_stride % 4 != 0 || d->in_batch_stride % 4 != 0) return false;
  if ((reinterpret_cast<uintptr_t>(x) & 15) != 0) return false;
  if (d->k != 1 && d->k != 3) return false;
  if (d->stride != 1 && d->stride != 2) return false;
  return true;
}

struct Tc3Lif {
  int T;
  NP np;
  const float* v0;
  const float* i0;
  float* v_out;
  float* i_out;
};

int conv_tc3_dispatch(const soda_conv_desc_t* d, const float* x, const float* packed, const float* shift,
                      const float* residual, float* y, const Tc3Lif* lif, void* stream) {
  if (!conv_tc3_supported(d, x))
    return fail(SODA_ERR_UNSUPPORTED,
                "conv_tc: needs pixel-major input, Cin % 32 == 0, 16-byte aligned strides, k in {1,3}, stride in {1,2}");
  Tc3Params p;
  memset(&p, 0, sizeof(p));
  {
    const char* e = getenv("SODA_TC_DBG");
    p.dbg = e ? atoi(e) : 0;
  }
  p.H = d->H, p.W = d->W, p.Cin = d->Cin, p.Cout = d->Cout, p.k = d->k, p.stride = d->stride;
  p.pad = d->k / 2;
  p.Ho = (d->H + 2 * p.pad - d->k) / d->stride + 1;
  p.Wo = (d->W + 2 * p.pad - d->k) / d->stride + 1;
  if (p.Ho <= 0 || p.Wo <= 0) return fail(SODA_ERR_BAD_ARG, "conv: empty output");
  p.taps = d->k * d->k;
  p.cblocks = d->Cin / T3_BK;
  p.kblocks = p.taps * p.cblocks;
  p.a_exact = (d->flags & (SODA_CONV_FLAG_INPUT_EXACT_TF32 | SODA_CONV_FLAG_INPUT_EXACT_BF16)) ? 1 : 0;
  {
    const char* e = getenv("SODA_TC_SPK");   // 0: keep the tf32 + correction scheme for bf16-exact inputs (A/B runs)
    p.spk = (d->flags & SODA_CONV_FLAG_INPUT_EXACT_BF16) && !(e && atoi(e) == 0) ? 1 : 0;
  }
  p.total_pix = (int64_t)d->N * p.Ho * p.Wo;
  p.out_bs = d->out_batch_stride, p.out_ps = d->out_pix_stride;
  p.res_bs = d->res_batch_stride, p.res_ps = d->res_pix_stride;
  p.shift = shift, p.residual = residual, p.y = y;
  p.vec_store = d->Cout % 4 == 0 && d->out_pix_stride % 4 == 0 && d->out_batch_stride % 4 == 0 &&
                (reinterpret_cast<uintptr_t>(y) & 15) == 0 &&
                (residual == nullptr || (d->res_pix_stride % 4 == 0 && d->res_batch_stride % 4 == 0 &&
                                         (reinterpret_cast<uintptr_t>(residual) & 15) == 0));
  const bool is_lif = lif != nullptr;
  if (is_lif) {
    if (lif->T <= 0 || d->N % lif->T != 0) return fail(SODA_ERR_BAD_ARG, "conv_lif: N must be T * batch");
    if (d->Cout % 16 != 0) return fail(SODA_ERR_UNSUPPORTED, "conv_lif: Cout must be a multiple of 16");
    if (!lif->v_out || !lif->i_out) return fail(SODA_ERR_BAD_ARG, "conv_lif: null state output");
    if ((lif->v0 == nullptr) != (lif->i0 == nullptr)) return fail(SODA_ERR_BAD_ARG, "conv_lif: v0 / i0 must come together");
    p.lif = 1, p.np = lif->np, p.v0 = lif->v0, p.i0 = lif->i0, p.v_out = lif->v_out, p.i_out = lif->i_out;
    p.T_inner = lif->T, p.Bsz = d->N / lif->T, p.n_outer = p.Bsz;
  } else {
    p.T_inner = 1, p.Bsz = d->N, p.n_outer = d->N;
  }
  const int BN = tc3_bn_for(d->Cout, is_lif);
  const int Cout_pad = tc3_cout_pad(d->Cout);
  p.nch = (d->Cout + BN - 1) / BN;
  const uint32_t b_stage = (p.spk ? 1u : 2u) * BN * 128u;
  p.b_stage_bytes = b_stage;

  CUtensorMap tmA, tmB;
  const int64_t HW = (int64_t)d->H * d->W;
  const bool frames_dense = d->N == 1 || d->in_batch_stride == HW * d->in_pix_stride;
  p.mode = d->stride == 1 ? 0 : 1;
  const int64_t HWo = (int64_t)p.Ho * p.Wo;
  const bool out_dense = (d->N == 1 || d->out_batch_stride == HWo * d->out_pix_stride) &&
                         (residual == nullptr || d->N == 1 || d->res_batch_stride == HWo * d->res_pix_stride);
  p.flat = (d->k == 1 && d->stride == 1 && frames_dense && out_dense && !is_lif && !(p.dbg & 32)) ? 1 : 0;   // dbg 32: tiled 1x1
  p.rows = T3_BM, p.WW = T3_BW, p.sbo = 1024;
  if (p.flat) {
    p.n_spatial = (int)ceil_div(p.total_pix, T3_BM);
    p.tiles_x = p.n_spatial, p.tiles_y = 1;
    p.n_outer = 1;
    const cuuint64_t dims[4] = {(cuuint64_t)d->Cin, (cuuint64_t)p.total_pix, 1, 1};
    const cuuint64_t str[3] = {(cuuint64_t)d->in_pix_stride * 4, (cuuint64_t)p.total_pix * d->in_pix_stride * 4,
                               (cuuint64_t)p.total_pix * d->in_pix_stride * 4};
    const cuuint32_t box[4] = {T3_BK, T3_BM, 1, 1};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    int rc = encode_map3(&tmA, x, 4, dims, str, box, es);
    if (rc) return rc;
  } else {
    p.tiles_y = (int)ceil_div(p.Ho, T3_BH), p.tiles_x = (int)ceil_div(p.Wo, T3_BW);
    p.n_spatial = p.tiles_x * p.tiles_y;
    const cuuint64_t dims[4] = {(cuuint64_t)d->Cin, (cuuint64_t)d->W, (cuuint64_t)d->H, (cuuint64_t)d->N};
    const cuuint64_t str[3] = {(cuuint64_t)d->in_pix_stride * 4, (cuuint64_t)d->W * d->in_pix_stride * 4,
                               (cuuint64_t)(d->N == 1 ? HW * d->in_pix_stride : d->in_batch_stride) * 4};
    if (p.mode == 0) {
      p.WW = T3_BW + 2 * p.pad;
      const int HH = T3_BH + 2 * p.pad;
      p.rows = p.WW * HH;
      p.sbo = (uint32_t)p.WW * 128u;
      const cuuint32_t box[4] = {T3_BK, (cuuint32_t)p.WW, (cuuint32_t)HH, 1};
      const cuuint32_t es[4] = {1, 1, 1, 1};
      int rc = encode_map3(&tmA, x, 4, dims, str, box, es);
      if (rc) return rc;
    } else {
      const cuuint32_t box[4] = {T3_BK, (cuuint32_t)(T3_BW * d->stride), (cuuint32_t)(T3_BH * d->stride), 1};
      const cuuint32_t es[4] = {1, (cuuint32_t)d->stride, (cuuint32_t)d->stride, 1};
      int rc = encode_map3(&tmA, x, 4, dims, str, box, es);
      if (rc) return rc;
    }
  }
  p.a_tx_bytes = (uint32_t)p.rows * 128u;
  p.plane_bytes = (p.a_tx_bytes + 1023u) & ~1023u;
  p.a_stage_bytes = 2u * p.plane_bytes;
  {
    const cuuint64_t dims[3] = {(cuuint64_t)d->Cin, (cuuint64_t)Cout_pad, (cuuint64_t)(3 * p.taps)};
    const cuuint64_t str[2] = {(cuuint64_t)d->Cin * 4, (cuuint64_t)Cout_pad * d->Cin * 4};
    const cuuint32_t box[3] = {T3_BK, (cuuint32_t)BN, 1};
    const cuuint32_t es[3] = {1, 1, 1};
    int rc = encode_map3(&tmB, packed, 3, dims, str, box, es);
    if (rc) return rc;
  }
  // shared-memory plan (in planes of plane_bytes): fp32 TMA ring of a_stages planes + c_stages correction planes,
  // weights resident if one chunk fits beside a useful ring, else a weight ring as well
  const int64_t b_total = (int64_t)p.kblocks * b_stage;
  const bool halo3 = p.mode == 0 && p.k == 3;
  const int64_t plane = p.plane_bytes;
  const char* e_as = getenv("SODA_TC_ASTAGES");   // experiments: cap on the fp32 ring depth
  const int a_cap = e_as ? std::max(2, std::min(T3_MAX_RING, atoi(e_as))) : T3_MAX_RING;
  const char* e_cs = getenv("SODA_TC_CSTAGES");
  const int c_env = e_cs ? std::max(2, std::min(T3_MAX_RING, atoi(e_cs))) : 0;
  if (b_total + 5 * plane <= T3_SMEM_LIMIT && b_total < (1 << 20)) {
    p.b_stationary = 1, p.b_stages = 1, p.mt = 1;
    p.b_region_bytes = (uint32_t)b_total;
    // two correction planes are enough (SODA_TC_CSTAGES sweep: 3-4 planes gain nothing in isolation and cost 15 % of
    // the whole step by shortening the fp32 ring)
    p.c_stages = c_env > 0 ? c_env : 2;
    p.a_stages = (int)std::min<int64_t>(a_cap, (T3_SMEM_LIMIT - b_total) / plane - p.c_stages);
  } else {
    // Streamed weights: every k-block (2 * BN * 128 B) comes from L2 once per work item, and at one 128-pixel tile
    // per item that stream alone exceeds what L2 can deliver to 148 SMs (measured floor: the load skeleton of the
    // stride-2 and 768-channel layers ran at ~12.7 TB/s of weight traffic).  So an item takes TWO M-tiles that
    // share each weight k-block (two accumulators), halving the stream.  Only where TMEM still double-buffers
    // (BN <= 64: 4 accumulator stages; at BN = 128 the pair would serialise MMA and epilogue - measured slower) and
    // not with LIF (state registers).
    p.b_stationary = 0;
    const char* e_mt = getenv("SODA_TC_MT");
    const bool want_pair = !is_lif && p.n_spatial > 1 && (e_mt ? atoi(e_mt) == 2 : BN <= 64);
    p.mt = want_pair ? 2 : 1;
    p.c_stages = c_env > 0 ? c_env : 2;
    // weight ring first (>= 4 stages if possible), the rest goes to the fp32 ring
    const int a_floor = p.mt == 2 ? 2 : 2;
    int bs = 4;
    while (bs > 2 && (int64_t)bs * b_stage + (a_floor + p.c_stages) * plane > T3_SMEM_LIMIT) --bs;
    int64_t a = (T3_SMEM_LIMIT - (int64_t)bs * b_stage) / plane - p.c_stages;
    if (a < a_floor) return fail(SODA_ERR_UNSUPPORTED, "conv_tc: shared-memory plan does not fit");
    const int a_pref = halo3 ? 3 : 6;   // beyond this, extra room serves the weight ring better
    p.a_stages = (int)std::min<int64_t>(std::min<int64_t>(a, a_cap), p.mt == 2 ? std::max(a_pref, 4) : a_pref);
    p.b_stages = (int)std::min<int64_t>(T3_MAX_RING, (T3_SMEM_LIMIT - (int64_t)(p.a_stages + p.c_stages) * plane) / b_stage);
    if (p.b_stages < 2) return fail(SODA_ERR_UNSUPPORTED, "conv_tc: shared-memory plan does not fit");
    p.b_region_bytes = (uint32_t)p.b_stages * b_stage;
  }
  p.n_sp_items = (p.n_spatial + p.mt - 1) / p.mt;
  {
    // measured on B200: no gain (-1 % .. +8 % time) - the hand-off chain, not HBM latency, paces the ring; off by default
    const char* e_pf = getenv("SODA_TC_PREFETCH");
    p.prefetch = e_pf ? atoi(e_pf) : 0;
  }
  p.n_items = p.nch * p.n_sp_items * p.n_outer;
  int grid = p.n_items < kNumSMs ? p.n_items : kNumSMs;
  grid = grid / p.nch * p.nch;
  if (grid < p.nch) grid = p.nch;
  p.it_step_sp = (grid / p.nch) % p.n_sp_items;
  p.it_step_outer = (grid / p.nch) / p.n_sp_items;
  const size_t smem = (size_t)p.b_region_bytes + (size_t)(p.a_stages + p.c_stages) * p.plane_bytes + 512 + T3_STAGE_BYTES + 1024;
  cudaStream_t st = (cudaStream_t)stream;
#ifdef SODA_TC_TIMELINE
  if (p.dbg & 64) {
    // debug only: synchronous timeline dump of CTA 0's first tiles (clock64 per role, relative to the first event)
    static unsigned long long* d_tl = nullptr;
    const size_t nb = (size_t)TL_TILES * 16 * sizeof(unsigned long long);
    if (d_tl == nullptr) cudaMalloc(&d_tl, nb);
    cudaMemsetAsync(d_tl, 0, nb, st);
    p.tl = d_tl;
    int rc;
    if (is_lif) rc = BN == 32 ? launch_tc3<32, true>(tmA, tmB, p, grid, smem, st) : launch_tc3<64, true>(tmA, tmB, p, grid, smem, st);
    else if (BN == 32) rc = launch_tc3<32, false>(tmA, tmB, p, grid, smem, st);
    else if (BN == 64) rc = launch_tc3<64, false>(tmA, tmB, p, grid, smem, st);
    else rc = launch_tc3<128, false>(tmA, tmB, p, grid, smem, st);
    static int dumps = 0;
    if (rc == 0 && dumps++ < 1) {
      unsigned long long h[TL_TILES * 16];
      cudaStreamSynchronize(st);
      cudaMemcpy(h, d_tl, nb, cudaMemcpyDeviceToHost);
      unsigned long long t0 = ~0ull;
      for (int i = 0; i < TL_TILES * 16; ++i) if (h[i] && h[i] < t0) t0 = h[i];
      fprintf(stderr, "conv_tc3 timeline (clk since first event) BN=%d mt=%d a_stages=%d b_stat=%d mode=%d items=%d grid=%d\n"
              "a_stages/c_stages/b_stages see header\n" "tile  prodTMA splitFul splitRdy mainFull mainCmt  corrRdy  corrCmt  epiFull  epiFree mainAccE epiStart  epiTmem   epiEnd\n",
              BN, p.mt, p.a_stages, p.b_stationary, p.mode, p.n_items, grid);
      for (int t = 0; t < TL_TILES; ++t) {
        fprintf(stderr, "%4d", t);
        for (int e = 0; e < 13; ++e) fprintf(stderr, " %8lld", h[t * 16 + e] ? (long long)(h[t * 16 + e] - t0) : -1LL);
        fprintf(stderr, "\n");
      }
    }
    return rc;
  }
#endif
  if (is_lif) {
    if (BN == 32) return launch_tc3<32, true>(tmA, tmB, p, grid, smem, st);
    return launch_tc3<64, true>(tmA, tmB, p, grid, smem, st);
  }
  if (BN == 32) return launch_tc3<32, false>(tmA, tmB, p, grid, smem, st);
  if (BN == 64) return launch_tc3<64, false>(tmA, tmB, p, grid, smem, st);
  return launch_tc3<128, false>(tmA, tmB, p, grid, smem, st);
}

int conv_tc3_lif_dispatch(const soda_conv_desc_t* d, const float* x, const float* packed, const float* shift,
                          const float* residual, int T, const soda_neuron_params_t* np, const float* v0,
                          const float* i0, float* y, float* v_out, float* i_out, void* stream) {
  Tc3Lif l;
  l.T = T, l.np = to_np(np), l.v0 = v0, l.i0 = i0, l.v_out = v_out, l.i_out = i_out;
  return conv_tc3_dispatch(d, x, packed, shift, residual, y, &l, stream);
}

}  // namespace soda
