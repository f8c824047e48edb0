// Anchor labelling for the detection loss (soda_roi_assign): SURVEY.md 8(f)-2, the step right after the forward in
// training (reference models/soda.py:259-266).
//
// Reference arithmetic, reproduced operation for operation:
//   utils/box.py:24-57    box_iou      : inter / ((area_a + area_g) - inter), areas = prod(b[2:] - b[:2])
//   utils/roi.py:65-109   _assign_anchor_to_box :
//        (1) every anchor takes its best ground-truth box (first maximal index) if that IoU >= threshold, else -1;
//        (2) G times: the globally largest remaining IoU (ties: smallest flattened index, torch.argmax) gives
//            (anchor, box); the anchor is re-labelled with the box, then the anchor's row and the box's column are
//            discarded.  Padded label rows (all -1, utils/datasets.py:127-135) take part exactly as in the reference:
//            their IoUs are all 0, so each claims the first anchor still free and labels it class 0 with mask 1.
//   utils/roi.py:38-60    class = label[box, 0].long() + 1 (0 = background), mask = (box >= 0) repeated over 4
//   utils/box.py:60-67    offset_boxes : 10 * (c_gt - c_anc) / wh_anc, 5 * log(1e-6 + wh_gt / wh_anc), times the mask
// The reference runs (2) as a Python loop of argmax over the [A, G] matrix plus two fills per box, and (1)-(3) once
// per sample; here one CTA per sample keeps the boxes in shared memory, recomputes IoUs on the fly (25 flops each,
// cheaper than an [A, G] matrix in HBM), does (2) from per-box cached maxima (one block-wide 64-bit max reduction per
// box, lazily redone when the cached anchor was claimed by another box) with the discarded rows in a shared-memory
// bitmap, and writes offsets / mask / classes in one final pass.  Integer/compare-bound: ~2 A G IoU evaluations.
#include "common.cuh"

namespace soda {

constexpr int ROI_THREADS = 1024;
constexpr int ROI_MAX_G = 256;

__device__ __forceinline__ float roi_iou(const float4& a, float area_a, const float4& g, float area_g) {
  const float lx = fmaxf(a.x, g.x), ly = fmaxf(a.y, g.y);
  const float hx = fminf(a.z, g.z), hy = fminf(a.w, g.w);
  const float iw = fmaxf(__fsub_rn(hx, lx), 0.0f), ih = fmaxf(__fsub_rn(hy, ly), 0.0f);
  const float inter = __fmul_rn(iw, ih);
  return __fdiv_rn(inter, __fsub_rn(__fadd_rn(area_a, area_g), inter));
}

// Orders like the reference's comparison of fp32 IoUs for every non-NaN value (IoUs are >= 0 or the -1 of a discarded
// entry, which is never a candidate here); a NaN (0/0: zero-area anchor against a zero-area box) maps above
// everything, as torch.argmax / torch.max treat it.
__device__ __forceinline__ uint32_t roi_order_bits(float v) {
  if (v != v) return 0xFFFFFFFFu;
  const uint32_t b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

__global__ void __launch_bounds__(ROI_THREADS, 1)
roi_assign_kernel(const float4* __restrict__ anchors, const float* __restrict__ labels, int A, int G, float thr,
                  float4* __restrict__ bbox_offset, float4* __restrict__ bbox_mask, long long* __restrict__ cls_out) {
  pdl_trigger();
  extern __shared__ uint32_t roi_smem[];
  __shared__ float4 s_box[ROI_MAX_G];
  __shared__ float s_area[ROI_MAX_G];
  __shared__ float s_cls[ROI_MAX_G];
  __shared__ unsigned char s_col_dead[ROI_MAX_G];
  __shared__ unsigned long long s_red[32];
  __shared__ unsigned long long s_best;
  __shared__ unsigned long long s_colkey[ROI_MAX_G];
  __shared__ unsigned char s_colzero[ROI_MAX_G];
  __shared__ int s_first_free;
  uint32_t* row_dead = roi_smem;                 // bitmap over anchors
  const int b = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* lab = labels + (int64_t)b * G * 5;
  long long* amap = cls_out + (int64_t)b * A;    // holds the anchor -> box map until the final pass
  for (int j = tid; j < G; j += ROI_THREADS) {
    const float4 g = make_float4(lab[j * 5 + 1], lab[j * 5 + 2], lab[j * 5 + 3], lab[j * 5 + 4]);
    s_box[j] = g;
    s_area[j] = __fmul_rn(__fsub_rn(g.z, g.x), __fsub_rn(g.w, g.y));
    s_cls[j] = lab[j * 5];
    s_col_dead[j] = 0;
  }
  for (int w = tid; w < (A + 31) / 32; w += ROI_THREADS) row_dead[w] = 0;
  __syncthreads();

  // (1) per-anchor best box above the threshold
  for (int a = tid; a < A; a += ROI_THREADS) {
    const float4 an = anchors[a];
    const float area_a = __fmul_rn(__fsub_rn(an.z, an.x), __fsub_rn(an.w, an.y));
    uint32_t best_bits = 0;
    float best = 0.f;
    int best_j = -1;
    for (int j = 0; j < G; ++j) {
      const float v = roi_iou(an, area_a, s_box[j], s_area[j]);
      const uint32_t ob = roi_order_bits(v);
      if (best_j < 0 || ob > best_bits) best_bits = ob, best = v, best_j = j;   // strict: first maximal index
    }
    amap[a] = (best_j >= 0 && best >= thr) ? best_j : -1;
  }
  __syncthreads();

  // (2) every box claims the best anchor still free.  Per box (column) the best free anchor is cached as a 64-bit key
  // (order bits of the IoU << 32 | ~(anchor * G + box): max = largest IoU, then smallest flattened index, torch.argmax's
  // rule).  Removing anchors can only lower a column's maximum, so a cached key is an upper bound: an iteration takes
  // the largest cached key among the free columns; if its anchor has been claimed meanwhile the column is rescanned
  // and the pick repeated.  A column whose maximum over the free anchors has dropped to 0 (padded -1 rows from the
  // start) stays at 0 for good and its best is simply the first free anchor - no rescan.
  auto scan_column = [&](int j) -> unsigned long long {   // block-wide; result valid in every thread
    unsigned long long key = 0;
    const float4 g = s_box[j];
    const float area_g = s_area[j];
    for (int a = tid; a < A; a += ROI_THREADS) {
      if (row_dead[a >> 5] & (1u << (a & 31))) continue;
      const float4 an = anchors[a];
      const float area_a = __fmul_rn(__fsub_rn(an.z, an.x), __fsub_rn(an.w, an.y));
      const float v = roi_iou(an, area_a, g, area_g);
      const uint32_t flat = (uint32_t)a * (uint32_t)G + (uint32_t)j;
      const unsigned long long k = ((unsigned long long)roi_order_bits(v) << 32) | (unsigned long long)(~flat);
      key = k > key ? k : key;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
      key = other > key ? other : key;
    }
    __syncthreads();   // s_red free again
    if (lane == 0) s_red[warp] = key;
    __syncthreads();
    key = s_red[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
      key = other > key ? other : key;
    }
    return key;
  };
  constexpr uint32_t ZERO_BITS = 0x80000000u;   // roi_order_bits(+0.0f)
  for (int j = 0; j < G; ++j) {
    const unsigned long long k = scan_column(j);
    if (tid == 0) s_colkey[j] = k, s_colzero[j] = (uint32_t)(k >> 32) == ZERO_BITS;
  }
  if (tid == 0) s_first_free = 0;
  __syncthreads();
  for (int it = 0; it < G; ++it) {
    bool exhausted = false;
    while (true) {
      if (warp == 0) {
        // largest key among the free columns; zero-class columns are keyed on the first free anchor
        const int a0 = s_first_free;
        unsigned long long key = 0;
        for (int j = lane; j < G; j += 32) {
          if (s_col_dead[j]) continue;
          unsigned long long k = s_colkey[j];
          if (s_colzero[j])
            k = a0 < A ? (((unsigned long long)ZERO_BITS << 32) | (unsigned long long)(~((uint32_t)a0 * (uint32_t)G + (uint32_t)j)))
                       : 0ull;
          key = k > key ? k : key;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
          key = other > key ? other : key;
        }
        if (lane == 0) s_best = key;
      }
      __syncthreads();
      const unsigned long long key = s_best;
      if (key == 0) {                           // no free anchor left (A < G)
        exhausted = true;
        break;
      }
      const uint32_t flat = ~(uint32_t)(key & 0xFFFFFFFFull);
      const int a = (int)(flat / (uint32_t)G), j = (int)(flat % (uint32_t)G);
      if (!(row_dead[a >> 5] & (1u << (a & 31)))) {
        __syncthreads();                        // everyone has read s_best / row_dead
        if (tid == 0) {
          amap[a] = j;
          row_dead[a >> 5] |= 1u << (a & 31);
          s_col_dead[j] = 1;
          int f = s_first_free;
          while (f < A && (row_dead[f >> 5] & (1u << (f & 31)))) ++f;
          s_first_free = f;
        }
        __syncthreads();
        break;
      }
      // stale: the cached anchor of column j was claimed by another box - rescan this column only
      const unsigned long long k = scan_column(j);
      if (tid == 0) s_colkey[j] = k, s_colzero[j] = (uint32_t)(k >> 32) == ZERO_BITS;
      __syncthreads();
    }
    if (exhausted) break;
  }
  __threadfence_block();
  __syncthreads();

  // (3) classes, mask, offsets
  for (int a = tid; a < A; a += ROI_THREADS) {
    const long long j = amap[a];
    float4 off = make_float4(0.f, 0.f, 0.f, 0.f), m = off;
    long long c = 0;
    if (j >= 0) {
      const float4 an = anchors[a], g = s_box[j];
      const float acx = __fdiv_rn(__fadd_rn(an.x, an.z), 2.0f), acy = __fdiv_rn(__fadd_rn(an.y, an.w), 2.0f);
      const float aw = __fsub_rn(an.z, an.x), ah = __fsub_rn(an.w, an.y);
      const float gcx = __fdiv_rn(__fadd_rn(g.x, g.z), 2.0f), gcy = __fdiv_rn(__fadd_rn(g.y, g.w), 2.0f);
      const float gw = __fsub_rn(g.z, g.x), gh = __fsub_rn(g.w, g.y);
      off.x = __fdiv_rn(__fmul_rn(10.0f, __fsub_rn(gcx, acx)), aw);
      off.y = __fdiv_rn(__fmul_rn(10.0f, __fsub_rn(gcy, acy)), ah);
      off.z = __fmul_rn(5.0f, logf(__fadd_rn(1e-6f, __fdiv_rn(gw, aw))));
      off.w = __fmul_rn(5.0f, logf(__fadd_rn(1e-6f, __fdiv_rn(gh, ah))));
      m = make_float4(1.f, 1.f, 1.f, 1.f);
      c = (long long)s_cls[j] + 1;
    }
    bbox_offset[(int64_t)b * A + a] = off;
    bbox_mask[(int64_t)b * A + a] = m;
    amap[a] = c;
  }
}

}  // namespace soda

extern "C" int soda_roi_assign(const float* anchors, const float* labels, int B, int A, int G, float iou_threshold,
                               float* bbox_offset, float* bbox_mask, long long* class_labels, void* stream) {
  using namespace soda;
  if (!anchors || !bbox_offset || !bbox_mask || !class_labels || (!labels && B > 0 && G > 0))
    return fail(SODA_ERR_BAD_ARG, "roi_assign: null pointer");
  if (B < 0 || A <= 0 || G < 0) return fail(SODA_ERR_BAD_ARG, "roi_assign: bad sizes");
  if (G > ROI_MAX_G) return fail(SODA_ERR_UNSUPPORTED, "roi_assign: more than 256 label rows per sample");
  if ((int64_t)A * (G > 0 ? G : 1) >= (1LL << 32)) return fail(SODA_ERR_UNSUPPORTED, "roi_assign: A * G >= 2^32");
  const size_t smem = (size_t)((A + 31) / 32) * 4;
  if (smem > 160 * 1024) return fail(SODA_ERR_UNSUPPORTED, "roi_assign: more than 1.3 M anchors");
  if ((reinterpret_cast<uintptr_t>(anchors) & 15) || (reinterpret_cast<uintptr_t>(bbox_offset) & 15) ||
      (reinterpret_cast<uintptr_t>(bbox_mask) & 15))
    return fail(SODA_ERR_BAD_ARG, "roi_assign: anchors and outputs must be 16-byte aligned");
  if (B == 0) return 0;
  {   // the opt-in is per device: set on every launch (cheap), not once per process
    cudaError_t e = cudaFuncSetAttribute(roi_assign_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    if (e != cudaSuccess) {
      set_error("roi_assign: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return (int)SODA_ERR_CUDA;
    }
  }
  roi_assign_kernel<<<B, ROI_THREADS, smem, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4*>(anchors), labels, A, G, iou_threshold, reinterpret_cast<float4*>(bbox_offset),
      reinterpret_cast<float4*>(bbox_mask), class_labels);
  return after_launch("roi_assign_kernel");
}
