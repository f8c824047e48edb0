// Event2D decode + binary-occupancy binning (soda_bin_events).
//
// Reference arithmetic: prophesee_toolbox/src/io/dat_events_tools.py:100-108 (unpack) and
// utils/datasets.py:420-433 / :326-336 (time bin + features[ts, p, y, x] = 1).
//
// HBM-bound integer work.  Algorithmic bytes: 8 B per event read + one fp32 frame written.
//
// Sorted path (the .dat contract): one CTA owns (sample b, step t, row band).  It locates the
// step's record range with a warp-wide 32-ary search on the time stamps, sets bits of a
// shared-memory occupancy bitmap with shared atomics (atomicOr is idempotent, which is exactly
// the reference's "= 1" semantics), then expands the bitmap to fp32 {0,1} with 16-byte streaming
// stores.  Every output byte is written exactly once and no zero-fill pass over HBM is needed.
//
// Unsorted path: fill kernel + one scattered 4-byte store per event.
#include "common.cuh"

namespace soda {

struct Ev {
  uint32_t t;
  uint32_t w;
};

__device__ __forceinline__ void decode(uint32_t w, int& x, int& y, int& p) {
  x = (int)(w & 0x3FFFu);           // w & 16383
  y = (int)((w >> 14) & 0x3FFFu);   // (w & 268419072) >> 14
  p = (int)((w >> 28) & 1u);        // (w & 268435456) >> 28
}

// lower_bound over ev[a..b).t for the first record with t >= key (key as int64); whole warp.
__device__ int64_t warp_lower_bound(const Ev* __restrict__ ev, int64_t a, int64_t b, int64_t key) {
  const int lane = threadIdx.x & 31;
  while (true) {
    int64_t len = b - a;
    if (len <= 32) {
      bool pred = false;
      if (lane < len) pred = (int64_t)__ldg(&ev[a + lane].t) < key;
      unsigned m = __ballot_sync(0xffffffffu, pred);
      return a + __popc(m);
    }
    int64_t pos = a + ((len * (lane + 1)) >> 5) - 1;
    bool pred = (int64_t)__ldg(&ev[pos].t) < key;
    unsigned m = __ballot_sync(0xffffffffu, pred);
    int cnt = __popc(m);
    int64_t pos_hi = __shfl_sync(0xffffffffu, pos, cnt < 32 ? cnt : 31);
    int64_t pos_lo = __shfl_sync(0xffffffffu, pos, cnt > 0 ? cnt - 1 : 0);
    if (cnt == 32) return b;  // last probe is b-1 and it is < key
    if (cnt > 0) a = pos_lo + 1;
    b = pos_hi;
  }
}

template <bool kVec4>
__global__ void __launch_bounds__(256)
bin_bitmap_kernel(const Ev* __restrict__ ev, const int64_t* __restrict__ offs,
                  const int64_t* __restrict__ t0_us, int64_t step_us, int T, int H, int W,
                  int band_rows, int clip_x, float* __restrict__ frames, int64_t stride_t,
                  int64_t stride_b, int* __restrict__ n_dropped) {
  extern __shared__ uint32_t bits[];  // [2][band_rows][W] bits, padded to 128-bit words
  __shared__ int64_t s_range[2];

  const int band = blockIdx.x, t = blockIdx.y, b = blockIdx.z;
  const int y0 = band * band_rows;
  const int rows = min(band_rows, H - y0);
  const int plane_bits = band_rows * W;
  const int n_words = (2 * plane_bits + 31) >> 5;

  for (int i = threadIdx.x; i < n_words; i += blockDim.x) bits[i] = 0u;

  const int64_t ev_a = offs[b], ev_b = offs[b + 1];
  const int64_t lo_key = t0_us[b] + (int64_t)t * step_us;
  const int warp = threadIdx.x >> 5;
  if (warp < 2) {
    int64_t r = warp_lower_bound(ev, ev_a, ev_b, lo_key + (warp ? step_us : 0));
    if ((threadIdx.x & 31) == 0) s_range[warp] = r;
  }
  __syncthreads();
  const int64_t r0 = s_range[0], r1 = s_range[1];

  int dropped = 0;
  for (int64_t e = r0 + threadIdx.x; e < r1; e += blockDim.x) {
    const uint2 raw = __ldg(reinterpret_cast<const uint2*>(ev) + e);
    int x, y, p;
    decode(raw.y, x, y, p);
    if (clip_x) x = min(x, W - 1);
    if (x >= W || y >= H) {
      dropped += (band == 0);
      continue;
    }
    const int yy = y - y0;
    if (yy < 0 || yy >= rows) continue;
    const int bit = p * plane_bits + yy * W + x;
    atomicOr(&bits[bit >> 5], 1u << (bit & 31));
  }
  if (n_dropped != nullptr && dropped) atomicAdd(n_dropped, dropped);
  __syncthreads();

  float* out = frames + (int64_t)t * stride_t + (int64_t)b * stride_b;
  const int row_elems = rows * W;  // per polarity, contiguous rows y0..y0+rows-1
  if (kVec4) {
    const int n4 = row_elems >> 2;  // W % 4 == 0 guaranteed by the launcher
    for (int p = 0; p < 2; ++p) {
      float4* dst = reinterpret_cast<float4*>(out + ((int64_t)p * H + y0) * W);
      const int base = p * plane_bits;
      for (int i = threadIdx.x; i < n4; i += blockDim.x) {
        const int bit = base + (i << 2);
        const uint32_t nib = (bits[bit >> 5] >> (bit & 31)) & 0xFu;  // 4 | bit offset, no straddle
        float4 v;
        v.x = (nib & 1u) ? 1.0f : 0.0f;
        v.y = (nib & 2u) ? 1.0f : 0.0f;
        v.z = (nib & 4u) ? 1.0f : 0.0f;
        v.w = (nib & 8u) ? 1.0f : 0.0f;
        __stcs(dst + i, v);
      }
    }
  } else {
    for (int p = 0; p < 2; ++p) {
      float* dst = out + ((int64_t)p * H + y0) * W;
      const int base = p * plane_bits;
      for (int i = threadIdx.x; i < row_elems; i += blockDim.x) {
        const int bit = base + i;
        dst[i] = ((bits[bit >> 5] >> (bit & 31)) & 1u) ? 1.0f : 0.0f;
      }
    }
  }
}

__global__ void fill_frames_kernel(float* __restrict__ frames, int64_t stride_t, int64_t stride_b,
                                   int64_t frame_elems, int B) {
  // grid: (chunks, T*B)
  const int tb = blockIdx.y;
  float* dst = frames + (int64_t)(tb / B) * stride_t + (int64_t)(tb % B) * stride_b;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < frame_elems;
       i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = 0.0f;
}

__global__ void scatter_events_kernel(const Ev* __restrict__ ev, const int64_t* __restrict__ offs,
                                      const int64_t* __restrict__ t0_us, int64_t step_us, int T,
                                      int H, int W, int clip_x, float* __restrict__ frames,
                                      int64_t stride_t, int64_t stride_b,
                                      int* __restrict__ n_dropped) {
  const int b = blockIdx.y;
  const int64_t a = offs[b], e_end = offs[b + 1];
  const int64_t t0 = t0_us[b];
  for (int64_t e = a + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < e_end;
       e += (int64_t)gridDim.x * blockDim.x) {
    const uint2 raw = __ldg(reinterpret_cast<const uint2*>(ev) + e);
    const int64_t dtm = (int64_t)raw.x - t0;
    if (dtm < 0) continue;
    const int64_t ts = dtm / step_us;
    if (ts >= T) continue;
    int x, y, p;
    decode(raw.y, x, y, p);
    if (clip_x) x = min(x, W - 1);
    if (x >= W || y >= H) {
      if (n_dropped != nullptr) atomicAdd(n_dropped, 1);
      continue;
    }
    frames[ts * stride_t + (int64_t)b * stride_b + ((int64_t)p * H + y) * W + x] = 1.0f;
  }
}

}  // namespace soda

extern "C" int soda_bin_events(const void* events, const int64_t* sample_offsets,
                               const int64_t* t0_us, int B, int64_t step_us, int T, int H, int W,
                               int clip_x, int sorted_by_time, float* frames, int64_t stride_t,
                               int64_t stride_b, int* n_dropped, void* stream) {
  using namespace soda;
  if (!sample_offsets || !t0_us || !frames) return fail(SODA_ERR_BAD_ARG, "bin_events: null pointer");
  if (B <= 0 || T <= 0 || H <= 0 || W <= 0 || step_us <= 0)
    return fail(SODA_ERR_BAD_ARG, "bin_events: non-positive size");
  if (H > 16384 || W > 16384) return fail(SODA_ERR_BAD_ARG, "bin_events: H/W exceed the 14-bit Event2D range");
  if ((reinterpret_cast<uintptr_t>(events) & 7) != 0)
    return fail(SODA_ERR_BAD_ARG, "bin_events: event buffer must be 8-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  const Ev* ev = reinterpret_cast<const Ev*>(events);

  if (sorted_by_time) {
    // Band height: enough CTAs for >= 4 waves of 148 SMs when possible, bitmap <= 32 KB.
    int64_t want = ceil_div(4 * kNumSMs, (int64_t)B * T);
    int bands = (int)(want < 1 ? 1 : want);
    if (bands > H) bands = H;
    int band_rows = (int)ceil_div(H, bands);
    const int max_rows = (32 * 1024 * 8) / (2 * W);
    if (band_rows > max_rows) band_rows = max_rows;
    if (band_rows < 1) return fail(SODA_ERR_UNSUPPORTED, "bin_events: row too wide for the bitmap");
    bands = (int)ceil_div(H, band_rows);
    const size_t smem = (size_t)(((2 * (int64_t)band_rows * W + 127) / 128) * 16);
    dim3 grid(bands, T, B);
    const bool vec4 = (W % 4 == 0) && (stride_t % 4 == 0) && (stride_b % 4 == 0) &&
                      ((reinterpret_cast<uintptr_t>(frames) & 15) == 0);
    if (vec4)
      bin_bitmap_kernel<true><<<grid, 256, smem, st>>>(ev, sample_offsets, t0_us, step_us, T, H, W,
                                                       band_rows, clip_x, frames, stride_t,
                                                       stride_b, n_dropped);
    else
      bin_bitmap_kernel<false><<<grid, 256, smem, st>>>(ev, sample_offsets, t0_us, step_us, T, H,
                                                        W, band_rows, clip_x, frames, stride_t,
                                                        stride_b, n_dropped);
    return after_launch("bin_bitmap_kernel");
  }

  const int64_t frame_elems = 2LL * H * W;
  dim3 fgrid((unsigned)std::min<int64_t>(ceil_div(frame_elems, 256 * 8), 1024), (unsigned)(T * B));
  fill_frames_kernel<<<fgrid, 256, 0, st>>>(frames, stride_t, stride_b, frame_elems, B);
  int rc = after_launch("fill_frames_kernel");
  if (rc) return rc;
  dim3 sgrid(kNumSMs * 4, B);
  scatter_events_kernel<<<sgrid, 256, 0, st>>>(ev, sample_offsets, t0_us, step_us, T, H, W, clip_x,
                                               frames, stride_t, stride_b, n_dropped);
  return after_launch("scatter_events_kernel");
}
