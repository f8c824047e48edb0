// Detection decode + greedy per-class NMS + row assembly (soda_multibox_detection): SURVEY.md 8(f)-1, the step
// right after the spiking forward in SODa.predict (reference models/soda.py:226-233).
//
// Reference arithmetic, reproduced operation for operation:
//   utils/box.py:72-91   offset_inverse : anchors -> centre form, xy = off_xy * anc_wh / 10 + anc_xy,
//                                         wh = exp(off_wh / 5) * anc_wh, back to corners
//   utils/box.py:24-45   box_iou        : inter / (area_a + area_b - inter), areas = prod(b[2:] - b[:2])
//   utils/box.py:94-119  nms            : per foreground class, candidates sorted by confidence (descending), greedy:
//                                         keep the best, drop every later box with iou > threshold (a NaN iou drops too)
//   utils/box.py:122-153 multibox_detection : conf, class = max(cls_probs) ; class - 1 (background = -1); rows ordered
//                                         kept boxes first (class by class, best first) then the others by anchor index
//                                         with class -1; conf < pos_threshold -> class -1, conf = 1 - conf
// The reference runs this as a Python while-loop with one device sync per kept box; here one CTA per sample does the
// whole thing on-chip: decode -> 64-bit key (class | descending confidence | anchor) -> bitonic sort in shared memory ->
// greedy suppression with the whole CTA computing IoUs against each kept box -> two block scans for the row order.
// Integer/compare-bound, a few hundred microseconds per sample at GEN1 size (13 545 anchors: keys in shared memory);
// larger sets (1Mpx: 170 280) keep the keys in the L2-resident workspace; anchors <= 262 143.
#include <mutex>

#include "common.cuh"

namespace soda {

constexpr int DET_THREADS = 1024;
constexpr int DET_MAX_A = 262143;        // anchor index field of the sort key: 18 bits (1Mpx: 170 280 anchors)
constexpr int DET_SMEM_N = 16384;        // up to this padded size the sort keys / flags live in shared memory
constexpr int DET_MAX_CLASSES = 64;

struct DetP {
  int B, A, C, n_pad;
  float nms_thr, pos_thr;
};

__device__ __forceinline__ float det_iou(const float4& a, const float4& b) {
  const float area_a = __fmul_rn(__fsub_rn(a.z, a.x), __fsub_rn(a.w, a.y));
  const float area_b = __fmul_rn(__fsub_rn(b.z, b.x), __fsub_rn(b.w, b.y));
  const float lx = fmaxf(a.x, b.x), ly = fmaxf(a.y, b.y);
  const float hx = fminf(a.z, b.z), hy = fminf(a.w, b.w);
  const float iw = fmaxf(__fsub_rn(hx, lx), 0.0f), ih = fmaxf(__fsub_rn(hy, ly), 0.0f);
  const float inter = __fmul_rn(iw, ih);
  return __fdiv_rn(inter, __fsub_rn(__fadd_rn(area_a, area_b), inter));
}

// exclusive block scan of one int per thread (1024 threads); returns the exclusive prefix, *total = block sum
__device__ __forceinline__ int det_block_scan(int v, int* warp_sums, int* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  if (lane == 31) warp_sums[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int w = warp_sums[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += n;
    }
    warp_sums[lane] = w;   // inclusive over warps
  }
  __syncthreads();
  const int base = warp == 0 ? 0 : warp_sums[warp - 1];
  *total = warp_sums[31];
  __syncthreads();
  return base + inc - v;
}

// per-sample global workspace: conf[A] f32 | cls[A] i32 | box[A] float4 (by anchor) | sbox[n_pad] float4 (sorted)
__host__ __device__ inline int64_t det_ws_per_sample(int A, int n_pad) {
  const int64_t Aa = (A + 3) & ~3;   // keeps the float4 arrays 16-byte aligned
  // + keys[n_pad] u64 | alive[n_pad] u8 | kept[Aa] u8 for anchor sets whose keys do not fit in shared memory
  const int64_t big = n_pad > DET_SMEM_N ? (int64_t)n_pad * 9 + ((Aa + 15) & ~(int64_t)15) : 0;
  return Aa * 8 + Aa * 16 + (int64_t)n_pad * 16 + big;
}

__global__ void __launch_bounds__(DET_THREADS, 1)
multibox_kernel(DetP p, const float* __restrict__ cls_probs, const float* __restrict__ offsets,
                const float* __restrict__ anchors, uint8_t* __restrict__ ws, float* __restrict__ out) {
  extern __shared__ __align__(16) uint8_t det_smem[];
  // keys [n_pad] (class | ~conf | anchor) sorted ascending; alive [n_pad] by sorted position: 1 candidate, 2 kept, 0 out;
  // kept_a [A] by anchor.  In shared memory up to 16384 (padded) anchors, else in the (L2-resident) workspace.
  uint8_t* kbase = det_smem;
  if (p.n_pad > DET_SMEM_N) {
    const int64_t Aa0 = (p.A + 3) & ~3;
    kbase = ws + (int64_t)blockIdx.x * det_ws_per_sample(p.A, p.n_pad) + Aa0 * 8 + Aa0 * 16 + (int64_t)p.n_pad * 16;
  }
  uint64_t* keys = reinterpret_cast<uint64_t*>(kbase);
  uint8_t* alive = kbase + (size_t)p.n_pad * 8;
  uint8_t* kept_a = alive + p.n_pad;
  __shared__ int warp_sums[32];
  __shared__ int s_next;
  __shared__ int seg_start[DET_MAX_CLASSES + 1];

  const int tid = threadIdx.x, b = blockIdx.x, A = p.A, C = p.C, n = p.n_pad;
  uint8_t* w = ws + (int64_t)b * det_ws_per_sample(A, n);
  float* conf = reinterpret_cast<float*>(w);
  const int64_t Aa = (A + 3) & ~3;
  int* cls = reinterpret_cast<int*>(w + Aa * 4);
  float4* box = reinterpret_cast<float4*>(w + Aa * 8);
  float4* sbox = box + Aa;
  const float* probs = cls_probs + (int64_t)b * A * C;
  const float4* off4 = reinterpret_cast<const float4*>(offsets) + (int64_t)b * A;
  const float4* anc4 = reinterpret_cast<const float4*>(anchors);

  // ---- 1. decode (utils/box.py:72-91, :136-139) -------------------------------------------------------------
  for (int a = tid; a < n; a += DET_THREADS) {
    uint64_t key = ~0ull;
    if (a < A) {
      const float* pr = probs + (int64_t)a * C;
      float best = pr[0];
      int arg = 0;
      for (int c = 1; c < C; ++c) {
        const float v = pr[c];
        if (v > best) best = v, arg = c;   // first maximum wins, as torch.max
      }
      const int cid = arg - 1;
      const float4 an = anc4[a], of = off4[a];
      const float acx = __fdiv_rn(__fadd_rn(an.x, an.z), 2.0f), acy = __fdiv_rn(__fadd_rn(an.y, an.w), 2.0f);
      const float aw = __fsub_rn(an.z, an.x), ah = __fsub_rn(an.w, an.y);
      const float cx = __fadd_rn(__fdiv_rn(__fmul_rn(of.x, aw), 10.0f), acx);
      const float cy = __fadd_rn(__fdiv_rn(__fmul_rn(of.y, ah), 10.0f), acy);
      const float bw = __fmul_rn(expf(__fdiv_rn(of.z, 5.0f)), aw);
      const float bh = __fmul_rn(expf(__fdiv_rn(of.w, 5.0f)), ah);
      const float4 bx = make_float4(__fsub_rn(cx, __fmul_rn(0.5f, bw)), __fsub_rn(cy, __fmul_rn(0.5f, bh)),
                                    __fadd_rn(cx, __fmul_rn(0.5f, bw)), __fadd_rn(cy, __fmul_rn(0.5f, bh)));
      conf[a] = best, cls[a] = cid, box[a] = bx;
      kept_a[a] = 0;
      const uint64_t order = cid < 0 ? 0x3FFFull : (uint64_t)cid;
      key = (order << 50) | ((uint64_t)(0xFFFFFFFFu - __float_as_uint(fmaxf(best, 0.0f))) << 18) | (uint64_t)a;
    }
    keys[a] = key;
  }
  __syncthreads();

  // ---- 2. bitonic sort of the keys in shared memory (class ascending, confidence descending) ----------------
  for (int k = 2; k <= n; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < n; i += DET_THREADS) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const uint64_t x = keys[i], y = keys[ixj];
          const bool up = (i & k) == 0;
          if ((x > y) == up) keys[i] = y, keys[ixj] = x;
        }
      }
      __syncthreads();
    }
  }

  // ---- 3. sorted boxes, candidate flags, class segments -----------------------------------------------------
  for (int pos = tid; pos < n; pos += DET_THREADS) {
    const uint64_t key = keys[pos];
    const bool real = key != ~0ull;
    if (real) sbox[pos] = box[(int)(key & 0x3FFFFull)];
    alive[pos] = (real && (key >> 50) != 0x3FFFull) ? 1 : 0;
  }
  if (tid <= C - 1 && tid <= DET_MAX_CLASSES) {
    // lower bound of class `tid` (tid = C - 1 gives the end of the last foreground class)
    int lo = 0, hi = n;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if ((keys[mid] >> 50) < (uint64_t)tid) lo = mid + 1; else hi = mid;
    }
    seg_start[tid] = lo;
  }
  __syncthreads();

  // ---- 4. greedy NMS per foreground class (utils/box.py:94-119) ---------------------------------------------
  for (int c = 0; c < C - 1; ++c) {
    const int hi = seg_start[c + 1];
    int i = seg_start[c];
    while (true) {
      if (tid == 0) s_next = 0x7FFFFFFF;
      __syncthreads();
      for (int pos = i + tid; pos < hi; pos += DET_THREADS) {
        if (alive[pos] == 1) {
          atomicMin(&s_next, pos);
          break;
        }
      }
      __syncthreads();
      const int pn = s_next;
      if (pn == 0x7FFFFFFF) break;
      const float4 top = sbox[pn];
      if (tid == 0) {
        alive[pn] = 2;
        kept_a[(int)(keys[pn] & 0x3FFFFull)] = 1;
      }
      for (int pos = pn + 1 + tid; pos < hi; pos += DET_THREADS) {
        if (alive[pos] == 1) {
          const float iou = det_iou(top, sbox[pos]);
          if (!(iou <= p.nms_thr)) alive[pos] = 0;   // survivors: iou <= threshold (NaN drops, as in torch)
        }
      }
      __syncthreads();
      i = pn + 1;
    }
    __syncthreads();
  }

  // ---- 5. rows: kept boxes first in (class, confidence) order, then the rest by anchor index ------------------
  float* o = out + (int64_t)b * A * 6;
  auto write_row = [&](int row, int a, bool kept) {
    float cf = conf[a];
    float ci = kept ? (float)cls[a] : -1.0f;
    if (cf < p.pos_thr) ci = -1.0f, cf = __fsub_rn(1.0f, cf);
    const float4 bx = box[a];
    float* r = o + (int64_t)row * 6;
    r[0] = ci, r[1] = cf, r[2] = bx.x, r[3] = bx.y, r[4] = bx.z, r[5] = bx.w;
  };
  int n_keep = 0;
  for (int base = 0; base < n; base += DET_THREADS) {
    const int pos = base + tid;
    const int flag = (pos < n && alive[pos] == 2) ? 1 : 0;
    int total;
    const int ex = det_block_scan(flag, warp_sums, &total);
    if (flag) write_row(n_keep + ex, (int)(keys[pos] & 0x3FFFFull), true);
    n_keep += total;
  }
  int n_rest = 0;
  for (int base = 0; base < A; base += DET_THREADS) {
    const int a = base + tid;
    const int flag = (a < A && kept_a[a] == 0) ? 1 : 0;
    int total;
    const int ex = det_block_scan(flag, warp_sums, &total);
    if (flag) write_row(n_keep + n_rest + ex, a, false);
    n_rest += total;
  }
}

}  // namespace soda

extern "C" int64_t soda_multibox_workspace_bytes(int B, int A) {
  if (B <= 0 || A <= 0) return 0;
  int n_pad = 1024;
  while (n_pad < A) n_pad <<= 1;
  return (int64_t)B * soda::det_ws_per_sample(A, n_pad);
}

extern "C" int soda_multibox_detection(const float* cls_probs, const float* offsets, const float* anchors, int B, int A,
                                       int C, float nms_threshold, float pos_threshold, void* workspace, float* out,
                                       void* stream) {
  using namespace soda;
  if (!cls_probs || !offsets || !anchors || !workspace || !out) return fail(SODA_ERR_BAD_ARG, "multibox: null pointer");
  if (B <= 0 || A <= 0 || C < 2) return fail(SODA_ERR_BAD_ARG, "multibox: non-positive size");
  if (A > DET_MAX_A || C > DET_MAX_CLASSES)
    return fail(SODA_ERR_UNSUPPORTED, "multibox: at most 262143 anchors and 64 classes in this build");
  if ((reinterpret_cast<uintptr_t>(offsets) & 15) || (reinterpret_cast<uintptr_t>(anchors) & 15) ||
      (reinterpret_cast<uintptr_t>(workspace) & 15))
    return fail(SODA_ERR_BAD_ARG, "multibox: offsets, anchors and workspace must be 16-byte aligned");
  DetP p;
  p.B = B, p.A = A, p.C = C, p.nms_thr = nms_threshold, p.pos_thr = pos_threshold;
  p.n_pad = 1024;
  while (p.n_pad < A) p.n_pad <<= 1;
  const size_t smem = p.n_pad > DET_SMEM_N ? 16 : (size_t)p.n_pad * 9 + (size_t)((A + 15) / 16 * 16);
  {   // the opt-in is per device: set on every launch (cheap), not once per process
    cudaError_t e = cudaFuncSetAttribute(multibox_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (e != cudaSuccess) {
      set_error("multibox: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
      return (int)SODA_ERR_CUDA;
    }
  }
  multibox_kernel<<<B, DET_THREADS, smem, (cudaStream_t)stream>>>(p, cls_probs, offsets, anchors,
                                                                   reinterpret_cast<uint8_t*>(workspace), out);
  return after_launch("multibox_kernel");
}

// ---- soda_multibox_detection_static: the same kernel with its scratch in a file-scope device array ------------------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart
// in the reference.  8 MiB per GPU that loads the library, held until exit: 14 samples at GEN1 size (13 545 anchors,
// 587 296 bytes each).  Every call shares the array: two calls that overlap on the GPU (different streams) race.
// Nothing in it has to start at zero - the kernel writes conf / cls / box / sbox (and keys / alive / kept for large
// anchor sets) before it reads them - so no clearing between calls.
namespace soda {
constexpr int64_t DET_STATIC_WS_BYTES = SODA_MULTIBOX_STATIC_WS_BYTES;
static __device__ __align__(256) uint8_t det_static_ws[DET_STATIC_WS_BYTES];

constexpr int DET_MAX_DEVICES = 64;
static uint8_t* det_static_ptr[DET_MAX_DEVICES];   // device address of det_static_ws on each device, looked up once
static std::mutex det_static_mu;

// The array's NAME in host code is a host-side stand-in, not a device address: ask the runtime, with the device current.
static int det_static_address(uint8_t** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= DET_MAX_DEVICES) {
    set_error("multibox_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return (int)SODA_ERR_CUDA;
  }
  std::lock_guard<std::mutex> g(det_static_mu);
  if (det_static_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, det_static_ws);
    if (e != cudaSuccess || ptr == nullptr) {
      set_error("multibox_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return (int)SODA_ERR_CUDA;
    }
    det_static_ptr[dev] = reinterpret_cast<uint8_t*>(ptr);
  }
  *out = det_static_ptr[dev];
  return 0;
}
}  // namespace soda

extern "C" int64_t soda_multibox_static_max_bytes(void) { return soda::DET_STATIC_WS_BYTES; }

extern "C" int soda_multibox_detection_static(const float* cls_probs, const float* offsets, const float* anchors, int B,
                                              int A, int C, float nms_threshold, float pos_threshold, float* out,
                                              void* stream) {
  using namespace soda;
  if (B <= 0 || A <= 0) return fail(SODA_ERR_BAD_ARG, "multibox_static: non-positive size");
  const int64_t need = soda_multibox_workspace_bytes(B, A);
  if (need > DET_STATIC_WS_BYTES) {
    set_error("multibox_static: B = %d samples of A = %d anchors need %lld scratch bytes, the built-in array holds %lld; "
              "split the batch or use soda_multibox_detection with a caller workspace",
              B, A, (long long)need, (long long)DET_STATIC_WS_BYTES);
    return (int)SODA_ERR_UNSUPPORTED;
  }
  uint8_t* ws = nullptr;
  const int rc = det_static_address(&ws);
  if (rc != 0) return rc;
  return soda_multibox_detection(cls_probs, offsets, anchors, B, A, C, nms_threshold, pos_threshold, ws, out, stream);
}
