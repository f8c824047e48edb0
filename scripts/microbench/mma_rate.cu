// Microbenchmark: issue rate of tcgen05.mma (cta_group::1, M=128) on B200 as a function of N, operand
// source (A in TMEM vs shared memory), kind (tf32 K=8 / bf16 K=16) and accumulator interleaving.
// One CTA per SM; one thread issues `reps` MMAs back to back, commits, waits; cycles by clock64.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu && ./mma_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../snn_for_object_detection_b200/csrc/tc_common.cuh"
using namespace soda::tc;

// MODE: 0 tf32 TS, 1 tf32 SS, 2 bf16 TS, 3 bf16 SS, 4 alternating groups of 4 tf32 SS / 4 bf16 SS (the conv kernel's
// two products of a k-block issued by ONE thread) ; NACC = accumulators cycled through (MODE 4: NACC = 1 both kinds on
// one accumulator, NACC = 2 one accumulator per kind)
template <int N, int MODE, int NACC>
__global__ void __launch_bounds__(128, 1) k(int reps, long long* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) tmem_alloc(&slot, 512);
  for (int i = threadIdx.x; i < 80 * 1024 / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0;
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tb = slot;
  if (threadIdx.x < 32) {
    constexpr uint32_t idesc = (MODE == 2 || MODE == 3) ? idesc_bf16(128, N) : idesc_tf32(128, N);
    constexpr uint32_t idesc_b = idesc_bf16(128, N);
    const uint64_t bdesc = smem_desc_sw128(smem_u32(smem));
    const uint64_t adesc = smem_desc_sw128(smem_u32(smem + 48 * 1024));
    long long t0 = 0, t1 = 0;
    for (int rep = 0; rep < 2; ++rep) {   // first pass warms up
      t0 = clock64();
      if (elect_one()) {
        for (int i = 0; i < reps; i += 8) {
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const uint32_t d = tb + (uint32_t)(u % NACC) * N;
            const uint32_t a_t = tb + 256 + (u & 3) * 64 + (u & 1) * 8;
            const uint64_t b = bdesc + 2 * (u & 3);
            if (MODE == 4) {
              const uint32_t dk = tb + (uint32_t)((u >> 2) % NACC) * N;
              if (u < 4) mma_tf32_ss(dk, adesc + 2 * (u & 3), b, idesc, 1);
              else mma_bf16_ss(dk, adesc + 2 * (u & 3), b, idesc_b, 1);
            } else if (MODE == 0) mma_tf32_ts(d, a_t, b, idesc, 1);
            else if (MODE == 1) mma_tf32_ss(d, adesc + 2 * (u & 3), b, idesc, 1);
            else if (MODE == 2) mma_bf16_ts(d, a_t, b, idesc, 1);
            else mma_bf16_ss(d, adesc + 2 * (u & 3), b, idesc, 1);
          }
        }
        mma_commit(&bar);
      }
      __syncwarp();
      mbar_wait(&bar, rep & 1);
      t1 = clock64();
    }
    if (threadIdx.x == 0 && blockIdx.x == 0) *out = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tb, 512); }
}

template <int N, int MODE, int NACC>
void run(long long* d) {
  const char* names[5] = {"tf32 TS", "tf32 SS", "bf16 TS", "bf16 SS", "tf32+bf16 SS"};
  const int reps = 1024;
  cudaFuncSetAttribute(k<N, MODE, NACC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  k<N, MODE, NACC><<<148, 128, 100 * 1024>>>(reps, d);
  long long c = 0; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
  cudaError_t e = cudaGetLastError();
  printf("%s N=%3d nacc=%d : %7.1f cycles/MMA  (floor N/2 = %d)%s\n", names[MODE], N, NACC, (double)c / reps, N / 2,
         e == cudaSuccess ? "" : cudaGetErrorString(e));
}
template <int MODE>
void run_mode(long long* d) {
  run<32, MODE, 1>(d); run<32, MODE, 2>(d); run<32, MODE, 4>(d);
  run<64, MODE, 1>(d); run<64, MODE, 2>(d); run<64, MODE, 4>(d);
  run<128, MODE, 1>(d); run<128, MODE, 2>(d);
  run<256, MODE, 1>(d);
}

int main() {
  long long* d; cudaMalloc(&d, 8);
  run_mode<0>(d); run_mode<1>(d); run_mode<2>(d); run_mode<3>(d);
  run<32, 4, 1>(d); run<32, 4, 2>(d); run<64, 4, 1>(d); run<64, 4, 2>(d); run<128, 4, 1>(d); run<128, 4, 2>(d);
  return 0;
}
