// Does a SWIZZLE_128B K-major tcgen05 shared-memory descriptor accept a start address that is shifted by
// whole 128-byte rows (not 1024-byte aligned), and what must the base-offset field [49,52) hold?
// Image: 18 x 16 "pixels" of 32 fp32 channels, 128 B per pixel, stored like TMA SWIZZLE_128B would store it
// (16-byte chunk c of pixel P at chunk position c ^ (P & 7), base 1024-aligned).  A operand for tap shift
// (sy, sx): row m = 8*g + i  <->  pixel (g + sy) * 16 + i + sx ; SBO = 2048 B.  B = [I | 0] so that
// D[m][n] = A[m][kk*8 + n] for n < 8.  Prints, per variant, how many of the 128x8 outputs match.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../snn_for_object_detection_b200/csrc/tc_common.cuh"
using namespace soda::tc;
#ifndef PITCH
#define PITCH 16
#endif

__device__ __forceinline__ float val(int P, int ch) { return (float)(P * 32 + ch); }   // exact in tf32 (< 2^11 * ...)

__global__ void __launch_bounds__(128, 1) k(int sy, int sx, int kk, int base_off_mode, float* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  float* img = (float*)smem;                 // 288 pixels * 128 B = 36 KB
  float* bt = (float*)(smem + 40 * 1024);    // B tile: 16 rows x 128 B, SW128
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) tmem_alloc(&slot, 32);
  for (int i = threadIdx.x; i < 288 * 32; i += blockDim.x) {
    const int P = i / 32, ch = i % 32, c = ch / 4, e = ch % 4;
    // values small enough to be exact in tf32: use (P % 64) * 32 + ch < 2048
    img[P * 32 + ((c ^ (P & 7)) * 4) + e] = (float)((P % 64) * 32 + ch);
  }
  for (int i = threadIdx.x; i < 16 * 32; i += blockDim.x) {
    const int n = i / 32, ch = i % 32, c = ch / 4, e = ch % 4;
    // B[n][k] = 1 if k == kk*8 + n (n < 8)
    bt[n * 32 + ((c ^ (n & 7)) * 4) + e] = (n < 8 && ch == kk * 8 + n) ? 1.0f : 0.0f;
  }
  fence_proxy_async_smem();
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tb = slot;
  if (threadIdx.x == 0) {
    const uint32_t a_addr = smem_u32(smem) + (uint32_t)((sy * PITCH + sx) * 128) + kk * 32;
    uint64_t ad = 0;
    ad |= (uint64_t)((a_addr & 0x3FFFFu) >> 4);
    ad |= (uint64_t)1 << 16;
    ad |= (uint64_t)((PITCH * 128) >> 4) << 32;      // SBO: 8-row groups are 16 pixels apart
    ad |= (uint64_t)1 << 46;
    ad |= (uint64_t)2 << 61;
    if (base_off_mode == 1) ad |= (uint64_t)((a_addr >> 7) & 7) << 49;
    const uint64_t bd = smem_desc_sw128(smem_u32(bt)) + 2 * kk;
    mma_tf32_ss(tb, ad, bd, idesc_tf32(128, 16), 0);
    mma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  uint32_t r[16];
  const int q = threadIdx.x >> 5;
  tmem_ld_32x16(tb + ((uint32_t)(q * 32) << 16), r);
  tmem_ld_wait();
  for (int j = 0; j < 16; ++j) out[threadIdx.x * 16 + j] = __uint_as_float(r[j]);
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tb, 32); }
}

int main() {
  float* d; cudaMalloc(&d, 128 * 16 * 4);
  float h[128 * 16];
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  for (int mode = 0; mode < 2; ++mode)
    for (int sy = 0; sy < 3; sy += 2)
      for (int sx = 0; sx < 3; ++sx)
        for (int kk = 0; kk < 4; kk += 3) {
          k<<<1, 128, 64 * 1024>>>(sy, sx, kk, mode, d);
          cudaError_t e = cudaDeviceSynchronize();
          cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
          int ok = 0;
          for (int m = 0; m < 128; ++m) {
            const int P = (m / 8 + sy) * PITCH + (m % 8) + sx;
            for (int n = 0; n < 8; ++n) ok += h[m * 16 + n] == (float)((P % 64) * 32 + kk * 8 + n);
          }
          printf("base_off_mode=%d sy=%d sx=%d kk=%d : %4d / 1024 match %s  (row0: %g %g, row9: %g)\n", mode, sy, sx, kk, ok,
                 e == cudaSuccess ? "" : cudaGetErrorString(e), h[0], h[1], h[9 * 16]);
        }
  return 0;
}
