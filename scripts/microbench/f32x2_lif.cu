// Is the packed-fp32 LIF step (FADD2 / FMUL2 / FFMA2) bit-identical to the scalar one?  nvcc -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../snn_for_object_detection_b200/csrc/neuron.cuh"
using namespace soda;
__global__ void k(NP p, const float* x, int T, int n, float* out, unsigned* bad) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n / 2) return;
  float v0 = p.v_leak, v1 = p.v_leak, c0 = 0, c1 = 0;
  f32x2_t v = pk2(p.v_leak, p.v_leak), c = pk2(0.f, 0.f);
  NP2 q = np_pairs(p);
  for (int t = 0; t < T; ++t) {
    float a = x[(size_t)t * n + 2 * i], b = x[(size_t)t * n + 2 * i + 1];
    float z0, z1, vd, y0, y1;
    lif_step(p, a, v0, c0, z0, vd);
    lif_step(p, b, v1, c1, z1, vd);
    lif_step2(q, pk2(a, b), v, c, y0, y1);
    float pv0, pv1, pc0, pc1;
    upk2(v, pv0, pv1), upk2(c, pc0, pc1);
    if (__float_as_uint(pv0) != __float_as_uint(v0)) atomicAdd(bad + 0, 1);
    if (__float_as_uint(pv1) != __float_as_uint(v1)) atomicAdd(bad + 0, 1);
    if (__float_as_uint(pc0) != __float_as_uint(c0)) atomicAdd(bad + 1, 1);
    if (__float_as_uint(pc1) != __float_as_uint(c1)) atomicAdd(bad + 1, 1);
    if (y0 != z0 || y1 != z1) atomicAdd(bad + 2, 1);
    if (__float_as_uint(pc0) != __float_as_uint(c0) && atomicAdd(bad + 3, 1) < 4)
      printf("t=%d x=%a i scalar %a packed %a\n", t, a, c0, pc0);
  }
  out[i] = v0 + c0;
}
int main() {
  const int n = 1 << 20, T = 32;
  float* x; float* out; unsigned* bad;
  cudaMallocManaged(&x, sizeof(float) * n * T); cudaMallocManaged(&out, sizeof(float) * n); cudaMallocManaged(&bad, 16);
  uint32_t s = 12345;
  for (size_t i = 0; i < (size_t)n * T; ++i) { s = s * 1664525u + 1013904223u; x[i] = ((int)(s >> 8) - (1 << 23)) * (3.0f / (1 << 23)); if ((s & 255) == 0) x[i] *= 1e-38f; }
  for (int j = 0; j < 4; ++j) bad[j] = 0;
  NP p{0.1f, 0.2f, 0.f, 1.f, 0.f, 100.f, 1, 0};
  k<<<n / 2 / 256, 256>>>(p, x, T, n, out, bad);
  printf("sync %d\n", (int)cudaDeviceSynchronize());
  printf("mismatches: v %u  i %u  z %u\n", bad[0], bad[1], bad[2]);
  return 0;
}
