// Shared helpers for the libsoda_b200 kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>

#include "../../include/soda_b200.h"

namespace soda {

// thread-local error text behind soda_last_error()
void set_error(const char* fmt, ...);
extern std::atomic<int64_t> g_launches;

inline int fail(soda_status_t code, const char* what) {
  set_error("%s", what);
  return (int)code;
}

// Called after every kernel launch: counts it and converts a launch error into a status.
inline int after_launch(const char* name) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("%s: %s (%d)", name, cudaGetErrorString(e), (int)e);
    return (int)SODA_ERR_CUDA;
  }
  return 0;
}

constexpr int kNumSMs = 148;  // B200

// Programmatic dependent launch (PDL).  pdl_trigger(): once every CTA of this grid has executed it (or exited), a
// dependent grid launched with cudaLaunchAttributeProgrammaticStreamSerialization may start.  pdl_wait(): in such a
// dependent grid, blocks until the preceding grid has completed and its memory is visible; no-op otherwise.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

__device__ __forceinline__ bool aligned16(const void* p) {
  return (reinterpret_cast<uintptr_t>(p) & 15) == 0;
}

}  // namespace soda
