// tcgen05 / TMEM / TMA implicit-GEMM convolution (soda_conv2d_fwd kind SODA_CONV_TC).
// Placeholder until the kernel lands: reports "unsupported" loudly, never falls back.
#include "common.cuh"

namespace soda {

int64_t conv_tc_packed_elems(int, int, int) { return 0; }

int conv_tc_pack(const float*, const float*, int, int, int, float*, cudaStream_t) {
  return fail(SODA_ERR_UNSUPPORTED, "conv_tc: tcgen05 kernel not built in this revision");
}

int conv_tc_dispatch(const soda_conv_desc_t*, const float*, const float*, const float*,
                     const float*, float*, void*) {
  return fail(SODA_ERR_UNSUPPORTED, "conv_tc: tcgen05 kernel not built in this revision");
}

}  // namespace soda
