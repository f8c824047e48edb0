// soda_conv_pack_weights / soda_conv2d_fwd: argument checking and dispatch between the FP32-pipe
// kernel (conv_ffma.cu), the tcgen05 kernel (conv_tc3.cu) and the direct event-frame kernel (conv_direct_lif.cu).
#include "common.cuh"

namespace soda {
int conv_ffma_pack(const float* w, const float* scale, int Cout, int Cin, int k, int cout_pad,
                   float* packed, cudaStream_t st);
int conv_ffma_dispatch(const soda_conv_desc_t* d, const float* x, const float* wp, int cout_pad,
                       const float* shift, const float* residual, float* y, void* stream);
struct NPfwd;
struct Tc3Lif;
int64_t conv_tc3_packed_elems(int Cout, int Cin, int k);
int conv_tc3_pack(const float* w, const float* scale, int Cout, int Cin, int k, float* packed,
                  cudaStream_t st);
int conv_tc3_dispatch(const soda_conv_desc_t* d, const float* x, const float* packed,
                      const float* shift, const float* residual, float* y, const Tc3Lif* lif,
                      void* stream);
int conv_tc3_lif_dispatch(const soda_conv_desc_t* d, const float* x, const float* packed,
                          const float* shift, const float* residual, int T,
                          const soda_neuron_params_t* np, const float* v0, const float* i0, float* y,
                          float* v_out, float* i_out, void* stream);
bool conv_tc3_supported(const soda_conv_desc_t* d, const float* x);
int conv_direct_lif_dispatch(const soda_conv_desc_t* d, const float* x, const float* wp, int cout_pad,
                             const float* shift, int T, const soda_neuron_params_t* np, const float* v0,
                             const float* i0, float* y, float* v_out, float* i_out, void* stream);
static inline int cout_pad_of(int Cout) { return (Cout + 31) / 32 * 32; }
}  // namespace soda

extern "C" int64_t soda_conv_packed_elems(int Cout, int Cin, int k, int kind) {
  using namespace soda;
  if (Cout <= 0 || Cin <= 0 || k <= 0) return 0;
  if (kind == SODA_CONV_TC) return conv_tc3_packed_elems(Cout, Cin, k);
  return (int64_t)k * k * Cin * cout_pad_of(Cout);
}

extern "C" int soda_conv_pack_weights(const float* weight, const float* scale, int Cout, int Cin,
                                      int k, int kind, float* packed, void* stream) {
  using namespace soda;
  if (!weight || !packed) return fail(SODA_ERR_BAD_ARG, "conv_pack: null pointer");
  if (Cout <= 0 || Cin <= 0 || k <= 0) return fail(SODA_ERR_BAD_ARG, "conv_pack: non-positive size");
  cudaStream_t st = (cudaStream_t)stream;
  if (kind == SODA_CONV_TC) return conv_tc3_pack(weight, scale, Cout, Cin, k, packed, st);
  if (kind != SODA_CONV_FP32) return fail(SODA_ERR_UNSUPPORTED, "conv_pack: unknown kind");
  return conv_ffma_pack(weight, scale, Cout, Cin, k, cout_pad_of(Cout), packed, st);
}

extern "C" int soda_conv2d_fwd(const soda_conv_desc_t* d, const float* x, const float* packed,
                               const float* shift, const float* residual, float* y, void* stream) {
  using namespace soda;
  if (!d || !x || !packed || !y) return fail(SODA_ERR_BAD_ARG, "conv2d: null pointer");
  if (d->N <= 0 || d->H <= 0 || d->W <= 0 || d->Cin <= 0 || d->Cout <= 0 || d->k <= 0 || d->stride <= 0)
    return fail(SODA_ERR_BAD_ARG, "conv2d: non-positive size");
  if (d->kind == SODA_CONV_TC) return conv_tc3_dispatch(d, x, packed, shift, residual, y, nullptr, stream);
  if (d->kind != SODA_CONV_FP32) return fail(SODA_ERR_UNSUPPORTED, "conv2d: unknown kind");
  return conv_ffma_dispatch(d, x, packed, cout_pad_of(d->Cout), shift, residual, y, stream);
}

extern "C" int soda_conv_tc_supported(const soda_conv_desc_t* d, const float* x) {
  if (!d || !x) return 0;
  return soda::conv_tc3_supported(d, x) ? 1 : 0;
}

extern "C" int soda_conv_lif_fwd(const soda_conv_desc_t* d, const float* x, const float* packed,
                                 const float* shift, const float* residual, int T,
                                 const soda_neuron_params_t* np, const float* v0, const float* i0,
                                 float* y, float* v_out, float* i_out, void* stream) {
  using namespace soda;
  if (!d || !x || !packed || !y || !np || !v_out || !i_out)
    return fail(SODA_ERR_BAD_ARG, "conv_lif: null pointer");
  if (d->N <= 0 || d->H <= 0 || d->W <= 0 || d->Cin <= 0 || d->Cout <= 0 || d->k <= 0 || d->stride <= 0)
    return fail(SODA_ERR_BAD_ARG, "conv_lif: non-positive size");
  if (d->kind == SODA_CONV_FP32) {
    if (residual != nullptr) return fail(SODA_ERR_UNSUPPORTED, "conv_lif(fp32): no residual input");
    return conv_direct_lif_dispatch(d, x, packed, cout_pad_of(d->Cout), shift, T, np, v0, i0, y, v_out, i_out, stream);
  }
  if (d->kind != SODA_CONV_TC) return fail(SODA_ERR_UNSUPPORTED, "conv_lif: unknown kind");
  return conv_tc3_lif_dispatch(d, x, packed, shift, residual, T, np, v0, i0, y, v_out, i_out, stream);
}
