/*
 * soda_b200.h - C ABI of the B200-native SODa hot path (libsoda_b200.so).
 *
 * The reference (KirillHit/snn_for_object_detection) has NO native/FFI boundary: its operator
 * interface is a Python protocol.  The entry points below are therefore the
 * device-side halves of the reference *call sites* they replace; each cites the reference
 * file:line whose arithmetic it reproduces.  The Python host side
 * (snn_for_object_detection_b200/) keeps the reference's LayerGen / BlockGen / SODa API on top.
 *
 * Conventions (all entry points):
 *   - plain C types only: device pointers, int/int64_t sizes, float parameters, a CUDA stream
 *     passed as void* (cudaStream_t);
 *   - returns 0 on success, a soda_status_t > 0 otherwise; soda_last_error() gives the
 *     thread-local message.  Never throws, never allocates device memory, never synchronises:
 *     work is enqueued on the caller's stream and every buffer is caller-owned;
 *   - activations are fp32, "pixel-major" (N, H, W, C) = torch channels_last, addressed as
 *         elem(n, y, x, c) = base[n * batch_stride + (y * W + x) * pix_stride + c]
 *     (strides in elements) so a layer can read or write a channel slice of a concat buffer;
 *   - there is no CPU implementation behind any of them.
 */
#ifndef SODA_B200_H
#define SODA_B200_H

#include <stdint.h>

#if defined(__GNUC__)
#define SODA_API __attribute__((visibility("default")))
#else
#define SODA_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  SODA_OK = 0,
  SODA_ERR_BAD_ARG = 1,      /* null pointer / non-positive size / misaligned buffer */
  SODA_ERR_UNSUPPORTED = 2,  /* shape or flag combination this build has no kernel for */
  SODA_ERR_CUDA = 3          /* a CUDA runtime/driver call failed (message has the code) */
} soda_status_t;

#define SODA_ABI_VERSION 2

struct soda_neuron_params_s; /* defined below (LIF / LI section) */

SODA_API int soda_abi_version(void);
SODA_API const char* soda_last_error(void);
/* Number of kernels launched by this library in this process (all threads). */
SODA_API int64_t soda_launch_count(void);

/* ---------------------------------------------------------------------------------------------
 * Event decode + occupancy binning  (K7)
 *   replaces: prophesee_toolbox/src/io/dat_events_tools.py:100-108 (Event2D unpack:
 *             x = w & 0x3FFF, y = (w >> 14) & 0x3FFF, p = (w >> 28) & 1) followed by
 *             utils/datasets.py:420-433 (ST) / :326-336 (MT):  ts = (t - t0) / step_us ;
 *             features[ts, p, y, x] = 1   (binary occupancy, duplicates idempotent).
 *   events        : device, n_total little-endian 8-byte records {u32 t_us, i32 w}, samples
 *                   concatenated; sample b owns records [sample_offsets[b], sample_offsets[b+1])
 *   sample_offsets: device int64[B+1]
 *   t0_us         : device int64[B]; events with t < t0 are dropped (datasets.py:416); events
 *                   whose bin is >= T are dropped (the reference never produces them: its loader
 *                   slices by time first, psee_loader.py:123-165)
 *   clip_x        : 1 = clamp x to W-1 as datasets.py:426 (ST); 0 = MT behaviour.  Events with
 *                   y >= H (or x >= W when !clip_x) raise IndexError in the reference; here they
 *                   are dropped and counted in *n_dropped (device int32, may be NULL).
 *   frames        : device fp32, element (t, b, p, y, x) at
 *                   frames[t*stride_t + b*stride_b + (p*H + y)*W + x]; every element of the
 *                   [T,B,2,H,W] block is written (0.0f or 1.0f) - no pre-zeroing needed.
 *   sorted_by_time: 1 = records of each sample are non-decreasing in t (the .dat contract that
 *                   psee_loader.py:157 itself relies on) -> single-pass shared-memory bitmap
 *                   kernel; 0 = any order -> clear + scatter kernels.
 * ------------------------------------------------------------------------------------------- */
SODA_API int soda_bin_events(const void* events, const int64_t* sample_offsets, const int64_t* t0_us,
                    int B, int64_t step_us, int T, int H, int W, int clip_x, int sorted_by_time,
                    float* frames, int64_t stride_t, int64_t stride_b, int* n_dropped,
                    void* stream);

/* ---------------------------------------------------------------------------------------------
 * Convolution (K1) with the per-channel affine of inference BatchNorm (K2), residual add and
 * concat-slice store (K6) in the epilogue.
 *   replaces: models/modules/layer_gen.py:129-136  nn.Conv2d(Cin, Cout, k, padding=k/2,
 *             stride, bias=False); :211-214 nn.BatchNorm2d in eval mode;
 *             models/generator.py:146 (Residual = sum of branches) and :158 (Dense = cat(dim=1)).
 *   y[n,oy,ox,co] = scale[co] * sum_{ky,kx,ci} x[n, oy*s+ky-k/2, ox*s+kx-k/2, ci] * w[co,ci,ky,kx]
 *                   + shift[co] (+ residual[n,oy,ox,co])
 *
 * Two steps, so that the weight re-layout and the BatchNorm fold are paid once per weight
 * update and not once per call:
 *   soda_conv_pack_weights : weight in the reference's own layout [Cout, Cin, k, k] (the
 *       state_dict tensor) * scale[Cout] (NULL = 1) -> an opaque packed buffer of
 *       soda_conv_packed_elems() floats owned by the caller.
 *   soda_conv2d_fwd        : consumes the packed buffer.
 *   kind: SODA_CONV_FP32 = FP32-pipe implicit GEMM, fp32-exact products (any shape);
 *         SODA_CONV_TC   = TMA-fed tcgen05/TMEM implicit GEMM, tf32 main pass + bf16 correction
 *                          pass on split operands (fp32-level accuracy); needs Cin % 32 == 0 and
 *                          pixel-major input.
 *   shift, residual : may be NULL.
 *   in_nchw: 1 = input is plain NCHW-contiguous [N,Cin,H,W] (the event frames); strides ignored
 * ------------------------------------------------------------------------------------------- */
#define SODA_CONV_FP32 0
#define SODA_CONV_TC 1
/* flags: the input tensor holds only values exactly representable in TF32 (spikes {0,1}, pooled
 * spike counts, event frames): the tensor-core kernel then skips the A-operand split pass. */
#define SODA_CONV_FLAG_INPUT_EXACT_TF32 1
/* flags: every input value has at most 8 significant bits (spikes {0,1}, sums / 2x2 sum pools of a few spike
 * maps: integers <= 256), i.e. is exact in bf16.  The tensor-core kernel then drops the tf32 pass altogether:
 * x (bf16, exact) times [bf16 rn(w) | bf16 rn(w - rn(w))] in two kind::f16 products with fp32 accumulation -
 * 16+ significant weight bits, half the tensor-core work and half the weight bytes of the tf32 + correction
 * scheme.  Implies SODA_CONV_FLAG_INPUT_EXACT_TF32. */
#define SODA_CONV_FLAG_INPUT_EXACT_BF16 2
/* flags (soda_conv_lif_fwd, direct event-frame kernel only): write the spikes as SODA_ACT_HALF / SODA_ACT_PAIR
 * (see "packed activation formats" below) instead of fp32; `y` and the out strides are then those of that format. */
#define SODA_CONV_FLAG_OUT_HALF 4
#define SODA_CONV_FLAG_OUT_PAIR 8

typedef struct {
  int N, H, W, Cin, Cout, k, stride;
  int in_nchw;
  int kind;
  int flags;
  int64_t in_batch_stride, in_pix_stride;
  int64_t out_batch_stride, out_pix_stride;
  int64_t res_batch_stride, res_pix_stride;
} soda_conv_desc_t;

SODA_API int64_t soda_conv_packed_elems(int Cout, int Cin, int k, int kind);
SODA_API int soda_conv_pack_weights(const float* weight, const float* scale, int Cout, int Cin, int k,
                           int kind, float* packed, void* stream);
/* 1 if the SODA_CONV_TC kernel takes this descriptor / input pointer, else 0 (use SODA_CONV_FP32). */
SODA_API int soda_conv_tc_supported(const soda_conv_desc_t* d, const float* x);
SODA_API int soda_conv2d_fwd(const soda_conv_desc_t* d, const float* x, const float* packed,
                    const float* shift, const float* residual, float* y, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Fused Conv -> BatchNorm(eval) -> LIF over T time steps (K1 + K2 + K3 + K6 in one kernel)
 *   replaces: the Conv / Norm / LIF triple emitted by models/tiny_yolo.py:53-58 (_conv) as executed
 *             once per time step by models/generator.py:184-198, including the Residual sum with
 *             the Pass branch of a bottleneck (models/tiny_yolo.py:60-65, generator.py:146).
 *   x        : [T*B, H, W, Cin] pixel-major, time-major frames (frame t*B + b); d->N = T*B
 *   y        : spikes {0,1} (+ residual), same frame order, written through out strides
 *   v0, i0   : [B, Ho, Wo, Cout] dense pixel-major incoming state, or both NULL (v = v_leak, i = 0)
 *   v_out, i_out : state after step T-1, same layout (never NULL)
 *   The work item of the kernel is (sample, pixel tile, channel chunk) and t runs inside it, so the
 *   membrane state lives in registers for all T steps.  LIF arithmetic is op-for-op soda_lif_fwd's.
 *   d->kind == SODA_CONV_TC : tcgen05 kernel, needs Cout % 16 == 0.
 *   d->kind == SODA_CONV_FP32 with in_nchw = 1 (the event frames, Cin = 2): direct FP32-pipe kernel
 *       for tiny K = Cin*k*k (Cin*k*k*Cout <= 11000, Cout % 8 == 0, no residual): one thread owns 8
 *       channels of one pixel for all T steps.
 * ------------------------------------------------------------------------------------------- */
SODA_API int soda_conv_lif_fwd(const soda_conv_desc_t* d, const float* x, const float* packed,
                      const float* shift, const float* residual, int T,
                      const struct soda_neuron_params_s* np, const float* v0, const float* i0,
                      float* y, float* v_out, float* i_out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * LIF / LI scans over T time steps (K3, K4)
 *   replaces: models/modules/layer_gen.py:233 norse.torch.LIFCell() and :253 LICell() (norse
 *             1.1.0, third-party, restated in oracle/norse_restated.py), called once per time
 *             step by models/generator.py:190-193; optional Tanh of layer_gen.py:283-284.
 *   x        : [T, n] input currents, time-major, any element order inside a step
 *   v0, i0   : [n] incoming state or NULL (= initial state v = v_leak, i = 0)
 *   out      : [T, n]  LIF: spikes {0,1};  LI: membrane potential (tanh applied if apply_tanh)
 *   v_out,i_out : [n] state after the last step (never NULL)
 *   v_dec_trace : NULL or [T, n], pre-reset membrane v_decayed (saved for backward)
 *   v_trace,i_trace : NULL or [T, n], per-step new state (what StateStorage records,
 *                     models/modules/common.py:117-122)
 *   Arithmetic is op-for-op that of the restatement (no FMA contraction), so results are
 *   bit-identical to it for identical inputs.  jump_first: 1 = norse 1.1.0 order (default).
 * ------------------------------------------------------------------------------------------- */
typedef struct soda_neuron_params_s {
  float dt_tau_mem; /* fp32(dt) * fp32(tau_mem_inv) = 0.1f  */
  float dt_tau_syn; /* fp32(dt) * fp32(tau_syn_inv) = 0.2f  */
  float v_leak, v_th, v_reset;
  float alpha;      /* SuperSpike steepness (backward only) */
  int jump_first;
  int detach_reset; /* backward only */
} soda_neuron_params_t;

SODA_API int soda_lif_fwd(const soda_neuron_params_t* p, const float* x, const float* v0, const float* i0,
                 int T, int64_t n, float* out, float* v_out, float* i_out, float* v_dec_trace,
                 float* v_trace, float* i_trace, void* stream);

/* LIF scan with the Residual sum and the Dense concat-slice store fused (one pass instead of
 * scan + add_n + slice copy).
 *   replaces: LIFCell (layer_gen.py:233) as the last layer of a Residual branch whose other branch
 *             is Pass (models/tiny_yolo.py:60-65; sum at models/generator.py:146), the result
 *             being one input of a Dense cat (generator.py:158).
 *   x, v0, i0, v_out, i_out : dense pixel-major, x = [T][n_pix][C], state = [n_pix][C]
 *   residual : NULL or element (t, q, c) at residual[(t*n_pix + q)*res_pix_stride + c]
 *   out      : element (t, q, c) at out[(t*n_pix + q)*out_pix_stride + c] = z (+ residual)
 *   Needs C % 4 == 0, strides % 4 == 0, 16-byte aligned pointers. */
SODA_API int soda_lif_res_fwd(const soda_neuron_params_t* p, const float* x, const float* v0, const float* i0,
                     int T, int64_t n_pix, int C, const float* residual, int64_t res_pix_stride,
                     float* out, int64_t out_pix_stride, float* v_out, float* i_out, void* stream);

/* LIF scan with the following 2x2 / stride-2 pooling of the spikes fused into the store (the generator block
 * Conv, Norm, LIF, Pool - reference README.md:57-64, models/modules/layer_gen.py:139-173, common.py:45-49):
 * the full-resolution spike map is never written.
 *   x            : dense pixel-major [T*B, H, W, C] input currents (frame t*B + b)
 *   out          : pooled spikes, element (t, b, oy, ox, c) at
 *                  out[(((t*B + b)*Ho + oy)*Wo + ox)*out_pix_stride + c],  Ho = H/2, Wo = W/2 (floor mode)
 *   v0, i0, v_out, i_out : dense [B, H, W, C] state at FULL resolution (NULL in = initial state)
 *   mode: 0 = sum, 1 = average, 2 = max.  Needs C % 4 == 0. */
SODA_API int soda_lif_pool_fwd(const soda_neuron_params_t* p, const float* x, const float* v0, const float* i0, int T,
                      int B, int H, int W, int C, int mode, float* out, int64_t out_pix_stride, float* v_out,
                      float* i_out, void* stream);

SODA_API int soda_li_fwd(const soda_neuron_params_t* p, const float* x, const float* v0, const float* i0,
                int T, int64_t n, int apply_tanh, float* out, float* v_out, float* i_out,
                float* i_trace, void* stream);

/* Saturable leaky integrator scan (LayerGen SLI): T steps of the reference's own sli_feed_forward_step,
 * models/modules/sli.py:110-126 (cell :80-107, LayerGen models/modules/layer_gen.py:331-347):
 *   i_jump = i + x * sigmoid(v_st - |v|);  v' = v + dt_tau_mem * ((v_leak - v) + i_jump);  i' = i_jump - dt_tau_syn * i_jump
 *   out = v' per step.  p->v_th carries v_st; layouts as soda_li_fwd; NULL v0 / i0 = initial state (v_leak, 0). */
SODA_API int soda_sli_fwd(const soda_neuron_params_t* p, const float* x, const float* v0, const float* i0, int T,
                 int64_t n, float* out, float* v_out, float* i_out, float* i_trace, void* stream);

/* Synaptic transmission scan (LayerGen Synapse): T steps of the reference's own synapse_feed_forward_step,
 * models/modules/synapse.py:73-103 (cell :36-70, LayerGen models/modules/layer_gen.py:321-328):
 *   tau = x > 0 ? tau_secretion : tau_dissociation;  p' = p + ((x - p) * tau) * dt
 *   g = sigma_inhibition != 0 ? 4 * sigma * (p' - sigma * p'^2) : p';  out = max(g, 0) per step
 *   x, out : [T, n];  p0 (NULL = zeros), p_out : [n]. */
SODA_API int soda_synapse_fwd(float tau_secretion, float tau_dissociation, float sigma_inhibition, float dt,
                     const float* x, const float* p0, int T, int64_t n, float* out, float* p_out, void* stream);

/* ConvLSTM gates (LayerGen LSTM): the elementwise part of the reference's models/modules/conv_lstm.py:62-75 after the
 * gate convolution (which runs on soda_conv2d_fwd: the input part W_x * X for all T steps at once, the recurrent part
 * W_h * h per step with the input part as the fused residual):
 *   pre    : [n_pix, 4 * hidden] pixel-major conv output, channel blocks (input | forget | out gate | in node)
 *   c_prev : [n_pix, hidden] or NULL (zeros);  c' = sigmoid(f) * c + sigmoid(i) * tanh(g);  h' = sigmoid(o) * tanh(c') */
SODA_API int soda_lstm_gates(const float* pre, const float* c_prev, int64_t n_pix, int hidden, float* h_out, float* c_out,
                    void* stream);

/* Stateless LayerGens: ReLU / SiLU / Tanh (models/modules/layer_gen.py:255-284; mode 0 / 1 / 2) over a dense tensor of
 * n floats, and Up with mode "nearest" and an integer scale (:176-194) over pixel-major frames
 * x [N, H, W, C] -> y [N, H*scale, W*scale, C]. */
SODA_API int soda_unary_fwd(const float* x, int64_t n, int mode, float* y, void* stream);
SODA_API int soda_upsample_nearest(const float* x, int64_t N, int H, int W, int C, int scale, float* y, void* stream);

/* Surrogate-gradient LIF backward (K8): fused reverse scan over T.
 *   replaces: autograd through T calls of LIFCell incl. norse SuperSpike.backward
 *             (g / (alpha*|v_dec - v_th| + 1)^2), models/soda.py:146-158.
 *   grad_out : [T, n] dL/dz ;  grad_v_T, grad_i_T : [n] or NULL (grad wrt final state)
 *   v_dec_trace : [T, n] from the forward
 *   grad_x : [T, n] ; grad_v0, grad_i0 : [n] or NULL */
SODA_API int soda_lif_bwd(const soda_neuron_params_t* p, const float* grad_out, const float* grad_v_T,
                 const float* grad_i_T, const float* v_dec_trace, int T, int64_t n, float* grad_x,
                 float* grad_v0, float* grad_i0, void* stream);

/* The same with grad_out read in place from a channel slice of a wider pixel-major tensor - the gradient of a Dense
 * concat (models/generator.py:158) as autograd hands it to a branch: element e of step t is channel e % C of pixel
 * t * (n / C) + e / C, pixels grad_out_pix_stride floats apart (>= C; n % C == 0).  Saves the .contiguous() copy of
 * every such slice in the training step.  Everything else as soda_lif_bwd. */
SODA_API int soda_lif_bwd_strided(const soda_neuron_params_t* p, const float* grad_out, int C, int64_t grad_out_pix_stride,
                 const float* grad_v_T, const float* grad_i_T, const float* v_dec_trace, int T, int64_t n,
                 float* grad_x, float* grad_v0, float* grad_i0, void* stream);

/* LI backward: out_t = v_t (pre-tanh; tanh' is applied by the caller's grad_out). */
SODA_API int soda_li_bwd(const soda_neuron_params_t* p, const float* grad_out, const float* grad_v_T,
                const float* grad_i_T, int T, int64_t n, float* grad_x, float* grad_v0,
                float* grad_i0, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Pre-split ("packed") activation formats and the convolution that consumes them (rows N1 / N2 of the scope table:
 * "spikes are written as packed maps that the next layer consumes directly", "LIF fused into the conv epilogue")
 *
 * Between two layers of a generated net (models/generator.py:184-198: the output of one LayerGen module is the
 * input of the next) an activation tensor need not be fp32: the PRODUCER's epilogue writes it in the form the
 * CONSUMER's tensor-core operands have, so the consumer's TMA loads are its MMA operands.
 *   SODA_ACT_F32  : fp32 pixel-major (the external contract: network inputs / outputs, neuron state, StateStorage)
 *   SODA_ACT_PAIR : per pixel and 32-channel group, 128 bytes = fp16 hi[32] | fp16 lo[32] with x = hi + lo
 *                   (hi = rn16(x), lo = rn16(x - hi): 22 significant bits).  Same bytes per value as fp32, hence the
 *                   same strides: element (n, y, x, c) of a PAIR tensor starts at byte
 *                   4 * (n * batch_stride + (y * W + x) * pix_stride + (c & ~31)) + 2 * (c & 31)  (+ 64 for lo).
 *                   Channel counts, channel-slice offsets and strides are multiples of 32; |x| <= 65504.
 *   SODA_ACT_HALF : fp16 pixel-major, for values exact in fp16: spikes {0,1} (layer_gen.py:217-235), sum-pooled
 *                   spike counts (common.py:45-49), sums of spike maps - integers <= 2048.  2 bytes per value.
 * Strides in every descriptor below are in ELEMENTS (channels) of the tensor's own format.
 * ------------------------------------------------------------------------------------------- */
#define SODA_ACT_F32 0
#define SODA_ACT_PAIR 1
#define SODA_ACT_HALF 2
/* flags: the PAIR input's lo halves are all zero (an exact tensor stored as PAIR): a_lo * w_hi is skipped */
#define SODA_PCONV_FLAG_INPUT_EXACT 1

typedef struct {
  int N, H, W, Cin, Cout, k, stride;
  int in_fmt, out_fmt, res_fmt;   /* in: PAIR / HALF; out, res: any */
  int flags;
  int64_t in_batch_stride, in_pix_stride;
  int64_t out_batch_stride, out_pix_stride;
  int64_t res_batch_stride, res_pix_stride;
} soda_pconv_desc_t;

/* Weights [Cout, Cin, k, k] (the state_dict tensor) * scale[Cout] (folded BatchNorm, NULL = 1) * *mult (device
 * scalar, a power of two that puts the largest weight near 2^11, NULL = 1) -> fp16 hi | fp16 lo planes in the
 * k-block order of the input format.  soda_pconv_packed_bytes() bytes, caller-owned.  The convolution multiplies its
 * accumulators by *descale = 1 / *mult. */
SODA_API int64_t soda_pconv_packed_bytes(int Cout, int Cin, int k);
SODA_API int soda_pconv_pack_weights(const float* weight, const float* scale, const float* mult, int Cout, int Cin,
                                     int k, int in_fmt, void* packed, void* stream);
/* 1 if soda_pconv_fwd takes this descriptor and these pointers (alignment), else 0. */
SODA_API int soda_pconv_supported(const soda_pconv_desc_t* d, const void* x, const void* y, const void* residual);
/*   replaces: nn.Conv2d (layer_gen.py:129-136) + eval BatchNorm (:211-214) + Residual sum (generator.py:146) +
 *             concat-slice store (:158), as soda_conv2d_fwd, on PAIR / HALF input.
 *   y = descale * sum(x * w) + shift (+ residual), written as d->out_fmt.  kind::f16 tensor-core products
 *   a_hi*w_hi + a_hi*w_lo + a_lo*w_hi with fp32 accumulation (PAIR) or a*w_hi + a*w_lo (HALF).
 *   status: NULL, or a device counter incremented for every output value beyond fp16's range (PAIR / HALF out). */
SODA_API int soda_pconv_fwd(const soda_pconv_desc_t* d, const void* x, const void* packed, const float* descale,
                            const float* shift, const void* residual, void* y, unsigned int* status, void* stream);
/* Fused Conv -> BatchNorm(eval) -> LIF over T steps on PAIR / HALF input (cf. soda_conv_lif_fwd): membrane state in
 * registers for all T steps, spikes (+ residual) written as d->out_fmt; v0 / i0 / v_out / i_out fp32 [B, Ho, Wo, Cout]. */
SODA_API int soda_pconv_lif_fwd(const soda_pconv_desc_t* d, const void* x, const void* packed, const float* descale,
                                const float* shift, const void* residual, int T, const soda_neuron_params_t* np,
                                const float* v0, const float* i0, void* y, float* v_out, float* i_out,
                                unsigned int* status, void* stream);

/* Conv(3x3, stride 1) -> BatchNorm(eval) -> LIF over T steps -> 2x2 sum pooling of the spikes in ONE kernel, for the
 * small-K blocks of generator nets (reference README.md:57-64: vgg_block(8) / Pool / vgg_block(32) / Pool; layers
 * models/modules/layer_gen.py:129-136, :211-214, :233, models/modules/common.py:45-49; time loop models/generator.py:184-198).
 *   x       : Cin = 2: plain NCHW fp32 frames [T*B, 2, H, W] (in_nchw = 1, the binned events);
 *             Cin = 8: SODA_ACT_HALF pixel-major [T*B, H, W, 8] (in_nchw = 0; the pooled counts of the previous block)
 *   packed  : soda_conv_pack_weights(kind = SODA_CONV_FP32) of the BatchNorm-folded filter;  shift: [Cout] or NULL
 *   y       : pooled spike counts {0..4}, element (t, b, py, px, c), in out_fmt (2 bytes per count as SODA_ACT_HALF)
 *   v0, i0, v_out, i_out : fp32 [B, H, W, Cout] membrane state at FULL resolution (NULL in = initial state)
 *   H, W even; Cout % 8 == 0.  The full-resolution spike map is never written. */
SODA_API int soda_conv_lif_pool_supported(int Cin, int Cout, int k, int stride, int H, int W, int in_nchw, int in_fmt);
SODA_API int soda_conv_lif_pool_fwd(const void* x, int in_fmt, int in_nchw, int64_t in_batch_stride, int64_t in_pix_stride,
                                    const float* packed, const float* shift, int T, int B, int H, int W, int Cin, int Cout,
                                    const soda_neuron_params_t* np, const float* v0, const float* i0, void* y, int out_fmt,
                                    int64_t out_batch_stride, int64_t out_pix_stride, float* v_out, float* i_out,
                                    void* stream);

/* Unfolds plain NCHW fp32 frames x [N, C, H, W] (C * 9 <= 32: the binned events, C = 2) for a 3x3 / pad-1 convolution
 * of the given stride into a pixel-major SODA_ACT_PAIR tensor out [N, Ho, Wo, 32]:
 *   out(n, oy, ox, (ky*3 + kx)*C + ci) = x(n, ci, oy*stride + ky - 1, ox*stride + kx - 1)   (0 outside the frame / beyond C*9)
 * so that the detector's first layer (models/tiny_yolo.py:16: Conv(64, 3, 2) on the event frames, K = 18) runs as a 1x1
 * layer on soda_pconv_lif_fwd.  Occupancy values {0,1} are exact in fp16: lo halves are zero. */
SODA_API int soda_im2col3_frames(const float* x, int N, int C, int H, int W, int stride, void* out, int64_t out_batch_stride,
                                 int64_t out_pix_stride, void* stream);

/* Format conversion / strided copy of a pixel-major tensor [n_frames][pix_per_frame][C] between any two formats
 * (decode to fp32, re-encode): the boundary between the packed region of a net and fp32 callers, and the copy of a
 * tensor into a concat slice of another format (generator.py:158).  C % 32 == 0 when PAIR is involved, else C % 8. */
SODA_API int soda_act_convert(const void* src, int src_fmt, int64_t src_batch_stride, int64_t src_pix_stride, void* dst,
                              int dst_fmt, int64_t dst_batch_stride, int64_t dst_pix_stride, int64_t n_frames,
                              int64_t pix_per_frame, int C, unsigned int* status, void* stream);

/* LIF scan (soda_lif_fwd's arithmetic) on an fp32 conv output with the Residual sum and the store fused, residual
 * and output in any format: x [T][n_pix][C] dense fp32; residual / out element (t, q, c) addressed through their
 * pixel strides (frames of n_pix pixels are dense: batch stride = n_pix * pix_stride); C % 32 == 0. */
SODA_API int soda_lif_fmt_fwd(const soda_neuron_params_t* p, const float* x, const float* v0, const float* i0, int T,
                              int64_t n_pix, int C, const void* residual, int res_fmt, int64_t res_pix_stride, void* out,
                              int out_fmt, int64_t out_pix_stride, float* v_out, float* i_out, unsigned int* status,
                              void* stream);

/* ---------------------------------------------------------------------------------------------
 * Pooling (K5)
 *   replaces: models/modules/common.py:45-49 SumPool2d (avg_pool2d * k * k) and
 *             layer_gen.py:163-168 AvgPool2d / MaxPool2d (floor mode, no padding).
 *   mode: 0 = sum, 1 = average, 2 = max.  Pixel-major in and out (see conventions).
 * ------------------------------------------------------------------------------------------- */
SODA_API int soda_pool2d_fwd(const float* x, int N, int H, int W, int C, int k, int stride, int mode,
                    int64_t in_batch_stride, int64_t in_pix_stride, float* y,
                    int64_t out_batch_stride, int64_t out_pix_stride, void* stream);

/* Per-channel affine on pixel-major data (stand-alone eval BatchNorm, layer_gen.py:211-214, for
 * generated nets where Norm does not follow a Conv) and strided copy into a concat slice. */
SODA_API int soda_channel_affine(const float* x, int64_t n_pix, int C, int64_t in_pix_stride,
                        const float* scale, const float* shift, float* y, int64_t out_pix_stride,
                        void* stream);

/* Sum of k pixel-major tensors of equal shape (Residual merge, generator.py:146). */
SODA_API int soda_add_n(const float* const* xs, int k, int64_t n, float* y, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Detection decode + per-class greedy NMS + row assembly (the step right after the forward in predict)
 *   replaces: utils/box.py:122-153 multibox_detection (with offset_inverse :72-91, box_iou :24-45, nms :94-119),
 *             called by models/soda.py:226-233 / :299-302.
 *   cls_probs : [B, A, C] class probabilities AFTER softmax (index 0 = background), dense
 *   offsets   : [B, A, 4] predicted box offsets;  anchors : [A, 4] corner-format anchors (16-byte aligned)
 *   out       : [B, A, 6] rows (class, conf, x1, y1, x2, y2): kept boxes first (class by class, best first), then
 *               the others in anchor order with class -1; conf < pos_threshold -> class -1, conf = 1 - conf
 *   workspace : soda_multibox_workspace_bytes(B, A) bytes of device memory, 16-byte aligned
 *   One CTA per sample; A <= 262143 anchors (GEN1: 13 545 - keys sorted in shared memory; 1Mpx: 170 280 - keys in the
 *   L2-resident workspace), C <= 64.
 * ------------------------------------------------------------------------------------------- */
SODA_API int64_t soda_multibox_workspace_bytes(int B, int A);
SODA_API int soda_multibox_detection(const float* cls_probs, const float* offsets, const float* anchors, int B, int A,
                            int C, float nms_threshold, float pos_threshold, void* workspace, float* out,
                            void* stream);

/* soda_multibox_detection without a caller workspace, for callers that cannot allocate device memory (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The scratch is a file-scope device array of SODA_MULTIBOX_STATIC_WS_BYTES (8 MiB)
 * that every GPU loading the library holds until exit; the call allocates nothing (no cudaMalloc / cudaFree).
 *   Same arguments, rows and limits as soda_multibox_detection, and additionally
 *   soda_multibox_workspace_bytes(B, A) <= soda_multibox_static_max_bytes()  (GEN1, 13 545 anchors: B <= 14),
 *   else SODA_ERR_UNSUPPORTED with the needed and available byte counts in soda_last_error().
 *   All calls share the one array: calls that overlap on the GPU (two streams, two threads) race - serialise them.
 *   Nothing in the array has to start at zero; it keeps the last call's contents. */
#define SODA_MULTIBOX_STATIC_WS_BYTES (8 << 20)
SODA_API int64_t soda_multibox_static_max_bytes(void);
SODA_API int soda_multibox_detection_static(const float* cls_probs, const float* offsets, const float* anchors, int B,
                            int A, int C, float nms_threshold, float pos_threshold, float* out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Anchor labelling for the detection loss (the step right after the forward in training)
 *   replaces: utils/roi.py:18-63 RoI.__call__ with _assign_anchor_to_box :65-109, utils/box.py:24-57 box_iou and
 *             :60-67 offset_boxes, called by models/soda.py:265 (_loss).
 *   anchors      : [A, 4] corner-format anchors (16-byte aligned)
 *   labels       : [B, G, 5] rows (class, x1, y1, x2, y2); padded rows are all -1 and take part as in the reference
 *   bbox_offset  : [B, A, 4] regression targets, already multiplied by the mask
 *   bbox_mask    : [B, A, 4] 1 for anchors with a box, else 0
 *   class_labels : [B, A] int64, 0 = background, else label class + 1
 *   Each anchor takes its best box if IoU >= iou_threshold; then every box claims the best anchor still free
 *   (ties: smallest anchor * G + box index).  One CTA per sample; G <= 256, A * G < 2^32.
 * ------------------------------------------------------------------------------------------- */
SODA_API int soda_roi_assign(const float* anchors, const float* labels, int B, int A, int G, float iou_threshold,
                             float* bbox_offset, float* bbox_mask, long long* class_labels, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Train-mode BatchNorm over T time steps (BPTT route)
 *   replaces: nn.BatchNorm2d in train mode, models/modules/layer_gen.py:197-214, called once per time step by the
 *             loop of models/soda.py:142 - batch statistics per step over (B, H, W), running statistics updated T
 *             times in sequence - and its autograd backward.
 *   x, y, grad_y, grad_x : [T * B, H, W, C] dense pixel-major (frame t * B + b), 16-byte aligned; n_pix = B * H * W
 *   C                    : 4 * a power of two, <= 1024
 *   gamma, beta          : [C] or NULL (1 / 0);  running_mean / running_var : [C] updated in place, or both NULL
 *   save_mean, save_invstd : [T, C], written by fwd, read by bwd
 *   grad_gamma, grad_beta  : [C] (summed over the T steps) or NULL
 *   workspace            : soda_bn_train_workspace_bytes(T, C) bytes, 16-byte aligned
 *   y = (x - mean_t) * rsqrt(var_t + eps) * gamma + beta, var_t biased; running_var takes var_t * n / (n - 1).
 *   Deterministic (fixed-order reductions, partials combined in fp64).
 * ------------------------------------------------------------------------------------------- */
SODA_API int64_t soda_bn_train_workspace_bytes(int T, int C);
SODA_API int soda_bn_train_fwd(const float* x, int T, int64_t n_pix, int C, const float* gamma, const float* beta,
                               float eps, float momentum, float* running_mean, float* running_var, float* y,
                               float* save_mean, float* save_invstd, void* workspace, void* stream);
SODA_API int soda_bn_train_bwd(const float* grad_y, const float* x, int T, int64_t n_pix, int C, const float* gamma,
                               const float* save_mean, const float* save_invstd, float* grad_x, float* grad_gamma,
                               float* grad_beta, void* workspace, void* stream);

/* Cross-rank (SyncBatchNorm, reference config.yaml:76) variant of the train-mode BatchNorm: statistics, then the
 * caller all-reduces the sums over the ranks (the only exchange; [T][2][C] fp64), then normalise with the global sums.
 *   soda_bn_train_stats        : sums[t][0][c] = sum x, sums[t][1][c] = sum x^2 of this rank (grad_y == NULL), or
 *                                sum gy and sum gy * xhat (grad_y != NULL; needs save_mean / save_invstd)
 *   soda_bn_train_fwd_from_sums: as soda_bn_train_fwd with the statistics taken from the (global) sums over n_total
 *                                pixels per time step; running_var uses n_total / (n_total - 1)
 *   soda_bn_train_bwd_from_sums: dx from the GLOBAL sums / n_total; grad_gamma / grad_beta from sums_local (this rank's
 *                                sums before the all-reduce; NULL = the global ones) */
SODA_API int soda_bn_train_stats(const float* x, const float* grad_y, const float* save_mean, const float* save_invstd,
                        int T, int64_t n_pix, int C, double* sums, void* workspace, void* stream);
SODA_API int soda_bn_train_fwd_from_sums(const float* x, const double* sums, int64_t n_total, int T, int64_t n_pix, int C,
                                const float* gamma, const float* beta, float eps, float momentum, float* running_mean,
                                float* running_var, float* y, float* save_mean, float* save_invstd, void* workspace,
                                void* stream);
SODA_API int soda_bn_train_bwd_from_sums(const float* grad_y, const float* x, const double* sums_global, int64_t n_total,
                                const double* sums_local, int T, int64_t n_pix, int C, const float* gamma,
                                const float* save_mean, const float* save_invstd, float* grad_x, float* grad_gamma,
                                float* grad_beta, void* workspace, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SODA_B200_H */
