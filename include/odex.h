/* odex.h -- C ABI of libodex.so: the sm_100a kernels behind the ODE-RL latent Neural-ODE rollout.
 *
 * The reference (jithendaraa/ODE-RL) has NO FFI / plugin interface on this path: its boundary is four Python
 * nn.Module classes (modules/DiffEqSolver.py:12-80, modules/ODEConvGRUCell.py:9-78, modules/ConvGRUCell.py:11-86)
 * whose forward() bodies are chains of ATen library calls.  Each entry point below names the reference call
 * site(s) (file:line under the reference root) whose ATen op chain it replaces.  The Python package
 * `ode_rl_b200` mirrors the four classes and binds these symbols with ctypes.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - activations are fp32 NHWC ("pixel rows"): element (b, y, x, c) of a tensor with row stride ld lives at
 *     ((b*H + y)*W + x)*ld + c; ld >= channels lets a kernel read/write a channel slice of a wider tensor;
 *   - times / step sizes / error ratios are fp64 and live in device memory so no entry point needs a host sync;
 *   - every entry takes the CUDA stream to enqueue on (cudaStream_t passed as void*), is re-entrant, never
 *     synchronises and never throws: it returns 0 on success, a negative cudaError_t on a CUDA failure and a
 *     positive ODEX_E_* code on an argument error.  odex_last_error() returns the message of the last failure
 *     on the calling thread;
 *   - `skip` (nullable) is a device int: when *skip != 0 at kernel start the whole launch is a no-op.  The
 *     adaptive solver enqueues attempts speculatively and sets it when the integration is finished.
 */
#ifndef ODEX_H
#define ODEX_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ODEX_VERSION 200

enum {
    ODEX_E_ARG = 1,         /* bad argument (null pointer, negative size ...) */
    ODEX_E_SHAPE = 2,       /* shape not supported by the kernels (see each entry) */
    ODEX_E_ALIGN = 3,       /* pointer / stride not 16-byte aligned */
    ODEX_E_UNSUPPORTED = 4, /* requested implementation not available for this shape */
    ODEX_E_DRIVER = 5       /* cuTensorMapEncodeTiled could not be resolved / failed */
};

enum { ODEX_ACT_NONE = 0, ODEX_ACT_RELU = 1, ODEX_ACT_TANH = 2 };
enum { ODEX_CONV_AUTO = 0, ODEX_CONV_SIMT = 1, ODEX_CONV_TC = 2 };

typedef void* odex_stream_t; /* cudaStream_t */

int odex_version(void);
const char* odex_last_error(void);
/* number of kernels this library has launched in the calling process (bench.py's `gpu_launches`). */
long long odex_launch_count(void);

/* ------------------------------------------------------------------------------------------------ weights
 * Repack a torch Conv2d weight (OIHW, [N][C][R][R], fp32) for the kernels.
 *   dgrad == 0: the forward operator.
 *   dgrad != 0: the operator of the input gradient (channels swapped, taps flipped), so the data-gradient of a
 *               convolution is computed by the same forward kernels.
 * SIMT layout: [tap][c_in][n_out] fp32.
 * TC layout  : per k-block (64-channel group, tap, 32-channel half): 2*n_out rows of 32 floats =
 *              [tf32 hi rows ; tf32 lo rows] (3xTF32 split), ready for SWIZZLE_128B TMA boxes.
 * F16 layout (tc == 2): per k-block (64-channel group, tap): 2*n_out rows of 64 halves = [fp16 hi rows ; fp16 lo rows] of
 *              w * s (lo scaled by 2^11; s = the power of two that puts max |w| in [2^12, 2^13)), followed by a 4-float
 *              tail {1/s, max_n sum_k |w[n,k]|, max |w|, 0} the kernels use to scale / bound their fp16 operands.
 * Sizes are in floats.  Replaces: nothing in the reference (cuDNN/oneDNN do their own packing). */
size_t odex_packed_weight_floats(int n_out, int c_in, int R, int tc);
int odex_pack_weights(const float* w_oihw, int N, int C, int R, int dgrad, int tc, float* out, odex_stream_t s);
/* odex_pack_weights(tc = 2) of BOTH orientations of one weight tensor (dgrad = 0 into out_fwd, dgrad = 1 into out_dgrad; each
 * odex_packed_weight_floats(N, C, R, 2) floats) in one pair of launches: a training step repacks every convolution after the
 * optimizer step, and the four tiny launches per convolution were 2.5 % of the config-A step.  N, C multiples of 64. */
int odex_pack_weights_f16_pair(const float* w_oihw, int N, int C, int R, float* out_fwd, float* out_dgrad, odex_stream_t s);

/* ------------------------------------------------------------------------------------------------ conv
 * out[p, n] = epi( bias[n] + sum_{r,s,c} X[p + (r,s) - pad, c] * W[n, c, r, s] ),  stride 1, pad (R-1)/2,
 * where X is the channel concatenation of up to two NHWC sources (torch.cat((x, h), 1) folded in) and
 *   v = act(acc + bias);  v *= act'(dact) if dact != NULL;  out = res + alpha * v if res != NULL else v,
 * alpha = alpha_host * (alpha_dev ? (float)*alpha_dev : 1).
 * Replaces: conv2d (+ReLU) of helpers/utils.py:167-177 via modules/DiffEqSolver.py:77; conv2d of
 * modules/ConvGRUCell.py:41-43,47-49 with the torch.cat of :72,79; the Euler increment of
 * modules/ODEConvGRUCell.py:54,57 (res/alpha epilogue); the 1x1 convs of modules/ODEConvGRUCell.py:27-30;
 * and, with dgrad-packed weights + dact, the autograd conv backward-data of all of those.
 * Shapes: R in {1,3,5}; W % 16 == 0, H % 8 == 0; c0, c1, N multiples of 4 (SIMT).  TC additionally needs
 * c0 (== c1 if present) % 64 == 0, N = 64 or a multiple of 128 (blocks of 128 output channels, one launch each),
 * W a power of two <= 128 with (128/W) | H. */
typedef struct odex_conv_desc {
    const float* x0; int ld0; int c0;
    const float* x1; int ld1; int c1;       /* x1 == NULL: single source */
    int B, H, W, R, N;
    const float* w_simt;                    /* odex_pack_weights(tc=0) or NULL */
    const float* w_tc;                      /* odex_pack_weights(tc=1) or NULL */
    const float* bias;                      /* [N] or NULL */
    int act;                                /* ODEX_ACT_* applied to acc + bias */
    const float* dact; int ldd; int dact_kind; /* saved activation a: v *= (a > 0) [RELU] / (1 - a*a) [TANH] */
    const float* res; int ldr;              /* optional residual */
    float alpha; const double* alpha_dev;
    float* out; int ldo;
    const int* skip;
    int impl;                               /* ODEX_CONV_* */
} odex_conv_desc;
int odex_conv2d(const odex_conv_desc* d, odex_stream_t s);
/* Measurement helper (bench.py): the same launch enqueued `reps` times back to back from C. */
int odex_conv2d_repeat(const odex_conv_desc* d, int reps, odex_stream_t s);

/* Weight / bias gradient of the same convolution: dw[n,c,r,s] += sum_p dy[p,n] * X[p + (r,s) - pad, c]
 * (OIHW, torch layout, ACCUMULATED into dw) and db[n] += sum_p dy[p,n] (db nullable).
 * Replaces: autograd conv backward-weight / bias of the call sites listed for odex_conv2d. */
int odex_conv2d_wgrad(const float* x0, int ld0, int c0, const float* x1, int ld1, int c1, const float* dy, int ldy,
                      int B, int H, int W, int R, int N, float* dw, float* db, odex_stream_t s);

/* Same gradient, accumulated into the kernels' own layout dwp[tap][n][c] ([R*R][N][c0+c1], c contiguous) so that the
 * tensor-core kernel can add its TMEM accumulators with 16-byte vector reductions and a backward pass that visits a
 * layer many times (one visit per f-eval) converts to OIHW only once, with odex_wgrad_unpack (dw[n][c][r][s] +=
 * dwp[r*R+s][n][c]).  impl: ODEX_CONV_AUTO / _SIMT / _TC.  The TC kernel (tcgen05 kind::tf32, 3xTF32; dy as the K-major A operand in
 * TMEM, the tap-shifted X atoms as MN-major B operands straight from the NHWC halo tile) needs c0, c1, N % 64 == 0, W % 16 == 0,
 * H % 4 == 0, R in {1,3,5}.  Replaces the same autograd ops as odex_conv2d_wgrad. */
int odex_conv2d_wgrad_packed(const float* x0, int ld0, int c0, const float* x1, int ld1, int c1, const float* dy,
                             int ldy, int B, int H, int W, int R, int N, float* dwp, float* db, int impl,
                             odex_stream_t s);
int odex_wgrad_unpack(const float* dwp, int N, int C, int R, float* dw, odex_stream_t s);
/* odex_conv2d_wgrad_packed on tcgen05 kind::f16 with the two-term fp16 operand split (x * s = fp16 hi + fp16 lo, s per tensor):
 * same result to fp32 rounding at twice the MMA rate.  The per-tensor scales come from a max-|.| pass over X and dy that this
 * entry enqueues first; scratch2 = 2 device floats it may use (contents irrelevant on entry, overwritten).  Same shapes as the TC
 * kernel (c0, c1, N % 64 == 0, W % 16 == 0, H % 4 == 0, R in {1,3,5}). */
int odex_conv2d_wgrad_f16(const float* x0, int ld0, int c0, const float* x1, int ld1, int c1, const float* dy, int ldy,
                          int B, int H, int W, int R, int N, float* dwp, float* db, float* scratch2, odex_stream_t s);
/* The same kernel when the caller already holds PER-IMAGE upper bounds of the operands' maxima: amax_x[b] >= max |X[b]| (over
 * both sources), amax_dy[b] >= max |dy[b]|, b < B, in device memory -- as odex_convnet / odex_conv16 leave them (amax_out) while
 * they produce or read those operands.  No max pass.  The kernel scales each image's operands by its own maxima and keeps only the
 * scale of the PRODUCT common to the launch (from max_b amax_x[b] * amax_dy[b]), so images whose magnitudes differ by any factor
 * can share a launch: an image loses operand bits only in proportion to how little it contributes to the sum. */
int odex_conv2d_wgrad_f16_scaled(const float* x0, int ld0, int c0, const float* x1, int ld1, int c1, const float* dy, int ldy,
                                 int B, int H, int W, int R, int N, float* dwp, float* db, const float* amax_x, const float* amax_dy,
                                 odex_stream_t s);

/* ------------------------------------------------------------------------------------------------ fused conv chain
 * A chain of L 3x3 C->C convolutions (C = 64) over 16x16 images in ONE launch, the image resident in shared memory
 * between layers:  a_0 = x;  v_l = act_l(conv_l(a_l) + bias_l);  v_l = v_l * act'(aux_l) [aux_kind 1 = ReLU', 2 = Tanh',
 * aux_l the saved activation]  or  v_l = aux_l + alpha * v_l [aux_kind 3, residual];  a_{l+1} = v_l;  out[l] = v_l when
 * out[l] != NULL (out[L-1] is required).  alpha = alpha * (alpha_dev ? (float)*alpha_dev : 1).
 * With forward-packed weights this is ODEFunc.forward (modules/DiffEqSolver.py:71-80 over the Sequential of
 * helpers/utils.py:158-183) including the saved activations and the fused `y + dt * f` increment of the solvers;
 * with dgrad-packed weights, aux = saved activations (masks) and out = every layer's input gradient, it is the
 * autograd backward-data chain of the same network.  Same 3xTF32 tcgen05 arithmetic as odex_conv2d(TC).
 * odex_convnet_supported returns 1 when the shape is taken (H = W = 16, C = 64, 1 <= L <= ODEX_NET_MAX_LAYERS, every
 * w_tc present, strides % 4 == 0), else 0; odex_convnet on an unsupported shape returns ODEX_E_UNSUPPORTED. */
#define ODEX_NET_MAX_LAYERS 6
typedef struct odex_convnet_desc {
    const float* x; int ldx;
    int B, H, W, C, L;
    const float* w_tc[ODEX_NET_MAX_LAYERS];     /* odex_pack_weights(tc=1) of each layer */
    const float* bias[ODEX_NET_MAX_LAYERS];     /* [C] or NULL */
    int act[ODEX_NET_MAX_LAYERS];               /* ODEX_ACT_* */
    const float* aux[ODEX_NET_MAX_LAYERS]; int ld_aux[ODEX_NET_MAX_LAYERS]; int aux_kind[ODEX_NET_MAX_LAYERS];
    float* out[ODEX_NET_MAX_LAYERS]; int ld_out[ODEX_NET_MAX_LAYERS];
    float alpha; const double* alpha_dev;
    const int* skip;
    int split;                                  /* 0: w_tc are odex_pack_weights(tc = 1) blobs, 3xTF32 on tcgen05 kind::tf32;
                                                   1: w_tc are odex_pack_weights(tc = 2) blobs, two-term fp16 split on kind::f16
                                                      (same 22 significant bits per operand, twice the MMA rate) */
    float* amax_out[ODEX_NET_MAX_LAYERS];       /* split = 1 only, each nullable: amax_out[l][b] <- max |a_l[b]| for every image b < B
                                                   (a_0 = x, a_l = v_{l-1}); B floats each.  These are the operands of the layers'
                                                   weight gradients: odex_conv2d_wgrad_f16_scaled takes them instead of running a
                                                   max pass.  split = 0 rejects a non-NULL entry. */
} odex_convnet_desc;
int odex_convnet(const odex_convnet_desc* d, odex_stream_t s);
int odex_convnet_supported(const odex_convnet_desc* d);

/* ------------------------------------------------------------------------------------------------ image-resident conv + ConvGRU gates
 * Stride-1 "same" convolution (R = 3 or 5) over 16x16 images with one or two 64-channel sources (torch.cat((x, h), 1) folded in
 * as two K ranges) and N output channels (multiple of 64), on tcgen05 kind::f16 with the two-term fp16 operand split of
 * odex_convnet(split = 1); one CTA = one image x 64 output channels, which at 16x16 holds whole GroupNorm groups, so the
 * ConvGRU gate math rides in the epilogue (north_star (c)):
 *   ODEX_C16_PLAIN      v = act(acc + bias); aux_kind 1 / 2: v *= relu' / tanh'(aux); 3: v = aux + alpha * v; out[p*ldo + n] = v.
 *                       Replaces odex_conv2d at these shapes (the ConvGRU data-gradient convolutions among them).
 *   ODEX_C16_GRU_GATES  N = 2C: g = acc + bias; a = sigmoid(GroupNorm(g; 32-channel groups, gamma, beta, eps)); channels [0, C):
 *                       z[p*C + c] = a; channels [C, 2C): rh[p*C + c - C] = a * h[p*ldh + c - C].  g_raw (nullable) receives g
 *                       [B,256,2C] for the backward; stats[b][group] = (mean, rstd).
 *                       Replaces conv_gates of modules/ConvGRUCell.py:40-44 + split / sigmoid / r*h of :72-79.
 *   ODEX_C16_GRU_OUT    N = C: c = tanh(GroupNorm(acc + bias)); hnew[p*ld_hnew + n] = (1 - z_in) * h + z_in * c; cand / g_raw
 *                       (nullable) receive c and the pre-norm convolution for the backward.
 *                       Replaces conv_can of modules/ConvGRUCell.py:46-50 + tanh / blend of :80-82.
 * w_f16 = odex_pack_weights(tc = 2) of the [N][64 * nsrc][R][R] operator. */
enum { ODEX_C16_PLAIN = 0, ODEX_C16_GRU_GATES = 1, ODEX_C16_GRU_OUT = 2 };
typedef struct odex_conv16_desc {
    const float* x0; int ld0;
    const float* x1; int ld1;               /* NULL: one source */
    int B, R, N;
    const float* w_f16;
    const float* bias;                      /* [N] or NULL */
    int mode;
    int act; const float* aux; int ld_aux; int aux_kind; float alpha; const double* alpha_dev;
    float* out; int ldo;
    const float* gamma; const float* beta; float eps;
    const float* h; int ldh; const float* z_in;
    float* z; float* rh; float* hnew; int ld_hnew; float* cand; float* g_raw; float* stats;
    const int* skip;
    float* out_hi; int ld_out_hi; int out_split; /* PLAIN, optional (out_split = 0: off): output channels [out_split, N) go to out_hi
                                               (row stride ld_out_hi, channel c at column c - out_split) instead of out, which
                                               then holds [0, out_split) only; out_split a multiple of 64.  The ConvGRU data
                                               gradient [dx | dh] lands in two dense tensors this way instead of one interleaved. */
    float* amax_out;                        /* nullable, B floats: amax_out[b] <- max(max |x0[b]|, max |x1[b]|) (the kernel measures it
                                               for its operand scale anyway).  For odex_conv2d_wgrad_f16_scaled: the sources of a
                                               forward launch are the X of that convolution's weight gradient, the source of its
                                               data-gradient launch (mode PLAIN over dy) is the dy. */
} odex_conv16_desc;
int odex_conv16(const odex_conv16_desc* d, odex_stream_t s);
int odex_conv16_supported(const odex_conv16_desc* d);

/* ------------------------------------------------------------------------------------------------ decoder tail
 * LeakyReLU(slope) -> ConvTranspose2d(C, 1, 4, 2, 1) of ConvDecode (models/ODEConvGRU.py:122-140, the activation and the
 * last layer), reading the PHASE-BLOCKED pre-activation output of the first transposed convolution as the tensor-core
 * path produces it: p[n][Y][X][(a*2+b)*C + ci] = channel ci at fine pixel (2Y+a, 2X+b).  out: [n][4*cells_h][4*cells_w].
 * w: [C][16] (torch weight [C,1,4,4]), b: [1] or NULL.  Backward: dp (same layout as p), dw[C][16] += , db[1] += (nullable).
 * Shapes taken: cells_h = cells_w = 16, C = 32 (config A); otherwise ODEX_E_UNSUPPORTED (the caller stays on torch). */
int odex_decode_tail_fwd(const float* p, int n, int cells_h, int cells_w, int C, const float* w, const float* b,
                         float slope, float* out, odex_stream_t s);
int odex_decode_tail_bwd(const float* p, const float* dout, int n, int cells_h, int cells_w, int C, const float* w,
                         float slope, float* dp, float* dw, float* db, odex_stream_t s);

/* ------------------------------------------------------------------------------------------------ encoder stem
 * Conv2d(1, C, 3, 2, 1) + LeakyReLU(slope) of ConvEncode (models/ODEConvGRU.py:101-119, the first layer and its activation)
 * on [n][H][W] single-channel frames, written in the space-to-depth layout the strided tensor-core convolution reads:
 * cells[n][Y][X][(a*2+b)*C + c] = leaky(conv(x) + b)[n][c][2Y+a][2X+b], Y < H/4, X < W/4.  w: [C][9] (torch [C,1,3,3]), b: [C]
 * or NULL.  Backward (no gradient for the frames): dw[C][9] +=, db[C] += (nullable) from dcells and the saved cells.
 * Shapes taken: H = W = 64, C = 16 (config A); otherwise ODEX_E_UNSUPPORTED (the caller stays on torch). */
int odex_encode_stem_fwd(const float* x, int n, int H, int W, int C, const float* w, const float* b, float slope,
                         float* cells, odex_stream_t s);
int odex_encode_stem_bwd(const float* x, const float* cells, const float* dcells, int n, int H, int W, int C, float slope,
                         float* dw, float* db, odex_stream_t s);

/* ------------------------------------------------------------------------------------------------ bilinear 2x upsample
 * nn.Upsample(scale_factor=2, mode='bilinear', align_corners=False) of the VidODE per-step decoder (models/VidODE.py:34),
 * NHWC fp32: x [n][H][W][C] -> out [n][2H][2W][C]; _bwd is its adjoint: dout [n][2H][2W][C] -> dx [n][H][W][C] (overwritten).
 * C must be a multiple of 4, pointers 16-byte aligned. */
int odex_upsample2x_fwd(const float* x, int n, int H, int W, int C, float* out, odex_stream_t s);
int odex_upsample2x_bwd(const float* dout, int n, int H, int W, int C, float* dx, odex_stream_t s);

/* ------------------------------------------------------------------------------------------------ VidODE torch-side remainder
 * BatchNorm2d (training mode: batch statistics over all M = N*H*W rows, biased variance, eps) fused with ReLU (relu != 0), NHWC
 * rows x[M][C] (C % 4 == 0, C <= 256): the `nn.Conv2d, nn.BatchNorm2d, nn.ReLU` triples of models/VidODE.py:16-19,36.
 * stats[3][C] <- (mean, rstd, biased variance) (the caller updates the module's running statistics from it);
 * scratch: odex_bn_scratch_floats(M, C) floats.  _bwd: dx, and dgamma_dbeta[2][C] <- (dbeta = sum g, dgamma = sum g * xhat),
 * g = dy * (y > 0).  Statistics are centred two-level sums merged in a fixed order (deterministic, no atomics). */
long long odex_bn_scratch_floats(long long M, int C);
int odex_bn_relu_fwd(const float* x, long long M, int C, const float* gamma, const float* beta, float eps, int relu, float* y,
                     float* stats, float* scratch, long long scratch_floats, odex_stream_t s);
int odex_bn_relu_bwd(const float* x, const float* y, const float* dy, long long M, int C, const float* gamma, const float* stats,
                     int relu, float* dx, float* dgamma_dbeta, float* scratch, long long scratch_floats, odex_stream_t s);
/* odex_bn_relu_fwd for callers that cannot allocate (code under CUDA-graph capture, serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  Same arguments and results as
 * odex_bn_relu_fwd minus the scratch buffer: the slab statistics go through a `static __device__` array inside the library
 * (odex_bn_static_scratch_floats() = 151 552 floats = 592 KiB, held by every GPU that loads the library until the process
 * exits), whose device address is resolved once per device with cudaGetSymbolAddress.  Calls no cudaMalloc / cudaFree or any
 * allocating routine.  Limit: 2 * nblocks * C <= 151 552 with nblocks = ceil(M / max(64, ceil(M / 592))), i.e. every M for
 * C <= 128 and M <= 18 944 rows (296 slabs) for C = 256; beyond it returns ODEX_E_UNSUPPORTED with a message.  All calls on a
 * device share the one array (nothing in it has to start at zero: a call writes every slot it reads): calls that overlap on
 * the GPU -- different streams, or graph branches -- race; serialise them on one stream. */
long long odex_bn_static_scratch_floats(void);
int odex_bn_relu_fwd_static(const float* x, long long M, int C, const float* gamma, const float* beta, float eps, int relu,
                            float* y, float* stats, odex_stream_t s);
/* One step of the recursive flow warp + mask composition of models/VidODE.py:121-131,162-186:
 *   warped = grid_sample(prev, grid + flow / ((size - 1) / 2), bilinear, padding_mode = 'border', align_corners = False)
 *   pred   = sigmoid(mask) * warped + (1 - sigmoid(mask)) * intermediate
 * prev / warped / pred: NCHW frames [B][C][H][W], C <= 3; head: the decoder head's NHWC output, row stride ldh floats, channels
 * (flow_x, flow_y, intermediate[C], mask logit); lin_x[W], lin_y[H]: the reference's torch.linspace(-1, 1, .) base grid. */
int odex_warp_compose_fwd(const float* prev, const float* head, int ldh, const float* lin_x, const float* lin_y, int B, int C,
                          int H, int W, float* warped, float* pred, odex_stream_t s);

/* ------------------------------------------------------------------------------------------------ ConvGRU gates
 * GroupNorm (groups of 32 channels, biased variance, eps, affine) fused with the gate non-linearities.
 * g: conv_gates output [B,H,W,2C] (bias included).  z = sigmoid(GN(g)[:, :C]); r = sigmoid(GN(g)[:, C:]);
 * rh = r * h.  stats[b][group][2] = (mean, rstd) is saved for the backward.
 * Replaces: GroupNorm of modules/ConvGRUCell.py:44 + split/sigmoid/mul of :75-79. */
int odex_gru_gates_fwd(const float* g, const float* h, int ldh, const float* gamma, const float* beta, float eps,
                       int B, int HW, int C, float* z, float* rh, float* stats, odex_stream_t s);
/* c = tanh(GN(cpre)); hnew = (1 - z) * h + z * c.  Replaces modules/ConvGRUCell.py:50,80,82. */
int odex_gru_out_fwd(const float* cpre, const float* z, const float* h, int ldh, const float* gamma,
                     const float* beta, float eps, int B, int HW, int C, float* hnew, float* cand, float* stats,
                     odex_stream_t s);
/* backward of odex_gru_out_fwd: given dhnew -> dcpre, dz, dh (dh = dhnew * (1 - z), written), dgamma/dbeta +=. */
int odex_gru_out_bwd(const float* dhnew, const float* cpre, const float* cand, const float* z, const float* h,
                     int ldh, const float* gamma, const float* stats, int B, int HW, int C, float* dcpre,
                     float* dz, float* dh, float* dgamma, float* dbeta, odex_stream_t s);
/* backward of odex_gru_gates_fwd: given dz, drh (row stride lddrh) -> dg [B,H,W,2C];
 * dh_out[p*lddh + c] = dh_in[p*C + c] + drh * r (dh_out may alias drh); dgamma/dbeta +=. */
int odex_gru_gates_bwd(const float* dz, const float* drh, int lddrh, const float* dh_in, const float* g,
                       const float* h, int ldh, const float* gamma, const float* beta, const float* stats,
                       int B, int HW, int C, float* dg, float* dh_out, int lddh, float* dgamma, float* dbeta,
                       odex_stream_t s);

/* ------------------------------------------------------------------------------------------------ Runge-Kutta
 * out = sum_{j<nin} c_j * in_j,  c_j = (float)coef_host[j] * (dtmask>>j & 1 ? (float)*dt_dev : 1).
 * Operands are summed j ascending in fp32 (torchdiffeq `k[..., :i+1].matmul(beta_i * dt)` order).
 * The stage planes are separate contiguous buffers (stage-major), never the reference's [...,7] interleave.
 * Replaces: every stage combination of torchdiffeq 0.2.1 rk_common._runge_kutta_step / fixed_grid
 * (euler `y + dt*f`, rk4_alt_step_func) called from modules/DiffEqSolver.py:45-46, and their transposes in
 * the backward.  nin <= 8; n % 4 == 0. */
int odex_lincomb(float* out, int nin, const float* const* in_host, const double* coef_host, uint32_t dtmask,
                 const double* dt_dev, long long n, const int* skip, odex_stream_t s);

/* out = dy * act'(a) with a the saved OUTPUT of the activation (ODEX_ACT_*): gradient through a trailing Tanh of
 * helpers/utils.py:179-181 (create_convnet final_act).  n % 4 == 0. */
int odex_act_bwd(float* out, const float* dy, const float* a, int kind, long long n, odex_stream_t s);

/* Device-resident controller of the adaptive solver: t, dt and the accept/reject decision never leave the GPU.
 * Layout is part of the ABI (the host only reads it back, once per forward, or polls a pinned copy).
 * The buffer the caller allocates must be odex_dopri_state_bytes() long: this header (padded to
 * ODEX_STATE_HEAD_BYTES) is followed by the reduction scratch of the norm kernels. */
#define ODEX_STATE_HEAD_BYTES 256
#define ODEX_RED_BLOCKS 592 /* 4 x 148 SMs */
typedef struct odex_dopri_state {
    double t0;          /* start of the last accepted step */
    double t1;          /* current time (end of the last accepted step) */
    double dt;          /* size of the next attempt */
    double dt_try;      /* size of the attempt being evaluated */
    double rtol, atol;
    double red[4];      /* reduction slots {sum of squares, aux, aux, element count}: per-rank partial sums; a
                           multi-GPU caller all-reduces (SUM) these 4 doubles between a norm kernel and the
                           controller kernel so every rank takes the step sequence of the global batch */
    double ratio;       /* error ratio of the last attempt */
    double h0, d0, d1, d2; /* initial-step heuristic scratch */
    int accept;         /* last attempt accepted */
    int done;           /* the integration is over (every output time decided, or an error bit set) -- this is what
                           `skip` points at */
    int adv_seq;        /* number (`seq`, given by the host) of the last ACCEPTED attempt: odex_dopri_advance(seq) runs
                           iff it finds its own number here; -1 before the first accepted step */
    int jnext;          /* next output index to produce */
    int jlo, jhi;       /* outputs [jlo, jhi) lie inside the step accepted by the last attempt */
    int n_out;          /* number of output times */
    int n_accept, n_reject, n_feval, n_attempt;
    int status;         /* bit0: underflow in dt, bit1: non-finite state, bit2: max attempts exceeded,
                           bit3: more accepted steps than backward checkpoints (save_cap) */
    int max_attempts;
    int save_cap;       /* capacity of the (y, f) checkpoint arrays handed to odex_dopri_advance; 0 = none kept */
    double dt_acc;      /* size of the last accepted step (what odex_dopri_advance interpolates with) */
} odex_dopri_state;
size_t odex_dopri_state_bytes(void);

/* state <- zeros, t0 = t1 = t[0], tolerances, n_out, save_cap; done if n_out <= 1.  (torchdiffeq solvers.py setup) */
int odex_dopri_init(odex_dopri_state* st, const double* t_dev, int n_out, double rtol, double atol,
                    int max_attempts, int save_cap, odex_stream_t s);
/* phase 0: red <- {sum (y0/scale)^2, sum (f0/scale)^2, 0, n}; scale = atol + rtol*|y0|.
 * phase 1: red <- {sum ((f1 - f0)/scale)^2, 0, 0, n}.
 * Replaces torchdiffeq misc._select_initial_step's three RMS norms.  fuse_ctrl != 0: the last block to finish also does
 * odex_dopri_initctrl(phase) (single-GPU mode: one launch instead of two). */
int odex_dopri_initnorm(odex_dopri_state* st, int phase, const float* y0, const float* f0, const float* f1,
                        long long n, int fuse_ctrl, odex_stream_t s);
/* phase 0: h0 from d0, d1 (st->h0; the host then forms y0 + h0*f0 with odex_lincomb(dt_dev = &st->h0)).
 * phase 1: dt from d1, d2, h0, latched as dt_try of the first attempt together with rk_common._adaptive_step's
 * `t + dt > t` assert (status bit0).  fp32 arithmetic like the reference. */
int odex_dopri_initctrl(odex_dopri_state* st, int phase, odex_stream_t s);
/* K4, single-GPU form -- ONE kernel for everything between the last f-eval of an attempt and the step decision:
 *   err = dt_try * sum_j c_error[j] k_j (k[1] never read, c_error[1] == 0);  ratio = RMS(err / (atol + rtol*max(|y0|,|y1|)))
 *   over the whole batch tensor (vectorised loads, warp-shuffle -> block -> last-block reduction in a fixed order);
 *   non-finite check of y0;  and, in the last block: accept = ratio <= 1, t / dt update (misc._optimal_step_size), counters,
 *   the range [jlo, jhi) of output times inside the accepted step, the attempt log (4 doubles per attempt: t0, dt, ratio,
 *   accept) into log_dev when non-NULL, `done`, and the asserts of the NEXT attempt (dt underflow, max attempts).
 * `seq` = the caller's running number of this attempt (0, 1, 2, ... per integration), see odex_dopri_advance.
 * Replaces rk_common._runge_kutta_step's error estimate + misc._compute_error_ratio + _optimal_step_size + the accept
 * branch of _adaptive_step (~15 launches and >= 4 D2H syncs in the reference). */
int odex_dopri_errctrl(odex_dopri_state* st, const double* t_dev, const float* y0, const float* y1,
                       const float* const* k7_host, long long n, double* log_dev, int log_cap, int seq, odex_stream_t s);
/* K4, multi-GPU form: odex_dopri_errnorm leaves the per-rank partial sums in st->red (the caller all-reduces those 4
 * doubles over the ranks), odex_dopri_ctrl then takes the decision exactly as odex_dopri_errctrl's last block does. */
int odex_dopri_errnorm(odex_dopri_state* st, const float* y0, const float* y1, const float* const* k7_host,
                       long long n, odex_stream_t s);
int odex_dopri_ctrl(odex_dopri_state* st, const double* t_dev, double* log_dev, int log_cap, int seq, odex_stream_t s);
/* K5: runs iff attempt `seq` was accepted (st->adv_seq == seq): quartic dense output (interp._interp_fit/_interp_evaluate)
 * for every output j in [jlo, jhi) into sol + j*n, then the FSAL advance y0 <- y1, k[0] <- k[6].  Otherwise a no-op.
 * ysave/fsave (nullable, together): [save_cap][n] checkpoints of (y, f) at the start of every accepted step (for the
 * backward replay); the controller stops with status bit3 before a step that would not fit. */
int odex_dopri_advance(odex_dopri_state* st, const double* t_dev, float* y0, const float* y1,
                       float* const* k7_host, float* sol, long long n, float* ysave, float* fsave, int save_cap, int seq,
                       odex_stream_t s);

/* *flag |= bit if x[0..n) holds a NaN (n % 4 == 0).  The encoder recurrence's per-frame `assert not torch.isnan(...)`
 * (modules/ODEConvGRUCell.py:56,59) as a sticky device word: a NaN produced in any iteration reaches the last hidden
 * state, so one pass over it per forward is equivalent, and the host reads the word once (or never, under graph capture). */
int odex_nan_flag(const float* x, long long n, int* flag, int bit, odex_stream_t s);

/* fp32 NCHW <-> NHWC repack of [B, C, HW] planes (tiled transpose). */
int odex_nchw_to_nhwc(const float* in, float* out, int B, int C, int HW, odex_stream_t s);
int odex_nhwc_to_nchw(const float* in, float* out, int B, int C, int HW, odex_stream_t s);

#ifdef __cplusplus
}
#endif
#endif /* ODEX_H */
