/*
 * dfq_b200.h -- C ABI of libdfq_b200.so (sm_100a), the drop-in boundary of the DFQ hot path.
 *
 * The reference (KadAMRN/Data_Free_Quantization) is pure Python/PyTorch: it has no FFI.  The
 * entry points below are what a maintainer binds (ctypes) underneath the
 * reference's own Python functions; each one names the reference code it replaces.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - tensors are contiguous fp32: weights OIHW [O, I/groups, kh, kw] (or [O, I]),
 *     activations NCHW;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - return value is a cudaError_t (0 = success); nothing allocates, nothing synchronises
 *     unless stated;
 *   - there is no CPU fallback: without a CUDA device every call returns an error.
 */
#ifndef DFQ_B200_H
#define DFQ_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DFQ_ABI_VERSION 23

#if defined(__GNUC__)
#pragma GCC visibility push(default)   /* the library is built with -fvisibility=hidden */
#endif

/* flags of DfqQuantDesc.flags */
#define DFQ_Q_SYMMETRIC   1   /* utils/quantize.py:51-60 */
#define DFQ_Q_RANGE_F32   2   /* scale formed in fp32 (min/max=None path, utils/quantize.py:27-37,66) */
#define DFQ_Q_PER_CHANNEL 4   /* one range per row of x.view(rows, -1) (new capability) */
#define DFQ_Q_CLAMP_OUT   8   /* clamp the dequantised output to [clip_lo, clip_hi]: clip_weight.py:29
                                 fused behind quantize_targ_layer, as main_dfq.py:214,227 orders them */

/* fp32 constants of the elementwise pass  x -> round(clamp((x+neg_min)/scale, qmin, qmax))*scale+add_back
 * (utils/quantize.py:72-76). */
typedef struct DfqQuantParams {
    float neg_min, scale, qmin, qmax, add_back;
} DfqQuantParams;

/* One tensor of a multi-tensor fake-quant launch (utils/layer_transform.py:288-305 loops these). */
typedef struct DfqQuantDesc {
    const float *x;      /* input                                                     */
    float *y;            /* dequantised output (may alias x)                          */
    void *codes;         /* optional integer codes q-qmin: uint8 if bits<=8 else uint16 */
    int64_t n;           /* elements                                                  */
    int64_t rows;        /* per-channel: number of rows (n % rows == 0); else 1       */
    int32_t bits;
    int32_t flags;
    float clip_lo, clip_hi;
    int64_t first_block; /* filled by dfq_quant_plan_host                             */
    int64_t range_slot;  /* filled by dfq_quant_plan_host: index into the range array */
} DfqQuantDesc;

/* One relation of cross-layer equalisation (utils/relation.py:6-33 + Cross_layer_equal.py:11-59). */
typedef struct DfqClePair {
    float *W1;           /* [C1, K1]  layer_first.weight                             */
    float *W2;           /* [O2, I2, KK] layer_second.weight                          */
    float *b1;           /* [C1] layer_first.bias (never NULL: reference adds zeros, :93-94) */
    float *bn_w;         /* [C1] BN fake_weight or NULL                               */
    float *bn_b;         /* [C1] BN fake_bias or NULL                                 */
    float *S;            /* [C1] Relation.S, multiplied by every iteration's s (utils/relation.py:26-30);
                            the caller fills it with ones when the relation has no S yet            */
    int32_t C1, K1, O2, I2, KK;
    int32_t stage;       /* position of the pair inside its chain (pairs of one stage are independent) */
    int32_t first_unit;  /* filled by dfq_cle_plan_host                               */
    int32_t n_units;     /* filled by dfq_cle_plan_host                               */
} DfqClePair;

/* One target layer whose weight enters the convergence metric (Cross_layer_equal.py:107). */
typedef struct DfqCleLayer {
    float *W;
    float *W_old;        /* scratch of the same size                                  */
    int64_t n;
    int32_t first_chunk; /* filled by dfq_cle_layers_plan_host                        */
    int32_t n_chunks;
} DfqCleLayer;

/* Streaming form of the same loop (csrc/cle2.cu): one weight tensor of a chain and one row tile of work. */
typedef struct DfqCleTensor {
    float *W;
    float *W_old;
    int32_t O, row_len;  /* rows and elements per row                                  */
    int32_t KK, I2;      /* taps and input channels per group (as W2 of its up pair)   */
    int32_t Og;          /* output rows per group of the up pair (O when ungrouped)    */
    int32_t up, down;    /* index of the pair where it is W2 / W1, or -1               */
    int32_t layer;       /* slot of the convergence metric                             */
} DfqCleTensor;
typedef struct DfqCleTask {
    int32_t tensor, row0, nrows;
    int32_t kind;        /* low byte 0: column scaling + row ranges, 1: row/column scaling + metric + next column
                            ranges; bits 8..: 64-column strip of a 2-D tile                                  */
} DfqCleTask;

/* Resident form of the same loop (csrc/cle3.cu): the weights of a chain stay in the shared memory of a group of SMs
 * for the whole loop.  One tensor of a chain (tensors of a chain are contiguous, in chain order): */
typedef struct DfqCle3Tensor {
    float *W;            /* [O, L] */
    float *S, *b1, *bn_w, *bn_b;   /* per-row vectors of the pair where it is W1 (NULL without one; bn_* may be NULL) */
    int32_t O, L;        /* rows, elements per row                                      */
    int32_t KK, C;       /* as W2 of its up pair: taps per input channel, number of columns (channels) */
    int32_t dw;          /* 1: depthwise, the row is its own column; 0: one group, column = input channel */
    int32_t up, down;    /* pair where it is W2 / W1, or -1                              */
    int32_t s_off, up_off;   /* channel offsets of the down / up pair in the s table      */
    int32_t cp_off;      /* its column-range partials in the scratch                     */
    int32_t dp_off;      /* its metric partials                                          */
    int32_t G;           /* CTAs of its chain                                            */
    int32_t layer;       /* index into the caller's layer table                          */
} DfqCle3Tensor;
typedef struct DfqCle3Chain { int32_t t0, len, cta0, ncta; } DfqCle3Chain;
typedef struct DfqCle3Plan {
    int32_t n_chains, n_tensors, grid, smem_bytes, total_C, cp_words, dp_doubles, fvec_floats, cp_sm_words, max_rows;
} DfqCle3Plan;

/* Result block of dfq_cle_run (device memory, dfq_cle_result_bytes(max_iters) = 16 + 8*max_iters):
 *   int32 iters; int32 stall; double last_diff; double diffs[max_iters]              */

/* One relation of bias absorption (bias_absorption.py:22-80). */
typedef struct DfqAbsorbPair {
    float *W2;           /* [O2, I2, KK]; clipped in place when clip is requested    */
    float *b1;           /* [C1] */
    float *b2;           /* [O2] */
    const float *bn_w;   /* [C1] fake_weight (gamma) */
    float *bn_b;         /* [C1] fake_bias (beta), updated */
    float *c;            /* [C1] scratch: absorbed vector */
    float *wc;           /* [O2] scratch: W2 . c */
    int32_t C1, O2, I2, KK;
    int32_t first_row;   /* filled by dfq_absorb_plan_host */
} DfqAbsorbPair;

/* One layer of bias correction (bias_correction.py:224-251). */
typedef struct DfqCorrDesc {
    const float *W;      /* [O, I2, KK] the FP weights (epsilon = Q(W) - W is recomputed in registers) */
    const float *v;      /* [I2 * max(groups,1)] per-input-channel vector (E[x]); NULL => 1/I2 (row mean of eps) */
    float *out;          /* [O] sum_i eps[o,i] * v[i]                                 */
    float *bias;         /* [O] optional: bias[o] += alpha * out[o]                   */
    float *fold;         /* [fold_mod] optional, zeroed by the call: fold[(o*I2+i) % fold_mod] += eps[o,i]
                            (the reference body's bias_prev.view(-1, C).mean(0), bias_correction.py:206-213) */
    int32_t O, I2, KK, bits;
    int32_t flags;       /* DFQ_Q_SYMMETRIC */
    float alpha;
    int32_t fold_mod;
    int32_t groups;      /* > 1: grouped conv, v has groups*I2 entries and row o reads v[(o/(O/groups))*I2 + i] */
    int64_t first_unit;  /* filled by dfq_corr_plan_host: first block of the reduction pass */
    int64_t first_block; /* filled by dfq_corr_plan_host: first block of the range pass     */
    int64_t range_slot;  /* filled by dfq_corr_plan_host: (min,max) pair in ranges_dev       */
    int64_t first_stream;/* filled by dfq_corr_plan_host: first 16 KB chunk of the streaming reduction (long
                            pointwise / linear rows); layers that do not qualify own zero chunks there and
                            `first_unit` units of the generic reduction instead                            */
} DfqCorrDesc;

/* One (conv, BatchNorm) pair of merge_batchnorm (utils/layer_transform.py:255-281). */
typedef struct DfqMergeDesc {
    float *W;            /* [O, K] layer weight, scaled in place                       */
    float *b;            /* [O] layer bias (never NULL: the reference adds zeros, :262-263) */
    float *gamma, *beta, *mean, *var;   /* [O] BatchNorm tensors; reset to 1, 0, 0, 1 when reset_bn != 0 (:277-280) */
    float *fake_weight;  /* [O] out: |gamma| (:273) */
    float *fake_bias;    /* [O] out: beta    (:274) */
    int64_t O, K;
    float eps;
    int32_t first_row;   /* filled by dfq_merge_plan_host */
} DfqMergeDesc;

/* One tensor of a multi-tensor clamp (clip_weight.py:27-29). */
typedef struct DfqClipDesc {
    float *w;
    int64_t n;
    int64_t first_block; /* filled by dfq_clip_plan_host */
} DfqClipDesc;

/* One tensor of the multi-tensor dequantisation (8-bit codes of an asymmetric grid -> fp32, the two roundings of
 * utils/quantize.py:76).  codes / slot may live in a peer GPU's memory (P2P). */
typedef struct DfqDequantDesc {
    const void *codes;   /* uint8 [n]                                                  */
    float *y;            /* fp32 [n]                                                   */
    int64_t n;
    const float *slot;   /* {min, max, scale, 0}                                       */
    int64_t first_block; /* filled by dfq_dequant_plan_host                            */
} DfqDequantDesc;

int dfq_abi_version(void);
const char *dfq_error_string(int err);
/* number of kernels launched by this library since load (bench.py's gpu_launches) */
int64_t dfq_launch_count(void);

/* ---- (a) uniform fake-quantisation: utils/quantize.py:24-78 ------------------------ */
/* host helper: the five fp32 constants from Python-float min/max (utils/quantize.py:51-68) */
void dfq_quant_params_host(double min_value, double max_value, int bits, int flags, DfqQuantParams *out);
/* global min/max of x -> out[0]=min, out[1]=max (float(W.min()), float(W.max())) */
int dfq_minmax_f32(const float *x, int64_t n, float *out_minmax, void *stream);
/* per-row min/max of x.view(rows, cols) -> out[2*r], out[2*r+1]
 * (QuantMeasure's per-sample extrema, utils/quantize.py:110-111; per-channel ranges) */
int dfq_minmax_rows_f32(const float *x, int64_t rows, int64_t cols, float *out_minmax, void *stream);
/* elementwise pass with host-supplied constants (QuantMeasure.forward, utils/quantize.py:123) */
int dfq_fake_quant_apply_f32(const float *x, float *y, void *codes, int64_t n, int bits,
                             const DfqQuantParams *params_host, void *stream);
/* plan + launch of the multi-tensor path (quantize_targ_layer, utils/layer_transform.py:288-305):
 *   descs_host[T] is filled in (first_block, range_slot) and the totals returned;
 *   the caller copies descs to the device and provides ranges_dev[4*total_slots] floats, which come back as
 *   {min, max, scale, 0} per slot (per tensor, or per row for DFQ_Q_PER_CHANNEL). */
int dfq_quant_plan_host(DfqQuantDesc *descs_host, int n_desc, int64_t *total_blocks, int64_t *total_slots);
int dfq_fake_quant_multi_f32(const DfqQuantDesc *descs_dev, int n_desc, int64_t total_blocks,
                             int64_t total_slots, float *ranges_dev, void *stream);
/* Inverse of the optional code output: y = code * scale + min with the fake-quant chain's own two roundings
 * (utils/quantize.py:76), slots_dev = {min, max, scale, 0} per tensor (rows == 1) or per row.  Used to rebuild the
 * quantised state dict from the all-gathered codes of the layer-sharded pass. */
int dfq_dequant_codes_f32(const void *codes, int code_bytes, float *y, int64_t n, int64_t rows,
                          const float *slots_dev, void *stream);

/* ---- QuantMeasure with the range resident on the device: utils/quantize.py:94-126 (no host round trip) -----
 * running_min / running_max: the module's [1] buffers (device), updated in place when update_stat != 0 with
 *   running_max = max(running_max, mean_n max_chw x), running_min = min(...)  (:110-111, Python max/min semantics);
 * slot_dev: 4 floats {min, max, scale, 0} written for dfq_fake_quant_apply_slot_f32;
 * stats_scratch: 2*n_samples floats (per-sample extrema), needed when update_stat != 0. */
int dfq_measure_update(const float *x, int64_t n_samples, int64_t chw, float *running_min, float *running_max,
                       float *slot_dev, int bits, int flags, int update_stat, float *stats_scratch, void *stream);
int dfq_fake_quant_apply_slot_f32(const float *x, float *y, void *codes, int64_t n, int bits, int flags,
                                  const float *slot_dev, void *stream);

/* dfq_measure_update without the stats_scratch argument and without any allocation.  No counterpart in the reference:
 * added for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops).  The per-sample extrema go to a file-scope device array of the library
 * (2 * DFQ_MEASURE_NOALLOC_MAX_SAMPLES floats = 32 KB, held by every GPU that loads the library until exit); the call
 * makes no cudaMalloc / cudaFree and calls no allocating routine.  n_samples > DFQ_MEASURE_NOALLOC_MAX_SAMPLES (with
 * update_stat != 0) is rejected with cudaErrorInvalidValue and a message on stderr.  Results are those of
 * dfq_measure_update bit for bit.  All calls share the array: calls that overlap on the GPU (different streams) race;
 * calls on one stream are ordered.  The array keeps the last call's contents; each call clears what it uses. */
#define DFQ_MEASURE_NOALLOC_MAX_SAMPLES 4096
int dfq_measure_update_noalloc(const float *x, int64_t n_samples, int64_t chw, float *running_min, float *running_max,
                               float *slot_dev, int bits, int flags, int update_stat, void *stream);

/* the same from already reduced per-sample statistics: stats = [n_samples, 2] floats (encoded == 0), or the
 * order-preserving integer encoding {enc(max(-y)), enc(max(y))} a producing dfq_q* kernel accumulated in its
 * epilogue (encoded != 0; the table must be zero before the producer runs). */
int dfq_measure_from_stats(const void *stats, int encoded, int64_t n_samples, float *running_min, float *running_max,
                           float *slot_dev, int bits, int flags, int update_stat, void *stream);

/* ---- (e) quantised eval forward, CUDA-core kernels: utils/quantize.py:225-256, 336-363; residual add and mean of
 * utils/layer_transform.py:18-32, 79-94.  Every kernel fake-quantises its input on load with the range in
 * in_slot (4 floats {min, max, scale, 0} on the device; NULL = input used as is), fuses bias + ReLU, and, when
 * out_stats != NULL, accumulates the per-sample min/max of its output (encoded, see dfq_measure_from_stats). */
typedef struct DfqConvDesc {
    const float *x;          /* [N, C_in, H, W]                                         */
    float *y;                /* [N, C_out, OH, OW]                                      */
    const float *w;          /* [C_out, C_in/groups, kh, kw]                            */
    const float *bias;       /* [C_out] or NULL                                         */
    const float *in_slot;    /* QuantMeasure range of the input or NULL                 */
    unsigned *out_stats;     /* [N, 2] or NULL                                          */
    int32_t N, C_in, H, W, C_out, OH, OW, kh, kw, stride, pad, dil, groups, relu, bits;
    int32_t reserved;
} DfqConvDesc;
int dfq_qconv2d_f32(const DfqConvDesc *desc_host, void *stream);
int dfq_qlinear_f32(const float *x, float *y, const float *w, const float *bias, int64_t N, int K, int O,
                    const float *in_slot, int bits, int relu, unsigned *out_stats, void *stream);
int dfq_qadd_f32(const float *a, const float *b, float *out, int64_t N, int64_t chw, const float *slot_a,
                 const float *slot_b, int bits_a, int bits_b, unsigned *out_stats, void *stream);
int dfq_qmean_f32(const float *x, float *out, int64_t N, int C, int HW, const float *slot, int bits,
                  unsigned *out_stats, void *stream);

/* Data-movement ops of the same forward, each ONE pass with the operand fake-quantised on load (in_slot, NULL = as
 * is) and the per-sample min/max of the result accumulated into out_stats (encoded, may be NULL):
 *   dfq_qinterp_bilinear_f32  F.interpolate(Q(x), size=(OH, OW), mode='bilinear', align_corners=...) as patched by
 *                             utils/layer_transform.py:98-110 (modeling/segmentation/deeplab.py:32, decoder.py:39,
 *                             aspp.py:71); source index and weights formed as ATen forms them, fp32
 *   dfq_qcopy_f32             y[n * dst_stride + k] = Q(x[n * src_stride + k]), k < count: one operand of
 *                             torch.cat(dim=1) as patched by utils/layer_transform.py:113-125 (y points at the
 *                             operand's channel offset inside the destination)
 *   dfq_pad2d_f32             F.pad(x, (left, right, top, bottom)), constant 0, never quantised
 *                             (utils/layer_transform.py ignore_op=['pad']; modeling/segmentation/backbone/mobilenet.py fixed_padding)
 * Per-sample element counts must stay below 2^31. */
int dfq_qinterp_bilinear_f32(const float *x, float *y, int64_t N, int C, int H, int W, int OH, int OW,
                             int align_corners, const float *in_slot, int bits, unsigned *out_stats, void *stream);
int dfq_qcopy_f32(const float *x, float *y, int64_t N, int64_t count, int64_t src_stride, int64_t dst_stride,
                  const float *in_slot, int bits, unsigned *out_stats, void *stream);
int dfq_pad2d_f32(const float *x, float *y, int64_t N, int C, int H, int W, int left, int right, int top, int bottom,
                  unsigned *out_stats, void *stream);

/* Tensor-core form of a pointwise (1x1, stride 1) convolution or a linear layer of the same forward
 * (utils/quantize.py:250-256, 358-363): s8 x s8 -> s32 tcgen05.mma.kind::i8, accumulator in TMEM.
 * Needs the weights on an integer grid of <= 8 bits: w[o,k] = s_w * code + min_w (what quantize_targ_layer
 * leaves, or what QuantConv2d re-derives every forward).  A "pixel" g = n * P + p is one GEMM row; element
 * (g, k) of the input is x[n*xs_n + k*xs_k + p*xs_p], output (g, o) is y[n*ys_n + o*ys_o + p*ys_p] (conv:
 * P = H*W, xs_k = ys_o = P, xs_p = ys_p = 1; linear: P = 1, xs_n = K, xs_k = 1, ys_n = O, ys_o = 1). */
typedef struct DfqPwI8Desc {
    const float *x;
    float *y;
    const int8_t *wpk;       /* device copy of dfq_pw_i8_pack_host's output (code - 128, slab-major) */
    const float *wr;         /* [O] s_w * sum_k (code[o,k] - 128)                              */
    const float *bias;       /* [O] or NULL                                                    */
    const float *in_slot;    /* QuantMeasure range of the input {min, max, scale, 0}; required */
    unsigned *out_stats;     /* [samples, 2] encoded per-sample extrema of y, or NULL          */
    int64_t G;               /* pixels in total                                                */
    int64_t xs_n, xs_k, xs_p, ys_n, ys_o, ys_p;   /* element strides                          */
    int32_t P, K, O, bits, relu;
    float s_w, c_w;          /* weight scale and centred offset min_w + 128 * s_w             */
    int32_t o_cap;           /* the o_cap wpk was packed with (0 = 256 channels per pass)      */
} DfqPwI8Desc;
int dfq_pw_i8_plan_host(int O, int K, int *n_pass, int *o_per_pass, int *npass, int *slabs, int64_t *packed_bytes);
int dfq_pw_i8_pack_host(const int8_t *codes_host, int O, int K, int8_t *packed_host);
/* The same with at most o_cap (16..256, 0 = 256) output channels per pass: a pass is one CTA's accumulator, so a
 * narrower pass spreads a layer with few pixel tiles over more SMs (the 1280 -> 1000 classifier at batch 256 is
 * 2 tiles: 8 CTAs at 256 channels per pass, 126 at 16).  Desc.o_cap must repeat the value used for packing. */
int dfq_pw_i8_plan_cap_host(int O, int K, int o_cap, int *n_pass, int *o_per_pass, int *npass, int *slabs,
                            int64_t *packed_bytes);
int dfq_pw_i8_pack_cap_host(const int8_t *codes_host, int O, int K, int o_cap, int8_t *packed_host);
int dfq_pw_i8(const DfqPwI8Desc *desc_host, void *stream);

/* Dense k x k convolution of the same forward (groups == 1; DeepLab's atrous 3x3 and decoder convolutions,
 * utils/quantize.py:250-256 with modeling/segmentation/aspp.py:50-53, decoder.py:22-30) as an implicit GEMM on the
 * same tcgen05 pipeline: x [N,C,H,W] fp32, y [N,O,OH,OW] fp32, zero padding `pad`, any stride / dilation.
 * Reduction order: tap-major, channels padded to a multiple of 16 -- dfq_conv_i8_k_host gives its length K;
 * lay the centred codes out as [O][kh*kw][C padded with zeros] and pack them with dfq_pw_i8_pack_host(codes, O, K).
 * wr as in DfqPwI8Desc; wrt[t][o] = s_w * (sum over channels of the centred codes of tap t): what a tap that falls
 * into the zero padding must not contribute. */
typedef struct DfqConvI8Desc {
    const float *x;
    float *y;
    const int8_t *wpk;
    const float *wr;         /* [O]                                                             */
    const float *wrt;        /* [kh*kw][O]                                                      */
    const float *bias;       /* [O] or NULL                                                     */
    const float *in_slot;    /* QuantMeasure range of the input {min, max, scale, 0}; required  */
    unsigned *out_stats;     /* [N, 2] encoded per-sample extrema of y, or NULL                 */
    int32_t N, C, H, W, O, OH, OW, kh, kw, stride, pad, dil, bits, relu;
    float s_w, c_w;
    int8_t *scratch;         /* NULL, or dfq_conv_i8_scratch_bytes(N, C, H, W) device bytes: the input is then
                              * quantised ONCE into s8 NHWC codes and every tap reads 16-byte words (k x k > 1) */
    int32_t o_cap;           /* the o_cap wpk was packed with (0 = 256 channels per pass)       */
    int32_t halo;            /* 1 (3 x 3, stride 1, pad 1, dilation 1, 8-bit activations, scratch given): wpk is packed
                              * over the [tap][chunk] reduction of dfq_conv_q8_halo_k_host and the halo kernel runs (the
                              * tile's input halo in shared memory, taps by shifted descriptors: see DfqConvQ8Desc.tma) */
} DfqConvI8Desc;
int dfq_conv_i8_k_host(int C, int kh, int kw);
int64_t dfq_conv_i8_scratch_bytes(int N, int C, int H, int W);
int dfq_conv_i8(const DfqConvI8Desc *desc_host, void *stream);

/* The dense k x k convolution of the STATIC-range plan (set_update_stat(False), utils/quantize.py:119-126; the
 * activation format of dfq_pw_q8 below): x is centred s8 codes [N, H, W, C padded to 16] on the grid in_slot, the output
 * is written as codes [N, OH, OW, O padded to 16] on the grid(s) of its consumer(s) -- no quantising pre-pass, no fp32
 * activation in HBM.  wpk / wr / wrt as for dfq_conv_i8 (same packing). */
typedef struct DfqConvQ8Desc {
    const int8_t *x;
    const int8_t *wpk;
    const float *wr;         /* [O]                                                             */
    const float *wrt;        /* [kh*kw][O]                                                      */
    const float *bias;       /* [O] or NULL                                                     */
    const float *in_slot;    /* {min, max, scale, 0} of the input's grid                        */
    int8_t *y0;              /* codes on out_slot0's grid                                       */
    const float *out_slot0;
    int8_t *y1;              /* NULL, or the same values on a second consumer's grid            */
    const float *out_slot1;
    int32_t N, C, H, W, O, OH, OW, kh, kw, stride, pad, dil, relu;
    float s_w, c_w;
    int32_t o_cap;
    int32_t tma;             /* 0: wpk packed by dfq_pw_i8_pack_cap_host (register-gather kernel, any stride);
                              * 1: wpk packed by dfq_conv_q8_pack_host (stride 1): the activation tile of every tap arrives
                              *    by a 4-D tensor-map TMA box of the code tensor, warp-specialised tcgen05 pipeline;
                              * 2: 3 x 3, stride 1, dilation 1, padding 1 only: the 18 x 10 input halo of a 16 x 8 pixel
                              *    tile is loaded once by TMA and all nine taps read it through shifted UMMA descriptors
                              *    (no im2col gather).  wpk packed by dfq_pw_i8_pack_cap_host over a reduction ordered
                              *    [tap][16-byte channel chunk] with the chunks of a tap zero-padded to a multiple of 4
                              *    (dfq_conv_q8_halo_k_host(C) = its length)                                             */
} DfqConvQ8Desc;
int dfq_conv_q8(const DfqConvQ8Desc *desc_host, void *stream);
/* weight layout of the tma form: reduction tap-major, then 128-byte channel slabs (zero padded); codes_host is
 * [O][kh*kw][C padded to 16] centred codes; *kq = padded reduction length, o_cap as for dfq_pw_q8 (0 = 240 per pass) */
int dfq_conv_q8_halo_k_host(int C);
int dfq_conv_q8_plan_host(int O, int C, int taps, int o_cap, int *kq, int64_t *packed_bytes);
int dfq_conv_q8_pack_host(const int8_t *codes_host, int O, int C, int taps, int o_cap, int8_t *packed_host);

/* ---- (e) static-range W8A8 eval (csrc/q8.cu): QuantMeasure with update_stat == False, utils/quantize.py:119-126 ----
 * Activations are CENTRED s8 CODES (code - 128 of utils/quantize.py:72-74 on the CONSUMER's grid), NHWC, channels
 * padded to a multiple of 16 with zero bytes; a range slot is {min, max, scale, 0} as fp32.  Every producer writes
 * the codes for up to two consumers (y0 / y1 with their slots; y1 may be NULL). */
/* Where a producer of the static-range plan puts the pixels of one of its outputs.  All zero: row g of the [pixels][Cp]
 * code tensor.  Otherwise the output is the INTERIOR of a tensor padded by F.pad(x, (pl, pr, pt, pb)) ahead of a
 * QuantMeasure (DeepLab's fixed_padding): pixel (n, oh, ow) of the OH x OW image goes to row
 * (n (OH+pt+pb) + oh + pt) (OW+pl+pr) + ow + pl; the border holds code(0.0) on that grid, written once by the plan. */
typedef struct DfqQ8OutMap { int32_t OH, OW, pl, pr, pt, pb; } DfqQ8OutMap;

typedef struct DfqPwQ8Desc {
    const int8_t *x;         /* [G][Kp] codes on this layer's grid (Kp = K rounded up to 16)            */
    const int8_t *wpk;       /* dfq_pw_q8_pack_host                                                      */
    const float *wr;         /* [O] s_w * sum_k (code_w - 128)                                           */
    const float *bias;       /* [O] or NULL                                                              */
    const float *in_slot;    /* this layer's QuantMeasure range                                          */
    int8_t *y0; const float *out_slot0;
    int8_t *y1; const float *out_slot1;
    float *yf;               /* optional fp32 output y[g * yf_g + o * yf_o] (the classifier's logits)    */
    int64_t yf_g, yf_o;
    int64_t G;               /* pixels (images x H x W) or rows of a linear layer                        */
    int32_t K, O, relu, o_cap;
    float s_w, c_w;          /* w_hat = s_w * (code_w - 128) + c_w                                       */
    DfqQ8OutMap map0, map1;  /* pixel placement of y0 / y1 (all zero = plain)                            */
} DfqPwQ8Desc;
int dfq_pw_q8_plan_host(int O, int K, int o_cap, int *n_pass, int *o_per_pass, int *npass, int *slabs,
                        int64_t *packed_bytes);
int dfq_pw_q8_pack_host(const int8_t *codes_host, int O, int K, int o_cap, int8_t *packed_host);
/* 1 x 1 convolution / linear layer: tensor-map TMA feed, tcgen05.mma.kind::i8, requantising epilogue. */
int dfq_pw_q8(const DfqPwQ8Desc *desc_host, void *stream);

typedef struct DfqDwQ8Desc {
    const int8_t *x;         /* [N][H][W][Cp]                                                            */
    const uint32_t *wrow;    /* [3][Cp] words (w(r,0), w(r,1), w(r,2), 0): centred s8 weight codes of kernel row r   */
    const float *bias;       /* [Cp] or NULL                                                             */
    const float *in_slot;
    int8_t *y0; const float *out_slot0;
    int8_t *y1; const float *out_slot1;
    int32_t N, H, W, C, stride, relu;
    float s_w, c_w;          /* w_hat = s_w * (code_w - 128) + c_w                                       */
    int32_t pad, dil;        /* zero padding per side (0, 1, ...) and dilation (0 is read as 1)           */
} DfqDwQ8Desc;
/* depthwise 3 x 3, stride 1 or 2; the dp4a kernel for padding 0 / 1 with dilation 1, a generic one otherwise */
int dfq_dw_q8(const DfqDwQ8Desc *desc_host, void *stream);

/* ---- data-movement ops of the static-range plan on code tensors (DeepLabV3+: utils/layer_transform.py:98-125) ----
 * F.interpolate(Q(x), (OH, OW), 'bilinear', align_corners): x codes [N,H,W,Cp] on in_slot's grid; the result leaves as
 * codes on the consumer grids (y0 / y1) and / or as fp32 NCHW (yf: the network's logits). */
int dfq_interp_q8(const int8_t *x, const float *in_slot, int N, int C, int H, int W, int OH, int OW, int align_corners,
                  int8_t *y0, const float *out_slot0, int8_t *y1, const float *out_slot1, float *yf, void *stream);
/* One operand of torch.cat(dim=1): G pixels x C channels of codes on in_slot's grid (its own QuantMeasure) -> the values
 * -> codes on the grid(s) of the concatenation's consumer, at channel offset c_off (multiple of 4) of rows cp_out
 * channels wide.  fill_tail != 0 (last operand): the row's remaining padding channels are cleared too. */
int dfq_regrid_q8(const int8_t *x, const float *in_slot, int64_t G, int C, int8_t *y0, const float *out_slot0,
                  int8_t *y1, const float *out_slot1, int cp_out, int c_off, int fill_tail, void *stream);
/* F.pad(x, (pl, pr, pt, pb)) ahead of a QuantMeasure (DeepLab's fixed_padding; utils/layer_transform.py ignore_op=['pad']:
 * the pad itself is not quantised, the consumer's measure quantises the padded tensor): x codes [N,H,W,Cp] on the
 * consumer's grid `slot` -> y [N,H+pt+pb,W+pl+pr,Cp], border = the code of 0.0 on that grid. */
int dfq_pad_q8(const int8_t *x, const float *slot, int N, int C, int H, int W, int pl, int pr, int pt, int pb, int8_t *y,
               void *stream);
/* nn.AdaptiveAvgPool2d(1) of the RAW fp32 tensor (not a patched op: modeling/segmentation/aspp.py:61-64), x [N][HW] rows
 * of C floats `ld` floats apart (the fp32 side output of dfq_pw_q8) -> codes [N][Cp] on the consumer grids. */
int dfq_pool_q8(const float *x, int64_t ld, int N, int HW, int C, int8_t *y0, const float *out_slot0, int8_t *y1,
                const float *out_slot1, void *stream);

typedef struct DfqStemQ8Desc {
    const float *x;          /* [N][3][H][W] fp32 image                                                  */
    const float *w;          /* [O][3][3][3]                                                             */
    const float *bias;
    const float *in_slot;
    int8_t *y0; const float *out_slot0;
    int8_t *y1; const float *out_slot1;
    int32_t N, H, W, O, stride, pad, relu, reserved;
    DfqQ8OutMap map0, map1;
    const int32_t *wq;       /* optional: [32][8] words of centred s8 weight codes, word k = taps 4k .. 4k + 3 in     */
    float s_w, c_w;          /* (c, kh, kw) order, rows >= O and taps >= 27 zero; w = s_w * code + c_w.  Non-NULL: the */
                             /* exact-integer (dp4a) kernel, else fp32 FMAs on the dequantised inputs                 */
    float *yf;               /* optional (needs wq; y0 / y1 NULL): the dynamic-range plan's form -- fp32 NCHW output   */
    uint32_t *out_stats;     /* and the encoded per-sample extrema [N][2] of it (as DfqConvDesc.out_stats), or NULL    */
} DfqStemQ8Desc;
/* dense 3 x 3 stem on 3 input channels, O <= 32 */
int dfq_stem_q8(const DfqStemQ8Desc *desc_host, void *stream);
/* residual add of two code tensors [pixels][Cp] on their own grids (utils/layer_transform.py:18-32) */
int dfq_add_q8(const int8_t *xa, const int8_t *xb, const float *slot_a, const float *slot_b, int8_t *y0,
               const float *out_slot0, int8_t *y1, const float *out_slot1, int64_t pixels, int C,
               const DfqQ8OutMap *map0, const DfqQ8OutMap *map1 /* host pointers, NULL = plain */, void *stream);
/* spatial mean of Q(x): codes [N][HW][Cp] -> codes [N][Cp] on out_slot's grid (y) and / or fp32 [N][C] (yf) */
int dfq_mean_q8(const int8_t *x, const float *in_slot, int8_t *y, const float *out_slot, float *yf, int64_t N, int C,
                int HW, void *stream);

/* ---- (b) cross-layer equalisation: Cross_layer_equal.py:63-115 ---------------------- */
/* sorts pairs by stage (stable) and fills first_unit/n_units */
int dfq_cle_plan_host(DfqClePair *pairs_host, int n_pairs, int *n_stages, int *total_units);
int dfq_cle_layers_plan_host(DfqCleLayer *layers_host, int n_layers, int *total_chunks);
int64_t dfq_cle_result_bytes(int max_iters);
/* Runs the whole `while diff > threshold and stall < count` loop in ONE cooperative launch.
 * pairs must be sorted by stage (dfq_cle_plan_host does it).  result_dev: see above.
 * smin/smax: s_min_max; is_signed: :34-35; eps: :42.  max_iters bounds the loop (the
 * reference is unbounded).  scratch_dev: 2*n_layers doubles, zero-initialised by the call. */
int dfq_cle_run(const DfqClePair *pairs_dev, int n_pairs, int n_stages, const DfqCleLayer *layers_dev,
                int n_layers, int total_chunks, double smin, double smax, int is_signed, float eps,
                double threshold, int count, int max_iters, double *scratch_dev, void *result_dev,
                void *stream);
/* Streaming form: every tensor is read twice per iteration in coalesced row tiles, one grid barrier per chain
 * position.  dfq_cle2_plan_host returns cudaErrorNotSupported (801) for relation sets that are not simple chains;
 * use dfq_cle_run then.  tensors_host: n_layers entries; tasks_host: max_tasks entries (n_tasks phase tasks followed
 * by n_init prologue tasks); phase_begin_host: 65 ints; chan_off_host: n_pairs + 1 ints.  r1_dev: 2 * total_C floats,
 * r2_dev: 4 * total_C uints, acc_dev: 2 * n_layers doubles (all zeroed by the call).  debug_dev: NULL, or
 * blocks * n_phases * 5 uint64 for per-phase timer stamps (tuning aid). */
int dfq_cle2_plan_host(const DfqClePair *pairs_host, int n_pairs, const DfqCleLayer *layers_host, int n_layers,
                       DfqCleTensor *tensors_host, DfqCleTask *tasks_host, int max_tasks, int *n_tasks, int *n_init,
                       int32_t *phase_begin_host, int *n_phases, int32_t *chan_off_host);
int dfq_cle2_run(const DfqClePair *pairs_dev, int n_pairs, const DfqCleTensor *tensors_dev, int n_tensors,
                 const DfqCleTask *tasks_dev, int n_tasks, int n_init, const int32_t *phase_begin_host, int n_phases,
                 const int32_t *chan_off_dev, int total_C, float *r1_dev, unsigned *r2_dev, double smin, double smax,
                 int is_signed, float eps, double threshold, int count, int max_iters, double *acc_dev,
                 void *result_dev, int blocks_per_sm, void *fat_dev, void *debug_dev, void *stream);
/* bytes of fat_dev (16-byte aligned scratch for the expanded task records) for n_tasks + n_init tasks */
int64_t dfq_cle2_fat_bytes(int n_tasks_total);

/* Resident form (csrc/cle3.cu), tried first: every chain gets a group of SMs sized by its element count, a CTA keeps
 * whole rows of its chain's tensors in shared memory for the whole loop, groups synchronise among themselves once
 * per chain position and the grid once per iteration.  dfq_cle3_plan_host returns cudaErrorNotSupported (801) when
 * the relations are not simple chains of ungrouped / depthwise layers or the chains do not fit the chip's shared
 * memory (use dfq_cle2_run then).  n_sm / smem_limit: 0 = query the current device.  tensors_host: n_layers entries,
 * chains_host: n_layers entries (n_chains used).  scratch_dev: dfq_cle3_scratch_bytes(plan), zeroed by the call.
 * result_dev as for dfq_cle_run.  debug_dev: NULL, or grid * 33 * 8 uint64 for globaltimer stamps (tuning aid). */
int dfq_cle3_plan_host(const DfqClePair *pairs_host, int n_pairs, const DfqCleLayer *layers_host, int n_layers,
                       int n_sm, int smem_limit, DfqCle3Tensor *tensors_host, DfqCle3Chain *chains_host,
                       DfqCle3Plan *plan);
int64_t dfq_cle3_scratch_bytes(const DfqCle3Plan *plan);
int dfq_cle3_run(const DfqCle3Tensor *tensors_dev, const DfqCle3Chain *chains_dev, const DfqCle3Plan *plan_host,
                 void *scratch_dev, double smin, double smax, int is_signed, float eps, double threshold, int count,
                 int max_iters, void *result_dev, void *debug_dev, void *stream);

/* all tensors of a table in ONE launch (the receiving side of the layer-sharded pass, dist.py) */
int dfq_dequant_plan_host(DfqDequantDesc *descs_host, int n_desc, int64_t *total_blocks);
int dfq_dequant_multi_f32(const DfqDequantDesc *descs_dev, int n_desc, int64_t total_blocks, void *stream);

/* ---- (c) bias absorption (+ optional weight clip): bias_absorption.py:9-80, clip_weight.py:4-33 */
int dfq_absorb_plan_host(DfqAbsorbPair *pairs_host, int n_pairs, int *total_rows);
/* max_dim = max over pairs of max(C1, O2) (sizes the per-element bias pass) */
int dfq_absorb_clip(const DfqAbsorbPair *pairs_dev, int n_pairs, int total_rows, int max_dim, float N,
                    int do_clip, float clip_lo, float clip_hi, void *stream);
int dfq_clip_f32(float *w, int64_t n, float lo, float hi, void *stream);
/* every target layer of a model in one launch */
int dfq_clip_plan_host(DfqClipDesc *descs_host, int n_desc, int64_t *total_blocks);
int dfq_clip_multi_f32(const DfqClipDesc *descs_dev, int n_desc, int64_t total_blocks, float lo, float hi,
                       void *stream);

/* ---- (d) bias correction: bias_correction.py:111-144, 224-251 ---------------------- */
int dfq_corr_plan_host(DfqCorrDesc *descs_host, int n_desc, int64_t *total_units, int64_t *total_blocks,
                       int64_t *total_stream);
/* ranges_dev: 4*n_desc floats, {min, max, scale, 0} per layer.  range_given != 0: the caller filled min and max
 * (4 B/elt); otherwise the call reduces them first (8 B/elt).  `out` must be non-NULL.      */
int dfq_bias_correct(const DfqCorrDesc *descs_dev, int n_desc, int64_t total_units, int64_t total_blocks,
                     int64_t total_stream, float *ranges_dev, int range_given, void *stream);

/* ---- BN folding: utils/layer_transform.py:240-285 ----------------------------------- */
int dfq_merge_bn(float *W, float *b, int64_t O, int64_t K, const float *gamma, const float *beta,
                 const float *mean, const float *var, float eps, float *fake_weight, float *fake_bias,
                 void *stream);
/* every (conv, BN) pair of a model in one launch; reset_bn also turns each BN into the identity */
int dfq_merge_plan_host(DfqMergeDesc *descs_host, int n_desc, int *total_rows);
int dfq_merge_bn_multi(const DfqMergeDesc *descs_dev, int n_desc, int total_rows, int reset_bn, void *stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif

#ifdef __cplusplus
}
#endif
#endif /* DFQ_B200_H */
