// Kernel (e), tensor-core part: pointwise (1x1) convolutions and linear layers of the quantised eval forward
// (utils/quantize.py:225-256, 336-363) as s8 x s8 -> s32 tcgen05.mma.kind::i8 with the accumulator in TMEM.
//
//   x_hat[p,k] = s_x * a[p,k] + c_x      a = code(x) - 128  (code = the integer of utils/quantize.py:74, computed
//   w_hat[o,k] = s_w * b[o,k] + c_w      b = code(w) - 128   on load with the consumer's QuantMeasure range;
//                                        grids of fewer than 8 bits are centred and scaled up to the s8 range)
//   y[p,o] = s_x s_w  SUM_k a b  +  s_x c_w SUM_k a[p,k]  +  c_x s_w SUM_k b[o,k]  +  K c_x c_w  +  bias[o]
//
// so the K-long dot products are exact 32-bit integer sums on the tensor pipe and the fp32 work is one FMA chain
// per output.  One CTA = 128 "pixels" (GEMM M, TMEM lanes) x one pass of <= 256 output channels (GEMM N, TMEM
// columns), K walked in 64-byte slabs through a 2-deep shared-memory ring:
//   * the weight codes of a slab arrive by the bulk async-copy engine (cp.async.bulk, TMA without a tensor map:
//     dfq_pw_i8_pack_host stores every slab already in the UMMA canonical K-major order) on an mbarrier;
//   * the activation slab is read as fp32 (coalesced along the pixel axis), quantised in registers, its code sum
//     accumulated with dp4a, and written to shared memory in the same canonical order with 16-byte stores;
//   * one thread issues the MMAs of the slab and commits them to the ring's "free" mbarrier.
// Epilogue: tcgen05.ld 32x32b (lane = pixel), the FMA chain above, bias, ReLU, per-sample min/max of the output
// for the next QuantMeasure (utils/quantize.py:110-111), coalesced fp32 stores.
#include <cstdlib>
#include "common.cuh"
#include "q8_pack.cuh"

namespace dfq {

constexpr int kI8M = 128;            // pixels per CTA (UMMA M)
constexpr int kI8Slab = 64;          // K bytes per ring stage (2 MMAs of K = 32)
constexpr int kI8ASlab = kI8M * kI8Slab;
constexpr int kI8Spin = 1 << 22;     // bounded mbarrier spin: a protocol bug traps instead of hanging the GPU

struct PwI8Args {
    const float *x; float *y; const int8_t *wpk; const float *wr; const float *bias; const RangeSlot *slot;
    unsigned *stats;
    int64_t G;                        // pixels in total (images x pixels per image)
    int64_t xs_n, xs_k, xs_p, ys_n, ys_o, ys_p;
    int P, K, O, bits, relu;
    int n_pass, o_per_pass, npass, slabs, tmem_cols;
    float s_w, c_w;
    // dense k x k form (kConv): K above is the PADDED reduction length taps * cchunks * 16, Kreal = taps * C
    const float *wrt;                 // [taps][O] s_w * (sum over channels of the centred codes of one tap)
    const int8_t *xq;                 // kQ: centred s8 codes [N][H][W][cchunks * 16]
    int C, H, W, OW, kh, kw, stride, pad, dil, taps, cchunks, Kreal;
    int64_t tiles;                    // warp-specialised form: 128-pixel tiles (work items = tiles * n_pass)
    // kOutQ (static-range plan, eval_q8.py): the output leaves as centred s8 codes, NHWC with cp_out channels per
    // pixel, on the grid(s) of the consumer(s) instead of fp32 NCHW
    int8_t *yq[2]; const RangeSlot *oslot[2]; int cp_out;
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok = 0;
    for (int spin = 0; spin < kI8Spin; ++spin) {
        asm volatile("{\n\t.reg .pred p;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) return;
    }
    __trap();
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// shared-memory matrix descriptor, no swizzle, K-major: 8-row x 16-byte core matrices (128 contiguous bytes);
// LBO = byte step between the two 16-byte K chunks of one MMA, SBO = byte step between 8-row groups
__device__ __forceinline__ uint64_t umma_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
           (1ull << 46);
}
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                        uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// The integer code of utils/quantize.py:72-74 in the low byte of the result.
// quant_code() of common.cuh with the magic-number add kept: 1.5 * 2^23 + q has q in its mantissa's low byte, so no
// float-to-int conversion is needed.
__device__ __forceinline__ uint32_t quant_code_bits(float x, const QuantConsts &c) {
    const float t = __fadd_rn(x, c.neg_min);
    if (c.fast) {                                           // block-uniform
        const float magic = 12582912.0f;
        const float cl = fminf(fmaxf(__fmul_rn(t, c.rscale), c.qmin), c.qmax);
        float m = __fadd_rn(cl, magic);
        if (fabsf(__fsub_rn(cl, __fadd_rn(m, -magic))) > c.amb) {          // within a few ulp of a tie: exact division
            const float tc = fminf(fmaxf(t, c.t_lo), c.t_hi);
            m = __fadd_rn(fminf(fmaxf(div_const_rn(tc, c.scale, c.rscale), c.qmin), c.qmax), magic);
        }
        return __float_as_uint(m);
    }
    return (uint32_t)__float2int_rn(fminf(fmaxf(__fdiv_rn(t, c.scale), c.qmin), c.qmax));
}

// fp32 NCHW activation -> centred s8 codes, NHWC with the channels padded to a multiple of 16 (pad bytes = s8 zero).
// A block owns 32 consecutive pixels of one image x 64 channels: coalesced fp32 reads along the pixels, transposed
// through shared memory, 16-byte stores along the channels.
__global__ void __launch_bounds__(kThreads)
quant_nhwc_s8_kernel(const float *__restrict__ x, int8_t *__restrict__ q, const RangeSlot *slot, int bits, int C,
                     int HW, int Cp) {
    // 64 pixels x 64 channels per block: 16 coalesced fp32 loads in flight per thread, one 16-byte store per thread
    __shared__ __align__(16) uint8_t tile[64][64 + 16];
    const int tid = threadIdx.x, px = tid & 63, cl = tid >> 6;
    const int p0 = blockIdx.x * 64, c0 = blockIdx.y * 64, n = blockIdx.z;
    const QuantConsts qc = consts_from_slot(*slot, bits, 0, 0.f, 0.f);
    const int sh = 8 - bits;
    const int p = p0 + px;
    const float *xn = x + ((int64_t)n * C) * HW;
    float v[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const int c = c0 + cl + 4 * i;
        v[i] = (p < HW && c < C) ? __ldg(xn + (int64_t)c * HW + p) : 0.0f;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const int c = c0 + cl + 4 * i;
        uint32_t b = 128u >> sh;
        if (p < HW && c < C) b = quant_code_bits(v[i], qc) & 0xffu;
        tile[px][cl + 4 * i] = (uint8_t)(((b << sh) ^ 0x80u) & 0xffu);
    }
    __syncthreads();
    {
        const int pp = tid >> 2, w = tid & 3;
        const int pg = p0 + pp, cb = c0 + 16 * w;
        if (pg < HW && cb < Cp)
            *reinterpret_cast<uint4 *>(q + ((int64_t)n * HW + pg) * Cp + cb) =
                *reinterpret_cast<const uint4 *>(&tile[pp][16 * w]);
    }
}

// sum of the per-tap constants of output channel `ol` over the out-of-bounds taps in `oob` (fixed tap order)
__device__ __forceinline__ float tap_corr(const float *tap_c, uint32_t oob, int taps, int npass, int ol) {
    float corr = 0.0f;
    for (int t = 0; t < taps; ++t)
        if (oob >> t & 1u) corr = __fadd_rn(corr, tap_c[t * npass + ol]);
    return corr;
}

// kPre: the activation loads of ring stage s + 1 are issued right after stage s has been quantised, so they overlap
// the barrier, the MMA issue and the wait for the ring slot.  It costs 11 more live registers (3 resident CTAs
// per SM instead of 4): a gain for deep-K layers on small feature maps whose grids do not fill the SMs anyway
// (14 x 14 / 7 x 7), a 4% loss for the large shallow ones -- dfq_pw_i8 picks per layer.
//
// kConv: dense k x k convolution (groups = 1, any stride / zero padding / dilation) as an implicit GEMM on the same
// pipeline.  The reduction index runs tap-major, channel-minor, channels padded to a multiple of 16, so every
// 16-byte K chunk belongs to ONE tap: a thread reads it at the tap's input pixel (coalesced along the row) or, when
// that pixel lies in the padding, stores s8 zeros.  Zero padding is an exact 0.0 in the reference (it quantises
// before F.conv2d pads) while an s8 zero stands for c_x, so the epilogue takes c_x * (s_w * sum_c b[o,c,t] + C c_w)
// back out for every out-of-bounds tap t of the pixel (a [taps][channels] table in shared memory; interior pixels
// skip it).
//
// kQ (with kConv): the input was quantised ONCE by quant_nhwc_s8_kernel into centred s8 codes laid out
// [N][H][W][C padded to 16], so a K chunk is one aligned 16-byte word of the tap's pixel: the producer is a 128-bit
// load, four dp4a for the code sum and the shared-memory store -- instead of 16 fp32 loads and 16 quantiser
// evaluations per tap (9x the same activation for a 3 x 3 kernel).
// kLin: the activation is row-major [G][K] with K contiguous (a linear layer): a thread's 16-float chunk is four
// 128-bit loads.  As 16 scalar loads the 32 lanes of every load instruction hit 32 different lines -- the LSU, not
// the latency, set the classifier's 7.5 us per slab (256 x 1280 -> 1000: 0.153 ms for 20 slabs).
template <bool kPre, bool kConv, bool kQ, bool kLin = false, bool kOutQ = false>
__global__ void __launch_bounds__(kThreads, kOutQ ? 2 : (kPre ? 3 : 4))
pw_i8_kernel(PwI8Args a) {
    extern __shared__ __align__(128) uint8_t smem[];
    // Weight ring: kBS stages, the bulk copy of slab s + kAhead is issued at step s.  With the two-stage ring the copy
    // of a slab was only issued in the step that consumes it, so its L2 latency (~1.5 k cycles for 16 KB) sat on every
    // step's critical path -- 3.5 k cycles per 64-byte slab on the decoder's 3 x 3 layers against 256 cycles of MMA.
    constexpr int kBS = kQ ? 4 : 2, kAhead = kBS - 2;
    __shared__ __align__(8) uint64_t bars[6];           // full[kBS] (weights landed), done[2] at [4] (MMAs of the stage retired)
    __shared__ uint32_t tmem_base_s;
    __shared__ int colsum_s[2][kI8M];
    __shared__ __align__(16) float chan_c[256];          // per output channel of this pass: c_x*wr[o] + K c_x c_w + bias[o]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int pass = blockIdx.x % a.n_pass;
    const int64_t tile = blockIdx.x / a.n_pass;
    uint8_t *sA = smem;                                  // [2][4 K-chunks][128 rows][16 B]
    uint8_t *sB = smem + 2 * kI8ASlab;                   // [2][4 K-chunks][npass rows][16 B]
    const uint32_t bslab = (uint32_t)a.npass * kI8Slab;
    float *tap_c = reinterpret_cast<float *>(sB + kBS * bslab);   // kConv: [taps][npass] c_x * (wrt[t][o] + C c_w s_x..)
    const uint32_t sA_u = smem_u32(sA), sB_u = smem_u32(sB);
    const uint32_t full0 = smem_u32(&bars[0]), done0 = smem_u32(&bars[4]);

    if (tid == 0) {
        for (int i = 0; i < kBS; ++i) mbar_init(full0 + 8 * i, 1);
        mbar_init(done0, 1); mbar_init(done0 + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32(&tmem_base_s)), "r"((uint32_t)a.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // ---- this thread's pixel (GEMM row) and K half -------------------------------------------------------
    const int r = tid & (kI8M - 1), h = tid >> 7;
    const int64_t g = tile * kI8M + r;
    const bool valid = g < a.G;
    // image and pixel of this row; 32-bit division (dfq_pw_i8 refuses G >= 2^31), shared with the epilogue: the
    // TMEM lane a warp may read, 32 * (warp % 4) + lane, is this thread's own row
    const uint32_t n32 = valid ? (uint32_t)g / (uint32_t)a.P : 0u;
    const int64_t n = n32;
    const int p = valid ? (int)((uint32_t)g - n32 * (uint32_t)a.P) : 0;
    const float *xb = a.x + n * a.xs_n + (kConv ? 0 : (int64_t)p * a.xs_p);
    const int xs_k = (int)a.xs_k;                        // K * P < 2^31 elements per image (kConv: H * W)
    int ih0 = 0, iw0 = 0;
    if (kConv) {
        const int oh = p / a.OW, ow = p - oh * a.OW;
        ih0 = oh * a.stride - a.pad; iw0 = ow * a.stride - a.pad;
    }
    // chunk q of the reduction (16 K bytes): is it live, may this thread read it, where, how many real channels
    auto chunk_src = [&](int q, bool &live, bool &inb, const float *&px, int &nch) {
        if (kConv) {
            const int t = q / a.cchunks, cc = q - t * a.cchunks;          // warp-uniform
            live = t < a.taps;
            const int i = t / a.kw, j = t - i * a.kw;
            const int ih = ih0 + i * a.dil, iw = iw0 + j * a.dil;
            inb = valid && live && (unsigned)ih < (unsigned)a.H && (unsigned)iw < (unsigned)a.W;
            nch = a.C - cc * 16;
            px = xb + ((int64_t)(cc * 16) * a.H + ih) * a.W + iw;         // dereferenced only when inb
        } else {
            live = q * 16 < a.K;
            inb = valid;
            nch = a.K - q * 16;
            px = xb + (int64_t)(q * 16) * xs_k;
        }
    };
    // kQ: the 16-byte word of chunk q at this thread's tap pixel (zeros in the padding / beyond the last tap)
    // The chunks a thread reads are q = h, h + 2, h + 4, ... in this order (load_stage is called for s = 0, 1, 2, ...),
    // so the tap and the channel chunk are STEPPED, not divided out: ncu charged 35 % of the kernel's instructions to
    // the two integer divisions per 16-byte load (profiles/r1zz_conv_i8_ncu.md).  q_tap is the word of channel chunk 0
    // at the current tap's pixel, or NULL in the padding / past the last tap.
    const int8_t *xqb = kQ ? a.xq + n * ((int64_t)a.H * a.W * a.cchunks * 16) : nullptr;
    int q_t = 0, q_cc = h, q_i = 0, q_j = 0;
    const int8_t *q_tap = nullptr;
    auto tap_ptr = [&]() {
        const int ih = ih0 + q_i * a.dil, iw = iw0 + q_j * a.dil;
        const bool inb = valid && q_t < a.taps && (unsigned)ih < (unsigned)a.H && (unsigned)iw < (unsigned)a.W;
        q_tap = inb ? xqb + ((int64_t)ih * a.W + iw) * (a.cchunks * 16) : nullptr;
    };
    if (kQ) {
        while (q_cc >= a.cchunks) { q_cc -= a.cchunks; ++q_t; if (++q_j == a.kw) { q_j = 0; ++q_i; } }
        tap_ptr();
    }
    auto chunk_q = [&](int) {
        const uint4 w = q_tap ? __ldg(reinterpret_cast<const uint4 *>(q_tap + q_cc * 16)) : make_uint4(0u, 0u, 0u, 0u);
        q_cc += 2;
        if (q_cc >= a.cchunks) {                                          // warp-uniform
            do { q_cc -= a.cchunks; ++q_t; if (++q_j == a.kw) { q_j = 0; ++q_i; } } while (q_cc >= a.cchunks);
            tap_ptr();
        }
        return w;
    };
    // activations of one ring stage: two 16-wide K chunks per thread, all loads in flight before any use
    float v[kQ ? 1 : 2][kQ ? 1 : 16];
    // kQ + kPre: the code words of kD stages ahead are in flight.  With one stage ahead the slab loop was ONE L2 round
    // trip per iteration (the loads of stage s + 1 leave after the stores of stage s and are needed by the stores of
    // stage s + 1): ~950 cycles per 256 cycles of MMA, whatever was done about the barrier or the weight stream
    // (profiles/r2_conv_q8_ncu.md).  The slab loop is unrolled by kD so that the ring index is a compile-time constant.
    constexpr int kD = (kQ && kPre) ? 4 : 1;
    uint4 vq[kD][2];
    auto load_stage = [&](int s, int u) {
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            if (kQ) {
                vq[kQ ? u : 0][c] = chunk_q(s * 4 + h + 2 * c);
            } else {
                bool live, inb; const float *px; int nch;
                chunk_src(s * 4 + h + 2 * c, live, inb, px, nch);
                if (kLin && live && inb && nch >= 16) {               // (rows past the end / a ragged last chunk: below)
                    const float4 *p4 = reinterpret_cast<const float4 *>(px);
#pragma unroll
                    for (int j4 = 0; j4 < 4; ++j4) {
                        const float4 t = __ldg(p4 + j4);
                        v[c][kQ ? 0 : 4 * j4] = t.x; v[c][kQ ? 0 : 4 * j4 + 1] = t.y;
                        v[c][kQ ? 0 : 4 * j4 + 2] = t.z; v[c][kQ ? 0 : 4 * j4 + 3] = t.w;
                    }
                } else if (live) {                                    // warp-uniform: padded chunks cost nothing
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[c][kQ ? 0 : j] = (inb && j < nch) ? __ldg(px + j * xs_k) : 0.0f;
                }
            }
        }
    };
    if (kPre) {                                         // in flight during the TMEM allocation and the barrier
#pragma unroll
        for (int u = 0; u < kD; ++u)
            if (u == 0 || u < a.slabs) load_stage(u, u);
    }
    const QuantConsts qc = consts_from_slot(*a.slot, a.bits, 0, 0.f, 0.f);
    const int sh = 8 - a.bits;                          // < 8-bit grids: a = (code - 2^(bits-1)) << sh, s_x >> sh
    const float sx = __fmul_rn(qc.scale, 1.0f / (float)(1 << sh));
    const float cx = __fmaf_rn(128.0f, sx, qc.add_back);
    const int o0 = pass * a.o_per_pass, o_end = min(a.O, o0 + a.o_per_pass);
    if (tid < a.npass) {
        const int o = o0 + tid;
        float c = 0.0f;
        if (o < o_end) {
            c = __fmaf_rn(cx, __ldg(a.wr + o), __fmul_rn(__fmul_rn((float)(kConv ? a.Kreal : a.K), a.c_w), cx));
            if (a.bias) c = __fadd_rn(c, __ldg(a.bias + o));
        }
        chan_c[tid] = c;
    }
    if (kConv) {                                          // what an out-of-bounds tap contributed through c_x
        const float ccw = __fmul_rn(__fmul_rn((float)a.C, a.c_w), cx);
        for (int idx = tid; idx < a.taps * a.npass; idx += kThreads) {
            const int t = idx / a.npass, ol = idx - t * a.npass, o = o0 + ol;
            tap_c[idx] = o < o_end ? __fmaf_rn(cx, __ldg(a.wrt + (int64_t)t * a.O + o), ccw) : 0.0f;
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    const int8_t *wsrc = a.wpk + (int64_t)pass * a.slabs * bslab;
    if (tid == 0) {
        for (int s0 = 0; s0 < kAhead && s0 < a.slabs; ++s0) {
            mbar_expect_tx(full0 + 8 * s0, bslab);
            bulk_g2s(sB_u + s0 * bslab, wsrc + (int64_t)s0 * bslab, bslab, full0 + 8 * s0);
        }
    }
    // instruction descriptor: D = s32, A = B = s8, both K-major, N = npass, M = 128
    const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(a.npass >> 3) << 17) |
                           ((uint32_t)(kI8M >> 4) << 24);
    int csum = 0;

    for (int sb = 0; sb < a.slabs; sb += kD) {
#pragma unroll
    for (int u = 0; u < kD; ++u) {
        const int s = sb + u;
        if (s >= a.slabs) break;                         // (uniform)
        const int buf = s & 1;
        if (!kPre) load_stage(s, 0);                     // 1. this stage's activations
        // 2. the ring stage is free once the MMAs issued two slabs ago have retired
        if (s >= 2) mbar_wait(done0 + 8 * buf, ((s >> 1) - 1) & 1);
        const int bbuf = s & (kBS - 1);
        if (tid == 0 && s + kAhead < a.slabs) {          // its stage was last read by slab s - 2, retired above
            const int s2 = s + kAhead, b2 = s2 & (kBS - 1);
            mbar_expect_tx(full0 + 8 * b2, bslab);
            bulk_g2s(sB_u + b2 * bslab, wsrc + (int64_t)s2 * bslab, bslab, full0 + 8 * b2);
        }
        // 3. quantise on load (utils/quantize.py:72-74), centre to s8, pack, canonical-order store
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            uint32_t w4[4] = {0u, 0u, 0u, 0u};
            if (kQ) {
                w4[0] = vq[kQ ? u : 0][c].x; w4[1] = vq[kQ ? u : 0][c].y; w4[2] = vq[kQ ? u : 0][c].z; w4[3] = vq[kQ ? u : 0][c].w;
#pragma unroll
                for (int q4 = 0; q4 < 4; ++q4) csum = __dp4a((int)w4[q4], 0x01010101, csum);
            } else {
                bool live, inb; const float *px; int nch;
                chunk_src(s * 4 + h + 2 * c, live, inb, px, nch);
                if (live && (!kConv || __any_sync(0xffffffffu, inb))) {
#pragma unroll
                    for (int q4 = 0; q4 < 4; ++q4) {
                        uint32_t b[4];
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            b[j] = (inb && 4 * q4 + j < nch) ? quant_code_bits(v[c][kQ ? 0 : 4 * q4 + j], qc) : (128u >> sh);
                        const uint32_t lo = __byte_perm(b[0], b[1], 0x0040), hi = __byte_perm(b[2], b[3], 0x0040);
                        w4[q4] = (__byte_perm(lo, hi, 0x5410) << sh) ^ 0x80808080u;
                        csum = __dp4a((int)w4[q4], 0x01010101, csum);
                    }
                }
            }
            *reinterpret_cast<uint4 *>(sA + buf * kI8ASlab + (h + 2 * c) * (kI8M * 16) + r * 16) =
                make_uint4(w4[0], w4[1], w4[2], w4[3]);
        }
        if (kPre && s + kD < a.slabs) load_stage(s + kD, u);
        // 4. generic-proxy stores -> visible to the tensor core's async proxy, then one thread issues
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            mbar_wait(full0 + 8 * bbuf, (s / kBS) & 1);
            tc_fence_after();
#pragma unroll
            for (int kk = 0; kk < 2; ++kk) {
                const uint64_t ad = umma_desc(sA_u + buf * kI8ASlab + kk * 2 * (kI8M * 16), kI8M * 16, 128);
                const uint64_t bd = umma_desc(sB_u + bbuf * bslab + kk * 2 * (a.npass * 16), a.npass * 16, 128);
                umma_i8(tmem, ad, bd, idesc, (s | kk) != 0);
            }
            umma_commit(done0 + 8 * buf);
        }
    }
    }
    colsum_s[h][r] = csum;
    {   // MMAs retire in order: the last stage's barrier covers the whole accumulation
        const int last = a.slabs - 1;
        mbar_wait(done0 + 8 * (last & 1), (last >> 1) & 1);
    }
    tc_fence_after();
    __syncthreads();

    // ---- epilogue: warp w reads TMEM lanes 32*(w%4).., 16-column chunks w/4, w/4+2, ... ---------------------
    const int er = r;
    const bool ev = valid;
    const int64_t en = n;
    const int ep = p;
    float *yb = a.y + en * a.ys_n + (int64_t)ep * a.ys_p;
    const int ys_o = (int)a.ys_o;
    const float fa = __fmul_rn(sx, a.s_w);
    const float row_c = __fmul_rn(__fmul_rn(sx, a.c_w), (float)(colsum_s[0][er] + colsum_s[1][er]));
    const float floor_v = a.relu ? 0.0f : -INFINITY;
    float lo = INFINITY, hi = -INFINITY;
    uint32_t oob = 0;                                             // kConv: taps of this pixel that fell in the padding
    if (kConv && ev) {
        for (int t = 0; t < a.taps; ++t) {
            const int i = t / a.kw, j = t - i * a.kw;
            const int ih = ih0 + i * a.dil, iw = iw0 + j * a.dil;
            if (!((unsigned)ih < (unsigned)a.H && (unsigned)iw < (unsigned)a.W)) oob |= 1u << t;
        }
    }
    QuantConsts oq0 = qc, oq1 = qc;
    if (kOutQ) {
        oq0 = consts_from_slot(*a.oslot[0], 8, 0, 0.f, 0.f);
        if (a.yq[1]) oq1 = consts_from_slot(*a.oslot[1], 8, 0, 0.f, 0.f);
    }
    for (int chunk = warp >> 2; chunk * 16 < a.npass; chunk += 2) {
        const int oc = o0 + chunk * 16;
        if (oc >= o_end) break;                                   // warp-uniform
        uint32_t d[16];
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, "
                     "%13, %14, %15}, [%16];"
                     : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]),
                       "=r"(d[8]), "=r"(d[9]), "=r"(d[10]), "=r"(d[11]), "=r"(d[12]), "=r"(d[13]), "=r"(d[14]),
                       "=r"(d[15])
                     : "r"(tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(chunk * 16)));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        const float4 *cc = reinterpret_cast<const float4 *>(chan_c + chunk * 16);
        const int nvalid = min(16, o_end - oc);                   // warp-uniform
        float *py = yb + (int64_t)oc * ys_o;
        if (!ev) {                                                // rows past the end (last tile only)
        } else if (nvalid == 16) {                                       // the common case: no per-column predicate
            // padding correction of this pixel for the chunk's 16 channels: tap-outer, channel-inner (the same
            // ascending tap order per channel as tap_corr), 128-bit shared loads -- the per-output form walked all
            // taps for every output of every warp that touches an image border (23 % of the instructions)
            float corr[16];
            if (kConv && oob) {
#pragma unroll
                for (int j = 0; j < 16; ++j) corr[j] = 0.0f;
                for (uint32_t mm = oob; mm; mm &= mm - 1) {
                    const float4 *tc = reinterpret_cast<const float4 *>(tap_c + (__ffs(mm) - 1) * a.npass + chunk * 16);
#pragma unroll
                    for (int j4 = 0; j4 < 4; ++j4) {
                        const float4 v4 = tc[j4];
                        corr[4 * j4] = __fadd_rn(corr[4 * j4], v4.x); corr[4 * j4 + 1] = __fadd_rn(corr[4 * j4 + 1], v4.y);
                        corr[4 * j4 + 2] = __fadd_rn(corr[4 * j4 + 2], v4.z); corr[4 * j4 + 3] = __fadd_rn(corr[4 * j4 + 3], v4.w);
                    }
                }
            }
            float y16[16];
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4) {
                const float4 c4 = cc[j4];
                const float cj[4] = {c4.x, c4.y, c4.z, c4.w};
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    float cst = __fadd_rn(row_c, cj[t]);
                    if (kConv && oob) cst = __fsub_rn(cst, corr[4 * j4 + t]);
                    const float yv = fmaxf(__fmaf_rn(fa, (float)(int)d[4 * j4 + t], cst), floor_v);
                    if constexpr (kOutQ) {
                        y16[4 * j4 + t] = yv;
                    } else {
                        *py = yv;
                        py += ys_o;
                        lo = fminf(lo, yv); hi = fmaxf(hi, yv);
                    }
                }
            }
            if constexpr (kOutQ) {                                 // 16 channels of one pixel = one 128-bit store per grid
                const int64_t off = (en * a.P + ep) * a.cp_out + oc;
                uint32_t w[4];
                pack_codes<16>(y16, oq0, w);
                *reinterpret_cast<uint4 *>(a.yq[0] + off) = make_uint4(w[0], w[1], w[2], w[3]);
                if (a.yq[1]) {
                    pack_codes<16>(y16, oq1, w);
                    *reinterpret_cast<uint4 *>(a.yq[1] + off) = make_uint4(w[0], w[1], w[2], w[3]);
                }
            }
        } else if (kOutQ) {                                        // ragged last chunk: channels >= nvalid are s8 zeros
            float y16[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                float cst = __fadd_rn(row_c, chan_c[chunk * 16 + j]);
                if (kConv && oob) cst = __fsub_rn(cst, tap_corr(tap_c, oob, a.taps, a.npass, chunk * 16 + j));
                y16[j] = j < nvalid ? fmaxf(__fmaf_rn(fa, (float)(int)d[j], cst), floor_v) : 0.0f;
            }
            const int64_t off = (en * a.P + ep) * a.cp_out + oc;
            for (int gi = 0; gi < 2; ++gi) {
                if (!a.yq[gi]) break;
                uint32_t w[4];
                pack_codes<16>(y16, gi ? oq1 : oq0, w);
#pragma unroll
                for (int q4 = 0; q4 < 4; ++q4) w[q4] = keep_left(w[q4], nvalid - 4 * q4);
                if (oc + 16 <= a.cp_out) *reinterpret_cast<uint4 *>(a.yq[gi] + off) = make_uint4(w[0], w[1], w[2], w[3]);
                else
                    for (int q4 = 0; q4 < 4 && oc + 4 * q4 < a.cp_out; ++q4)
                        *reinterpret_cast<uint32_t *>(a.yq[gi] + off + 4 * q4) = w[q4];
            }
        } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) {                        // compile-time indices keep d[] in registers
                if (j < nvalid) {
                    float cst = __fadd_rn(row_c, chan_c[chunk * 16 + j]);
                    if (kConv && oob) cst = __fsub_rn(cst, tap_corr(tap_c, oob, a.taps, a.npass, chunk * 16 + j));
                    const float yv = fmaxf(__fmaf_rn(fa, (float)(int)d[j], cst), floor_v);
                    *py = yv;
                    py += ys_o;
                    lo = fminf(lo, yv); hi = fmaxf(hi, yv);
                }
            }
        }
        __syncwarp();                                             // tcgen05.ld is warp-collective
    }
    if (!ev) { lo = INFINITY; hi = -INFINITY; }
    if (a.stats) {                                                // per-sample extrema, encoded (common.cuh)
        const int64_t n0 = __shfl_sync(0xffffffffu, en, 0);
        const bool uni = __all_sync(0xffffffffu, ev && en == n0);
        if (uni) {
            lo = warp_min(lo); hi = warp_max(hi);
            if (lane == 0 && hi >= lo) {
                atomicMax(a.stats + 2 * en, enc_ordered(-lo));
                atomicMax(a.stats + 2 * en + 1, enc_ordered(hi));
            }
        } else if (ev && hi >= lo) {
            atomicMax(a.stats + 2 * en, enc_ordered(-lo));
            atomicMax(a.stats + 2 * en + 1, enc_ordered(hi));
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)a.tmem_cols)
                     : "memory");
}


// ---------------------------------------------------------------------------------------------------------
// Persistent, role-split form of the pointwise / linear kernel (same arithmetic, same packed weights).
// A CTA walks the work items blockIdx.x, blockIdx.x + gridDim.x, ... (item = tile * n_pass + pass, gridDim.x a
// multiple of n_pass so that the pass is fixed per CTA).  Roles meet only at mbarriers:
//   warps 0-3 / 4-7  two worker groups taking alternate items.  Thread r of a group owns GEMM row r of its tile from
//              start to end: per 64-byte K slab it loads its 64 activations (two rounds of 32 loads in flight),
//              quantises, packs, stores the canonical-order rows of the group's ring stage and arrives on
//              full_a[group][stage] (count 128); then it waits for acc_full[group] and runs the epilogue of the
//              SAME row (its code sum never leaves its registers): tcgen05.ld, FMA chain, stores, per-sample
//              extrema.  While one group is in its epilogue the other one is producing: loads, tensor pipe and
//              stores of neighbouring tiles overlap whatever the layer's K : N ratio is.
//   warp 8     lane 0 issues the MMAs in item order: waits full_a / full_b, 2 x tcgen05.mma into the group's TMEM
//              accumulator, commits to empty_a[group][stage] and empty_b[stage]; after the last slab to acc_full.
//   warp 9     lane 0 streams the weight slabs (3-deep ring): waits empty_b, one bulk copy per slab onto full_b.
// The TMEM allocation (2 x npass columns), the barriers, the quantiser constants and the per-channel constants are
// set up once per CTA instead of once per tile.
constexpr int kWsBStages = 3;
constexpr int kWsThreads = 320;

__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

__global__ void __launch_bounds__(kWsThreads, 2)
pw_i8_ws_kernel(PwI8Args a) {
    extern __shared__ __align__(128) uint8_t smem[];
    // full_a[2][2], empty_a[2][2], acc_full[2], full_b[3], empty_b[3]
    __shared__ __align__(8) uint64_t bars[10 + 2 * kWsBStages];
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(16) float chan_c[256];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int pass = blockIdx.x % a.n_pass;
    const int64_t items = a.tiles * a.n_pass;
    const uint32_t bslab = (uint32_t)a.npass * kI8Slab;
    uint8_t *sA = smem;                                   // [group][2 stages][4 K-chunks][128 rows][16 B]
    uint8_t *sB = smem + 4 * kI8ASlab;                    // [3 stages][4 K-chunks][npass rows][16 B]
    const uint32_t sA_u = smem_u32(sA), sB_u = smem_u32(sB);
    const uint32_t full_a = smem_u32(&bars[0]), empty_a = smem_u32(&bars[4]), acc_full = smem_u32(&bars[8]),
                   full_b = smem_u32(&bars[10]), empty_b = smem_u32(&bars[10 + kWsBStages]);
    if (tid == 0) {
        for (int i = 0; i < 4; ++i) { mbar_init(full_a + 8 * i, kI8M); mbar_init(empty_a + 8 * i, 1); }
        for (int i = 0; i < 2; ++i) mbar_init(acc_full + 8 * i, 1);
        for (int i = 0; i < kWsBStages; ++i) { mbar_init(full_b + 8 * i, 1); mbar_init(empty_b + 8 * i, 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 8) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32(&tmem_base_s)), "r"((uint32_t)a.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    const QuantConsts qc = consts_from_slot(*a.slot, a.bits, 0, 0.f, 0.f);
    const int sh = 8 - a.bits;
    const float sx = __fmul_rn(qc.scale, 1.0f / (float)(1 << sh));
    const float cx = __fmaf_rn(128.0f, sx, qc.add_back);
    const int o0 = pass * a.o_per_pass, o_end = min(a.O, o0 + a.o_per_pass);
    if (tid < a.npass) {
        const int o = o0 + tid;
        float c = 0.0f;
        if (o < o_end) {
            c = __fmaf_rn(cx, __ldg(a.wr + o), __fmul_rn(__fmul_rn((float)a.K, a.c_w), cx));
            if (a.bias) c = __fadd_rn(c, __ldg(a.bias + o));
        }
        chan_c[tid] = c;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    const uint32_t acc_cols = (uint32_t)a.tmem_cols >> 1;        // columns per accumulator buffer (>= npass)

    if (warp < 8) {
        // ======================================================================== workers (two groups)
        const uint32_t grp = (uint32_t)warp >> 2;
        const int r = tid & (kI8M - 1);                          // GEMM row = TMEM lane 32 * (warp % 4) + lane
        const int xs_k = (int)a.xs_k, ys_o = (int)a.ys_o;
        const float fa = __fmul_rn(sx, a.s_w), sxcw = __fmul_rn(sx, a.c_w);
        const float floor_v = a.relu ? 0.0f : -INFINITY;
        const uint32_t fa_bar = full_a + 16 * grp, ea_bar = empty_a + 16 * grp, af_bar = acc_full + 8 * grp;
        uint8_t *sAg = sA + grp * 2 * kI8ASlab;
        uint32_t gs = 0, jt = 0;
        for (int64_t item = blockIdx.x + (int64_t)grp * gridDim.x; item < items; item += 2 * (int64_t)gridDim.x, ++jt) {
            const int64_t g = (item / a.n_pass) * kI8M + r;
            const bool valid = g < a.G;
            const uint32_t n32 = valid ? (uint32_t)g / (uint32_t)a.P : 0u;
            const int p = valid ? (int)((uint32_t)g - n32 * (uint32_t)a.P) : 0;
            const float *xb = a.x + (int64_t)n32 * a.xs_n + (int64_t)p * a.xs_p;
            int csum = 0;
            for (int s = 0; s < a.slabs; ++s, ++gs) {
                const uint32_t stage = gs & 1u, use = gs >> 1;
                if (use >= 1) mbar_wait(ea_bar + 8 * stage, (use - 1) & 1);
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    const int k0 = s * kI8Slab + half * 32;
                    float v[32];
                    if (k0 < a.K) {                                  // uniform
                        const float *px = xb + (int64_t)k0 * xs_k;
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = (valid && k0 + j < a.K) ? __ldg(px + j * xs_k) : 0.0f;
                    }
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        uint32_t w4[4] = {0u, 0u, 0u, 0u};
                        if (k0 + 16 * c < a.K) {
#pragma unroll
                            for (int q4 = 0; q4 < 4; ++q4) {
                                uint32_t b[4];
#pragma unroll
                                for (int j = 0; j < 4; ++j) {
                                    const int kk = 16 * c + 4 * q4 + j;
                                    b[j] = (valid && k0 + kk < a.K) ? quant_code_bits(v[kk], qc) : (128u >> sh);
                                }
                                const uint32_t lo = __byte_perm(b[0], b[1], 0x0040), hi = __byte_perm(b[2], b[3], 0x0040);
                                w4[q4] = (__byte_perm(lo, hi, 0x5410) << sh) ^ 0x80808080u;
                                csum = __dp4a((int)w4[q4], 0x01010101, csum);
                            }
                        }
                        *reinterpret_cast<uint4 *>(sAg + stage * kI8ASlab + (2 * half + c) * (kI8M * 16) + r * 16) =
                            make_uint4(w4[0], w4[1], w4[2], w4[3]);
                    }
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_arrive(fa_bar + 8 * stage);
            }
            // ---- epilogue of the same row
            float *yb = a.y + (int64_t)n32 * a.ys_n + (int64_t)p * a.ys_p;
            const float row_c = __fmul_rn(sxcw, (float)csum);
            mbar_wait(af_bar, jt & 1);
            tc_fence_after();
            float lo = INFINITY, hi = -INFINITY;
            for (int chunk = 0; chunk * 16 < a.npass; ++chunk) {
                const int oc = o0 + chunk * 16;
                if (oc >= o_end) break;                               // uniform
                uint32_t d[16];
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, "
                             "%12, %13, %14, %15}, [%16];"
                             : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]),
                               "=r"(d[7]), "=r"(d[8]), "=r"(d[9]), "=r"(d[10]), "=r"(d[11]), "=r"(d[12]), "=r"(d[13]),
                               "=r"(d[14]), "=r"(d[15])
                             : "r"(tmem + ((uint32_t)((warp & 3) * 32) << 16) + grp * acc_cols + (uint32_t)(chunk * 16)));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                const float4 *cc = reinterpret_cast<const float4 *>(chan_c + chunk * 16);
                const int nvalid = min(16, o_end - oc);               // uniform
                float *py = yb + (int64_t)oc * ys_o;
                if (!valid) {
                } else if (nvalid == 16) {
#pragma unroll
                    for (int j4 = 0; j4 < 4; ++j4) {
                        const float4 c4 = cc[j4];
                        const float cj[4] = {c4.x, c4.y, c4.z, c4.w};
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            const float yv = fmaxf(__fmaf_rn(fa, (float)(int)d[4 * j4 + t], __fadd_rn(row_c, cj[t])), floor_v);
                            *py = yv;
                            py += ys_o;
                            lo = fminf(lo, yv); hi = fmaxf(hi, yv);
                        }
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        if (j < nvalid) {
                            const float yv = fmaxf(__fmaf_rn(fa, (float)(int)d[j], __fadd_rn(row_c, chan_c[chunk * 16 + j])), floor_v);
                            *py = yv;
                            py += ys_o;
                            lo = fminf(lo, yv); hi = fmaxf(hi, yv);
                        }
                    }
                }
                __syncwarp();
            }
            tc_fence_before();      // our TMEM reads are done before the full_a arrival that lets the next MMA start
            if (a.stats) {
                const int64_t n = n32;
                const int64_t n0 = __shfl_sync(0xffffffffu, n, 0);
                const bool uni = __all_sync(0xffffffffu, valid && n == n0);
                if (uni) {
                    lo = warp_min(lo); hi = warp_max(hi);
                    if (lane == 0 && hi >= lo) {
                        atomicMax(a.stats + 2 * n, enc_ordered(-lo));
                        atomicMax(a.stats + 2 * n + 1, enc_ordered(hi));
                    }
                } else if (valid && hi >= lo) {
                    atomicMax(a.stats + 2 * n, enc_ordered(-lo));
                    atomicMax(a.stats + 2 * n + 1, enc_ordered(hi));
                }
            }
        }
    } else if (warp == 8) {
        // ==================================================================================== MMA issuer
        if (lane == 0) {
            const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(a.npass >> 3) << 17) |
                                   ((uint32_t)(kI8M >> 4) << 24);
            uint32_t gsa[2] = {0u, 0u}, gsb = 0, it = 0;
            for (int64_t item = blockIdx.x; item < items; item += gridDim.x, ++it) {
                const uint32_t grp = it & 1u;
                for (int s = 0; s < a.slabs; ++s, ++gsb) {
                    const uint32_t sta = gsa[grp] & 1u, usea = gsa[grp] >> 1;
                    const uint32_t stb = gsb % kWsBStages, useb = gsb / kWsBStages;
                    gsa[grp] += 1;
                    mbar_wait(full_b + 8 * stb, useb & 1);
                    mbar_wait(full_a + 16 * grp + 8 * sta, usea & 1);
                    tc_fence_after();
#pragma unroll
                    for (int kk = 0; kk < 2; ++kk) {
                        const uint64_t ad = umma_desc(sA_u + (grp * 2 + sta) * kI8ASlab + kk * 2 * (kI8M * 16), kI8M * 16, 128);
                        const uint64_t bd = umma_desc(sB_u + stb * bslab + kk * 2 * (a.npass * 16), a.npass * 16, 128);
                        umma_i8(tmem + grp * acc_cols, ad, bd, idesc, (s | kk) != 0);
                    }
                    umma_commit(empty_a + 16 * grp + 8 * sta);
                    umma_commit(empty_b + 8 * stb);
                }
                umma_commit(acc_full + 8 * grp);
            }
        }
    } else {
        // ================================================================================= weight streamer
        if (lane == 0) {
            const int8_t *wsrc = a.wpk + (int64_t)pass * a.slabs * bslab;
            uint32_t gsb = 0;
            for (int64_t item = blockIdx.x; item < items; item += gridDim.x) {
                for (int s = 0; s < a.slabs; ++s, ++gsb) {
                    const uint32_t stb = gsb % kWsBStages, useb = gsb / kWsBStages;
                    if (useb >= 1) mbar_wait(empty_b + 8 * stb, (useb - 1) & 1);
                    mbar_expect_tx(full_b + 8 * stb, bslab);
                    bulk_g2s(sB_u + stb * bslab, wsrc + (int64_t)s * bslab, bslab, full_b + 8 * stb);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)a.tmem_cols)
                     : "memory");
}

}  // namespace dfq

using namespace dfq;

// Split of the output channels into passes of <= 256 (one UMMA N each, padded to a multiple of 16) and the K
// padding to whole 64-byte slabs.  packed_bytes = n_pass * slabs * npass * 64.
// o_cap: most output channels one pass (= one CTA's accumulator) may hold, 16..256 (0 = 256).  A narrower pass means
// more CTAs for the same layer -- the A tile is quantised once per pass, so it only pays when the default grid
// leaves SMs idle (the 256 x 1280 -> 1000 classifier is 2 tiles x 4 passes = 8 CTAs at 256 channels per pass).
static int plan_cap(int O, int K, int o_cap, int *n_pass, int *o_per_pass, int *npass, int *slabs,
                    int64_t *packed_bytes) {
    if (O <= 0 || K <= 0 || o_cap < 0) return (int)cudaErrorInvalidValue;
    int cap = (o_cap == 0 || o_cap > 256) ? 256 : (o_cap < 16 ? 16 : o_cap / 16 * 16);
    int np = (O + cap - 1) / cap;
    int opp = ((O + np - 1) / np + 15) / 16 * 16;
    np = (O + opp - 1) / opp;
    *n_pass = np; *o_per_pass = opp; *npass = opp;
    *slabs = (K + kI8Slab - 1) / kI8Slab;
    *packed_bytes = (int64_t)np * *slabs * opp * kI8Slab;
    return 0;
}
extern "C" int dfq_pw_i8_plan_host(int O, int K, int *n_pass, int *o_per_pass, int *npass, int *slabs,
                                   int64_t *packed_bytes) {
    return plan_cap(O, K, 0, n_pass, o_per_pass, npass, slabs, packed_bytes);
}
extern "C" int dfq_pw_i8_plan_cap_host(int O, int K, int o_cap, int *n_pass, int *o_per_pass, int *npass, int *slabs,
                                       int64_t *packed_bytes) {
    return plan_cap(O, K, o_cap, n_pass, o_per_pass, npass, slabs, packed_bytes);
}

// codes: [O, K] centred weight codes (code - 128) on the host.  packed: the layout the kernel's bulk copies read,
// [pass][slab][K chunk of 16][row within pass][16 B], zero padded (a zero code adds nothing to the dot product).
extern "C" int dfq_pw_i8_pack_cap_host(const int8_t *codes, int O, int K, int o_cap, int8_t *packed);
extern "C" int dfq_pw_i8_pack_host(const int8_t *codes, int O, int K, int8_t *packed) {
    return dfq_pw_i8_pack_cap_host(codes, O, K, 0, packed);
}
extern "C" int dfq_pw_i8_pack_cap_host(const int8_t *codes, int O, int K, int o_cap, int8_t *packed) {
    int np, opp, npass, slabs; int64_t bytes;
    int e = plan_cap(O, K, o_cap, &np, &opp, &npass, &slabs, &bytes);
    if (e) return e;
    for (int64_t i = 0; i < bytes; ++i) packed[i] = 0;
    for (int o = 0; o < O; ++o) {
        const int pass = o / opp, ol = o - pass * opp;
        for (int k = 0; k < K; ++k) {
            const int slab = k / kI8Slab, kc = (k % kI8Slab) / 16, kb = k % 16;
            packed[((((int64_t)pass * slabs + slab) * 4 + kc) * npass + ol) * 16 + kb] = codes[(int64_t)o * K + k];
        }
    }
    return 0;
}

static int pw_i8_launch(PwI8Args &a, bool conv, int64_t G, cudaStream_t stream) {
    a.tmem_cols = 32;
    while (a.tmem_cols < a.npass) a.tmem_cols *= 2;
    const size_t smem = 2 * (size_t)kI8ASlab + (conv && a.xq ? 4 : 2) * (size_t)a.npass * kI8Slab +
                        (conv ? (size_t)a.taps * a.npass * sizeof(float) : 0);
    if (smem > 100 * 1024) return (int)cudaErrorInvalidValue;
    static bool attr_set = false;
    if (!attr_set) {
        const void *fns[] = {(const void *)pw_i8_kernel<false, false, false>, (const void *)pw_i8_kernel<true, false, false>,
                             (const void *)pw_i8_kernel<false, true, false>, (const void *)pw_i8_kernel<true, true, false>,
                             (const void *)pw_i8_kernel<false, true, true>, (const void *)pw_i8_kernel<true, true, true>,
                             (const void *)pw_i8_kernel<false, false, false, true>,
                             (const void *)pw_i8_kernel<true, false, false, true>,
                             (const void *)pw_i8_kernel<false, true, true, false, true>,
                             (const void *)pw_i8_kernel<true, true, true, false, true>};
        for (const void *fn : fns) {
            cudaError_t ce = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
            if (ce != cudaSuccess) return (int)ce;
        }
        attr_set = true;
    }
    const int64_t tiles = (G + kI8M - 1) / kI8M;
    if (tiles * a.n_pass > 0x7fffffffll || G > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    static int sms = 0;
    if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
    const bool pre = a.slabs >= 2 && tiles * a.n_pass <= (int64_t)8 * sms;
    const unsigned grid = (unsigned)(tiles * a.n_pass);
    // The persistent role-split kernel is OPT-IN (DFQ_PW_WS=<min work items per SM>, or =force for every layer):
    // at batch 256 it wins on the one tiny-K / wide-N layer (16 -> 96 at 112 x 112: 0.374 -> 0.336 ms) and loses
    // 10-140 % everywhere else -- 2 resident CTAs x 8 worker warps against 4 x 8 warps of the one-tile kernel, and
    // 64 serial loads per thread and slab.  Kept with its parity test as the base for round 2.
    static int ws_mode = -1;
    static bool ws_force = false;
    if (ws_mode < 0) {
        const char *e = getenv("DFQ_PW_WS");
        ws_force = e && e[0] == 'f';
        ws_mode = ws_force ? 1 : (e ? atoi(e) : 0);
    }
    if (!conv && ws_mode && a.npass <= 256 && (ws_force || tiles * a.n_pass >= (int64_t)ws_mode * sms)) {
        PwI8Args w = a;
        w.tiles = tiles;
        w.tmem_cols = 32;
        while (w.tmem_cols < 2 * a.npass) w.tmem_cols *= 2;            // two accumulator buffers (<= 512 columns)
        const size_t wsmem = 4 * (size_t)kI8ASlab + (size_t)kWsBStages * a.npass * kI8Slab;
        static bool ws_attr = false;
        if (!ws_attr) {
            cudaError_t ce = cudaFuncSetAttribute(pw_i8_ws_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
            if (ce != cudaSuccess) return (int)ce;
            ws_attr = true;
        }
        int per_sm = min(2, 512 / w.tmem_cols);
        int64_t g2 = min(tiles * a.n_pass, (int64_t)per_sm * sms);
        g2 = max((int64_t)a.n_pass, g2 / a.n_pass * a.n_pass);
        pw_i8_ws_kernel<<<(unsigned)g2, kWsThreads, wsmem, stream>>>(w);
        g_launches += 1;
        return (int)cudaGetLastError();
    }
    if (conv && a.xq) {
        // a 256-column accumulator admits two CTAs per SM whatever the register count, so the prefetching form's 79
        // registers cost nothing there (DFQ_CONV_PRE=0/1 overrides for A/B runs)
        static int conv_pre = -2;
        if (conv_pre == -2) { const char *e = getenv("DFQ_CONV_PRE"); conv_pre = e ? atoi(e) : -1; }
        const bool pre_q = conv_pre >= 0 ? (conv_pre != 0 && a.slabs >= 2) : (pre || (a.slabs >= 2 && a.tmem_cols >= 256));
        if (a.yq[0]) {
            if (pre_q) pw_i8_kernel<true, true, true, false, true><<<grid, kThreads, smem, stream>>>(a);
            else pw_i8_kernel<false, true, true, false, true><<<grid, kThreads, smem, stream>>>(a);
        } else if (pre_q) pw_i8_kernel<true, true, true><<<grid, kThreads, smem, stream>>>(a);
        else pw_i8_kernel<false, true, true><<<grid, kThreads, smem, stream>>>(a);
    } else if (false) {
        if (pre) pw_i8_kernel<true, true, true><<<grid, kThreads, smem, stream>>>(a);
        else pw_i8_kernel<false, true, true><<<grid, kThreads, smem, stream>>>(a);
    } else if (conv) {
        if (pre) pw_i8_kernel<true, true, false><<<grid, kThreads, smem, stream>>>(a);
        else pw_i8_kernel<false, true, false><<<grid, kThreads, smem, stream>>>(a);
    } else if (a.xs_k == 1 && a.K % 4 == 0 && a.xs_n % 4 == 0 && (a.P == 1 || a.xs_p % 4 == 0) &&
               ((uintptr_t)a.x & 15) == 0) {                           // K-contiguous rows, 16-byte aligned chunks
        if (pre) pw_i8_kernel<true, false, false, true><<<grid, kThreads, smem, stream>>>(a);
        else pw_i8_kernel<false, false, false, true><<<grid, kThreads, smem, stream>>>(a);
    } else {
        if (pre) pw_i8_kernel<true, false, false><<<grid, kThreads, smem, stream>>>(a);
        else pw_i8_kernel<false, false, false><<<grid, kThreads, smem, stream>>>(a);
    }
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_pw_i8(const DfqPwI8Desc *d, void *stream) {
    if (!d || !d->x || !d->y || !d->wpk || !d->wr || !d->in_slot || d->G <= 0 || d->bits < 1 || d->bits > 8)
        return (int)cudaErrorInvalidValue;
    PwI8Args a = {};
    int64_t bytes;
    int e = plan_cap(d->O, d->K, d->o_cap, &a.n_pass, &a.o_per_pass, &a.npass, &a.slabs, &bytes);
    if (e) return e;
    a.x = d->x; a.y = d->y; a.wpk = d->wpk; a.wr = d->wr; a.bias = d->bias; a.slot = (const RangeSlot *)d->in_slot;
    a.stats = d->out_stats; a.G = d->G;
    a.xs_n = d->xs_n; a.xs_k = d->xs_k; a.xs_p = d->xs_p; a.ys_n = d->ys_n; a.ys_o = d->ys_o; a.ys_p = d->ys_p;
    a.P = d->P; a.K = d->K; a.O = d->O; a.bits = d->bits; a.relu = d->relu;
    a.s_w = d->s_w; a.c_w = d->c_w;
    return pw_i8_launch(a, false, d->G, (cudaStream_t)stream);
}

// Reduction length of the dense k x k form: taps * (C rounded up to 16); feed it to dfq_pw_i8_plan_host /
// dfq_pw_i8_pack_host with the codes laid out [O][tap][C padded with zeros].
extern "C" int dfq_conv_i8_k_host(int C, int kh, int kw) { return kh * kw * ((C + 15) / 16 * 16); }
// Bytes of DfqConvI8Desc.scratch (the input as s8 codes, NHWC with C padded to 16).
extern "C" int64_t dfq_conv_i8_scratch_bytes(int N, int C, int H, int W) {
    return (int64_t)N * H * W * ((C + 15) / 16 * 16);
}

namespace dfq {
int conv_q8_tma(const DfqConvQ8Desc *d, void *stream);       // q8.cu
int conv_q8_halo(const DfqConvQ8Desc *d, void *stream);      // conv_halo.cu
int conv_i8_halo(const DfqConvI8Desc *d, const int8_t *codes, void *stream);
}

extern "C" int dfq_conv_i8(const DfqConvI8Desc *d, void *stream) {
    if (!d || !d->x || !d->y || !d->wpk || !d->wr || !d->wrt || !d->in_slot || d->N <= 0 || d->bits < 1 || d->bits > 8 ||
        d->kh <= 0 || d->kw <= 0 || d->kh * d->kw > 32 || d->stride <= 0 || d->dil <= 0 || d->pad < 0)
        return (int)cudaErrorInvalidValue;
    const int OH = (d->H + 2 * d->pad - d->dil * (d->kh - 1) - 1) / d->stride + 1;
    const int OW = (d->W + 2 * d->pad - d->dil * (d->kw - 1) - 1) / d->stride + 1;
    if (OH != d->OH || OW != d->OW || OH <= 0 || OW <= 0) return (int)cudaErrorInvalidValue;
    if ((int64_t)d->C * d->H * d->W > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    PwI8Args a = {};
    int64_t bytes;
    a.taps = d->kh * d->kw; a.cchunks = (d->C + 15) / 16; a.K = a.taps * a.cchunks * 16; a.Kreal = a.taps * d->C;
    int e = plan_cap(d->O, a.K, d->o_cap, &a.n_pass, &a.o_per_pass, &a.npass, &a.slabs, &bytes);
    if (e) return e;
    a.x = d->x; a.y = d->y; a.wpk = d->wpk; a.wr = d->wr; a.wrt = d->wrt; a.bias = d->bias;
    a.slot = (const RangeSlot *)d->in_slot; a.stats = d->out_stats;
    a.P = OH * OW; a.G = (int64_t)d->N * a.P;
    a.xs_n = (int64_t)d->C * d->H * d->W; a.xs_k = (int64_t)d->H * d->W; a.xs_p = 1;
    a.ys_n = (int64_t)d->O * a.P; a.ys_o = a.P; a.ys_p = 1;
    a.O = d->O; a.bits = d->bits; a.relu = d->relu; a.s_w = d->s_w; a.c_w = d->c_w;
    a.C = d->C; a.H = d->H; a.W = d->W; a.OW = OW; a.kh = d->kh; a.kw = d->kw; a.stride = d->stride; a.pad = d->pad;
    a.dil = d->dil;
    if (d->scratch) {                       // quantise once, then every tap reads 16-byte code words
        const int Cp = a.cchunks * 16, HW = d->H * d->W;
        dim3 grid((unsigned)((HW + 63) / 64), (unsigned)((Cp + 63) / 64), (unsigned)d->N);
        if (d->N > 65535 || grid.y > 65535) return (int)cudaErrorInvalidValue;
        quant_nhwc_s8_kernel<<<grid, kThreads, 0, (cudaStream_t)stream>>>(d->x, d->scratch, a.slot, d->bits, d->C, HW, Cp);
        g_launches += 1;
        a.xq = d->scratch;
        if (d->halo) return conv_i8_halo(d, d->scratch, stream);     // wpk in the halo layout (3 x 3, stride 1, pad 1)
    } else if (d->halo) {
        return (int)cudaErrorInvalidValue;
    }
    return pw_i8_launch(a, true, a.G, (cudaStream_t)stream);
}

// Static-range form (eval_q8.py): the input already IS centred s8 codes, NHWC with the channels padded to 16 (what every
// producer of the u8 plan writes), so there is no quantising pre-pass, and the output leaves as codes on the grid(s) of
// its consumer(s).  Same pipeline, same packed weights, same padding correction as dfq_conv_i8.
extern "C" int dfq_conv_q8(const DfqConvQ8Desc *d, void *stream) {
    if (!d || !d->x || !d->y0 || !d->out_slot0 || !d->wpk || !d->wr || !d->wrt || !d->in_slot || d->N <= 0 ||
        d->kh <= 0 || d->kw <= 0 || d->kh * d->kw > 32 || d->stride <= 0 || d->dil <= 0 || d->pad < 0 ||
        (d->y1 && !d->out_slot1))
        return (int)cudaErrorInvalidValue;
    const int OH = (d->H + 2 * d->pad - d->dil * (d->kh - 1) - 1) / d->stride + 1;
    const int OW = (d->W + 2 * d->pad - d->dil * (d->kw - 1) - 1) / d->stride + 1;
    if (OH != d->OH || OW != d->OW || OH <= 0 || OW <= 0) return (int)cudaErrorInvalidValue;
    if ((int64_t)d->C * d->H * d->W > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    if (d->tma == 2) return conv_q8_halo(d, stream);   // 3 x 3, stride 1, pad 1: halo tile in shared memory, no gather
    if (d->tma) {                           // wpk in the dfq_conv_q8_pack_host layout: TMA-fed, warp-specialised kernel
        if (d->stride != 1) return (int)cudaErrorInvalidValue;
        return conv_q8_tma(d, stream);
    }
    PwI8Args a = {};
    int64_t bytes;
    a.taps = d->kh * d->kw; a.cchunks = (d->C + 15) / 16; a.K = a.taps * a.cchunks * 16; a.Kreal = a.taps * d->C;
    int e = plan_cap(d->O, a.K, d->o_cap, &a.n_pass, &a.o_per_pass, &a.npass, &a.slabs, &bytes);
    if (e) return e;
    a.wpk = d->wpk; a.wr = d->wr; a.wrt = d->wrt; a.bias = d->bias;
    a.slot = (const RangeSlot *)d->in_slot;
    a.P = OH * OW; a.G = (int64_t)d->N * a.P;
    a.O = d->O; a.bits = 8; a.relu = d->relu; a.s_w = d->s_w; a.c_w = d->c_w;
    a.C = d->C; a.H = d->H; a.W = d->W; a.OW = OW; a.kh = d->kh; a.kw = d->kw; a.stride = d->stride; a.pad = d->pad;
    a.dil = d->dil;
    a.xq = d->x;
    a.yq[0] = d->y0; a.yq[1] = d->y1;
    a.oslot[0] = (const RangeSlot *)d->out_slot0; a.oslot[1] = (const RangeSlot *)d->out_slot1;
    a.cp_out = (d->O + 15) / 16 * 16;
    return pw_i8_launch(a, true, a.G, (cudaStream_t)stream);
}
