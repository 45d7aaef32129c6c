// Kernel (e), static-range form: the quantised eval forward with every QuantMeasure in `update_stat == False` mode
// (utils/quantize.py:119-126: the range is running_min / running_max as set_quant_minmax left them), so a consumer's
// grid is known before its producer runs.  Activations never exist as fp32 in HBM: every producer requantises in its
// epilogue to the grid(s) of its consumer(s) and writes CENTRED s8 CODES, NHWC with the channels padded to 16
// (a = code - 128, x_hat = s * code + min; pad bytes are s8 zeros), 4x fewer activation bytes than the dynamic-range
// plan (eval_plan.py) and no quantiser in front of the tensor cores:
//
//   pwq8_kernel   1 x 1 convolutions / linear layers: persistent, warp-specialised tcgen05 GEMM.
//                 warp 0   one lane: the A tile (128 pixels x 128 K bytes) arrives by TENSOR-MAP TMA
//                          (cp.async.bulk.tensor.2d, SWIZZLE_128B) straight in the UMMA K-major layout, the weight slab
//                          by a bulk copy of the pre-packed canonical layout, both on the stage's `full` mbarrier;
//                 warp 1   one lane: tcgen05.mma.kind::i8 (s8 x s8 -> s32) into one of TWO TMEM accumulators,
//                          tcgen05.commit frees the stage / publishes the accumulator;
//                 warps 2-9 epilogue: tcgen05.ld, y = s_x s_w acc + s_x c_w SUM_k a + c_x s_w SUM_k b + K c_x c_w + bias
//                          (SUM_k a comes out of the MMA itself: the packed weights carry a row of ones), ReLU, the
//                          consumer's exact quantiser (utils/quantize.py:72-74), 16 codes per 128-bit store; or fp32
//                          (the classifier's logits).  The epilogue of tile i overlaps the loads and MMAs of tile i+1.
//   dwq8_kernel   depthwise 3 x 3 (pad 1, stride 1 / 2) on CUDA cores: codes -> the exact fp32 values of the
//                 reference's quantiser, 9 FMAs per output in the order of the fp32 kernels, requantise, pack.
//   stemq8_kernel 3 -> O dense 3 x 3 stem from the fp32 NCHW image (quantised on load with the first measure's range).
//   addq8_kernel  residual add of two code tensors on different grids (utils/layer_transform.py:18-32).
//   meanq8_kernel spatial mean (utils/layer_transform.py:79-95) feeding the classifier's grid.
#include <cuda.h>
#include <vector>
#include "common.cuh"
#include "q8_pack.cuh"
#include "q8_ptx.cuh"

namespace dfq {
namespace {

constexpr int kQM = 128;                 // pixels per tile (UMMA M, TMEM lanes)
constexpr int kQSlab = 128;              // K bytes per ring stage (one SWIZZLE_128B row; up to 4 MMAs of K = 32)
constexpr int kQASlab = kQM * kQSlab;    // 16 KB
constexpr int kQThreads = 320;           // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue
constexpr int kQMaxStages = 8;

struct PwQ8Args {
    const int8_t *wpk; const float *wr; const float *bias; const RangeSlot *in_slot;
    const RangeSlot *out_slot[2]; int8_t *y[2]; int cp_out;
    Q8OutMap omap[2];
    float *yf; int64_t yf_g, yf_o;
    int64_t G, tiles;
    int K, Kp, O, relu, n_pass, o_per_pass, npass, slabs, tmem_cols, stages;
    float s_w, c_w;
    // dense k x k form (CONV): a tile is a BH x BW patch (= 128 pixels) of one image's output; the K loop runs tap-major
    // over `taps` x `cslabs` 128-byte channel slabs, each one 4-D TMA box of the code tensor shifted by the tap's offset
    // (out-of-image pixels arrive as zero bytes).  K above = taps * C (real reduction length).
    const float *wrt;                 // [taps][O] s_w * (sum over channels of the centred weight codes of one tap)
    int N, C, Cp, H, W, OH, OW, kh, kw, pad, dil, taps, cslabs, BW, BH, tiles_x, tiles_y;
};

template <int U> struct TmemLd;
template <> struct TmemLd<8> {
    static __device__ __forceinline__ void ld(uint32_t (&d)[8], uint32_t addr) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                     : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7])
                     : "r"(addr));
    }
};
template <> struct TmemLd<16> {
    static __device__ __forceinline__ void ld(uint32_t (&d)[16], uint32_t addr) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, "
                     "%14, %15}, [%16];"
                     : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]),
                       "=r"(d[8]), "=r"(d[9]), "=r"(d[10]), "=r"(d[11]), "=r"(d[12]), "=r"(d[13]), "=r"(d[14]), "=r"(d[15])
                     : "r"(addr));
    }
};

struct EpiCtx {
    const PwQ8Args *a; const float *chan_c; uint32_t tbase; int64_t g; int o0, n_real, half; bool valid, small_acc;
    float fa, fr, floor_v;
    uint32_t oob; const float *tap_c;      // CONV: taps of this pixel that fell in the zero padding, and what each stands for
    int gp[2];                             // row of this pixel in output 0 / 1 (plain: g; padded consumer: its interior)
};

// One accumulator row (this thread's pixel) of one tile: units of U columns alternate between the two warps of the
// lane group; SUM_k a and the first unit arrive behind ONE wait, the next unit is in flight behind the arithmetic.
template <int U, bool CONV, bool MAPPED>
__device__ __forceinline__ void epi_tile(const EpiCtx &e, const QuantConsts &q0, const QuantConsts &q1) {
    const PwQ8Args &a = *e.a;
    const bool has0 = a.y[0] != nullptr, has1 = a.y[1] != nullptr;
    uint32_t rs, d[U];
    int unit = e.half;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(rs) : "r"(e.tbase + (uint32_t)e.n_real));
    if (unit * U < e.n_real) TmemLd<U>::ld(d, e.tbase + (uint32_t)(unit * U));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    const float row_c = __fmul_rn(e.fr, (float)(int)rs);                     // s_x c_w SUM_k a[p,k]
    for (; unit * U < e.n_real; unit += 2) {
        const int nv = min(U, e.n_real - unit * U);                          // warp-uniform
        float yv[U];
        if (CONV && e.oob) {     // border pixel: zero padding is an exact 0.0 in the reference (the tap is absent),
                                 // while the s8 zero the TMA filled in stands for c_x: take that back out, tap by tap
            float corr[U];
#pragma unroll
            for (int j = 0; j < U; ++j) corr[j] = 0.0f;
            for (uint32_t mm = e.oob; mm; mm &= mm - 1) {
                const float *tc = e.tap_c + (__ffs(mm) - 1) * a.npass + unit * U;
#pragma unroll
                for (int j = 0; j < U; ++j) corr[j] = __fadd_rn(corr[j], tc[j]);
            }
#pragma unroll
            for (int j = 0; j < U; ++j)
                yv[j] = fmaxf(__fmaf_rn(e.fa, (float)(int)d[j],
                                        __fsub_rn(__fadd_rn(row_c, e.chan_c[unit * U + j]), corr[j])), e.floor_v);
        } else if (e.small_acc) {       // |acc| <= K * 2^14 < 2^22: exact conversion on the full-rate pipes (I2F is 1/4 rate)
#pragma unroll
            for (int j = 0; j < U; ++j)
                yv[j] = fmaxf(__fmaf_rn(e.fa, __fadd_rn(__int_as_float(0x4B400000 + (int)d[j]), -12582912.0f),
                                        __fadd_rn(row_c, e.chan_c[unit * U + j])), e.floor_v);
        } else {
#pragma unroll
            for (int j = 0; j < U; ++j)
                yv[j] = fmaxf(__fmaf_rn(e.fa, (float)(int)d[j], __fadd_rn(row_c, e.chan_c[unit * U + j])), e.floor_v);
        }
        if ((unit + 2) * U < e.n_real) TmemLd<U>::ld(d, e.tbase + (uint32_t)((unit + 2) * U));
        if (e.valid) {
            if (a.yf) {
                float *py = a.yf + e.g * a.yf_g + (int64_t)(e.o0 + unit * U) * a.yf_o;
#pragma unroll
                for (int j = 0; j < U; ++j)
                    if (j < nv) py[(int64_t)j * a.yf_o] = yv[j];
            }
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                if (!(t ? has1 : has0)) continue;
                const QuantConsts &qc = t ? q1 : q0;
                uint32_t w[U / 4];
                pack_codes<U>(yv, qc, w);
                if (nv < U) {
#pragma unroll
                    for (int j4 = 0; j4 < U / 4; ++j4) w[j4] = keep_left(w[j4], nv - 4 * j4);
                }
                int8_t *dst = a.y[t] + (MAPPED ? (int64_t)e.gp[t] : e.g) * a.cp_out + e.o0 + unit * U;
                if (U == 16) *reinterpret_cast<uint4 *>(dst) = make_uint4(w[0], w[1], w[U / 4 - 2], w[U / 4 - 1]);
                else *reinterpret_cast<uint2 *>(dst) = make_uint2(w[0], w[1]);
            }
        }
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    }
    // channels between the layer's last one and the padded row end that no unit covered: s8 zero
    if (e.valid && e.half == 0) {
        const int done = (e.n_real + U - 1) / U * U, row_end = min(a.o_per_pass, a.cp_out - e.o0);
        for (int cpad = done; cpad < row_end; cpad += 8) {
#pragma unroll
            for (int t = 0; t < 2; ++t)
                if (t ? has1 : has0) *reinterpret_cast<uint2 *>(a.y[t] + (MAPPED ? (int64_t)e.gp[t] : e.g) * a.cp_out + e.o0 + cpad) = make_uint2(0u, 0u);
        }
    }
}

template <bool CONV, bool MAPPED>
__global__ void __launch_bounds__(kQThreads, 2)
pwq8_kernel(const __grid_constant__ CUtensorMap tmA, PwQ8Args a) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bars[2 * kQMaxStages + 4];   // full[stages] | empty[stages] | tfull[2] | tempty[2]
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(16) float chan_c[256];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t bslab = (uint32_t)a.npass * kQSlab;
    const uint32_t sA_u = smem_u32(smem), sB_u = sA_u + (uint32_t)a.stages * kQASlab;
    const uint32_t full0 = smem_u32(&bars[0]), empty0 = full0 + 8 * kQMaxStages, tfull0 = empty0 + 8 * kQMaxStages,
                   tempty0 = tfull0 + 16;
    const int pass = blockIdx.x % a.n_pass;
    const int64_t tile0 = blockIdx.x / a.n_pass, tstep = gridDim.x / a.n_pass;
    const int o0 = pass * a.o_per_pass, n_real = min(a.o_per_pass, a.O - o0);

    if (tid == 0) {
        for (int i = 0; i < a.stages; ++i) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, 1); }
        mbar_init(tfull0, 1); mbar_init(tfull0 + 8, 1);
        mbar_init(tempty0, 8); mbar_init(tempty0 + 8, 8);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32(&tmem_base_s)), "r"((uint32_t)a.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    const RangeSlot in = *a.in_slot;
    const float sx = in.scale;
    const float cx = __fmaf_rn(128.0f, sx, in.mn);           // x_hat = sx * a + cx with a = code - 128
    if (tid < 256) {
        float c = 0.0f;
        if (tid < n_real) {
            const int o = o0 + tid;
            c = __fmaf_rn(cx, __ldg(a.wr + o), __fmul_rn(__fmul_rn((float)a.K, a.c_w), cx));
            if (a.bias) c = __fadd_rn(c, __ldg(a.bias + o));
        }
        chan_c[tid] = c;
    }
    // CONV: what one out-of-bounds tap contributed through c_x, per output channel of this pass
    float *tap_c = reinterpret_cast<float *>(smem + (size_t)a.stages * (kQASlab + bslab));
    if (CONV) {
        const float ccw = __fmul_rn(__fmul_rn((float)a.C, a.c_w), cx);
        for (int i = tid; i < a.taps * a.npass; i += kQThreads) {
            const int t = i / a.npass, ol = i - t * a.npass;
            tap_c[i] = ol < n_real ? __fmaf_rn(cx, __ldg(a.wrt + (int64_t)t * a.O + o0 + ol), ccw) : 0.0f;
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;

    if (warp == 0) {
        // ===== producer: TMA tensor-map loads of A, bulk copies of the packed weight slabs
        if (lane == 0) {
            const int8_t *wsrc = a.wpk + (int64_t)pass * a.slabs * bslab;
            int stage = 0; uint32_t phase = 0;
            for (int64_t tile = tile0; tile < a.tiles; tile += tstep) {
                int tn = 0, oh0 = 0, ow0 = 0;
                if (CONV) {
                    const int per = a.tiles_x * a.tiles_y;
                    tn = (int)(tile / per);
                    const int r = (int)(tile - (int64_t)tn * per);
                    oh0 = (r / a.tiles_x) * a.BH; ow0 = (r % a.tiles_x) * a.BW;
                }
                int tap = 0, cs = 0;
                for (int s = 0; s < a.slabs; ++s) {
                    mbar_wait(empty0 + 8 * stage, phase ^ 1);
                    mbar_expect_tx(full0 + 8 * stage, kQASlab + bslab);
                    if (CONV) {
                        const int ti = tap / a.kw, tj = tap - ti * a.kw;
                        tma_load_4d(sA_u + stage * kQASlab, &tmA, cs * kQSlab, ow0 + tj * a.dil - a.pad,
                                    oh0 + ti * a.dil - a.pad, tn, full0 + 8 * stage);
                        if (++cs == a.cslabs) { cs = 0; ++tap; }
                    } else
                    tma_load_2d(sA_u + stage * kQASlab, &tmA, s * kQSlab, (int)(tile * kQM), full0 + 8 * stage);
                    bulk_g2s(sB_u + stage * bslab, wsrc + (int64_t)s * bslab, bslab, full0 + 8 * stage);
                    if (++stage == a.stages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer
        if (lane == 0) {
            // D = s32, A = B = s8, both K-major, N = npass, M = 128
            const uint32_t idesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(a.npass >> 3) << 17) |
                                   ((uint32_t)(kQM >> 4) << 24);
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t aphase = 0;
            for (int64_t tile = tile0; tile < a.tiles; tile += tstep) {
                mbar_wait(tempty0 + 8 * acc, aphase ^ 1);
                tc_fence_after();
                const uint32_t d = tmem + (uint32_t)(acc * a.npass);
                for (int s = 0; s < a.slabs; ++s) {
                    mbar_wait(full0 + 8 * stage, phase);
                    tc_fence_after();
                    const int kb = CONV ? min(kQSlab, a.Cp - (s % a.cslabs) * kQSlab) : min(kQSlab, a.Kp - s * kQSlab);
                    const int nk = (kb + 31) >> 5;
                    for (int kk = 0; kk < nk; ++kk) {
                        const uint64_t ad = desc_sw128(sA_u + stage * kQASlab + kk * 32);
                        const uint64_t bd = desc_plain(sB_u + stage * bslab + kk * 2 * (a.npass * 16), a.npass * 16, 128);
                        umma_i8(d, ad, bd, idesc, (s | kk) != 0);
                    }
                    umma_commit(empty0 + 8 * stage);
                    if (++stage == a.stages) { stage = 0; phase ^= 1; }
                }
                umma_commit(tfull0 + 8 * acc);
                if (++acc == 2) { acc = 0; aphase ^= 1; }
            }
        }
    } else {
        // ===== epilogue: warp w reads TMEM lanes 32 * (w % 4) .., the two warps of a lane group take alternate chunks
        const int lg = warp & 3, half = (warp - 2) >> 2;
        const int row = 32 * lg + lane;
        QuantConsts q0, q1;
        const bool has0 = a.y[0] != nullptr, has1 = a.y[1] != nullptr;
        if (has0) q0 = consts_from_slot(*a.out_slot[0], 8, 0, 0.f, 0.f);
        if (has1) q1 = consts_from_slot(*a.out_slot[1], 8, 0, 0.f, 0.f);
        const float fa = __fmul_rn(sx, a.s_w), fr = __fmul_rn(sx, a.c_w);
        const float floor_v = a.relu ? 0.0f : -INFINITY;
        const bool small_acc = !CONV && a.Kp <= 255;
        int acc = 0; uint32_t aphase = 0;
        for (int64_t tile = tile0; tile < a.tiles; tile += tstep) {
            int64_t g = tile * kQM + row;
            bool valid = g < a.G;
            uint32_t oob = 0;
            if (CONV) {
                const int per = a.tiles_x * a.tiles_y;
                const int tn = (int)(tile / per);
                const int r = (int)(tile - (int64_t)tn * per);
                const int oh = (r / a.tiles_x) * a.BH + row / a.BW, ow = (r % a.tiles_x) * a.BW + row % a.BW;
                valid = oh < a.OH && ow < a.OW;
                g = ((int64_t)tn * a.OH + oh) * a.OW + ow;
                for (int t = 0; t < a.taps; ++t) {
                    const int ih = oh + (t / a.kw) * a.dil - a.pad, iw = ow + (t % a.kw) * a.dil - a.pad;
                    if (!((unsigned)ih < (unsigned)a.H && (unsigned)iw < (unsigned)a.W)) oob |= 1u << t;
                }
                if (!valid) oob = 0;
            }
            mbar_wait(tfull0 + 8 * acc, aphase);
            tc_fence_after();
            const uint32_t tbase = tmem + ((uint32_t)(32 * lg) << 16) + (uint32_t)(acc * a.npass);
            const EpiCtx ctx = {&a, chan_c, tbase, g, o0, n_real, half, valid, small_acc, fa, fr, floor_v, oob, tap_c,
                                {MAPPED && valid ? (int)q8_out_pixel(a.omap[0], g) : 0,
                                 MAPPED && valid ? (int)q8_out_pixel(a.omap[1], g) : 0}};
            if (n_real <= 16) epi_tile<8, CONV, MAPPED>(ctx, q0, q1);      // a 16-channel layer keeps all eight epilogue warps busy
            else epi_tile<16, CONV, MAPPED>(ctx, q0, q1);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty0 + 8 * acc);
            if (++acc == 2) { acc = 0; aphase ^= 1; }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)a.tmem_cols)
                     : "memory");
}

// ---------------------------------------------------------------------------------------------------------
// Depthwise 3 x 3, pad 1, stride 1 or 2, NHWC codes in and out, exact integer sums on dp4a.
//   y = s_x s_w SUM a b + s_x c_w SUM a + c_x s_w SUM_inb b + n_inb c_x c_w + bias        (a, b centred s8 codes;
// zero padding is an exact 0.0 in the reference, i.e. the tap is absent: a = 0 there, and the two constant terms count
// the in-bound taps only).
// A thread owns 4 channels (one code word per pixel) of one output row and walks it in strips of 12 input columns.
// Each 4-pixel x 4-channel block of a row is transposed in registers (8 PRMT) into one word per CHANNEL holding four
// horizontally adjacent pixels; the three taps of a row then are a byte window of that string (1 PRMT) against the
// row's packed weights (w0, w1, w2, 0): ONE dp4a per row for the dot product and one for SUM a, instead of three
// code -> fp32 conversions and three FMAs.  ~30 instructions per output, about half of the fp32 formulation.
struct DwQ8Args {
    const int8_t *x; const uint32_t *wrow; const float *bias; const RangeSlot *in_slot;
    const RangeSlot *out_slot[2]; int8_t *y[2];
    int N, H, W, OH, OW, C, Cp, stride, relu, strips;
    int segs, spp;                    // a row is walked by `segs` threads, `spp` strips each (small batches: more threads)
    float s_w, c_w;
    int pad, dil;                // generic kernel only (the dp4a kernel is pad 1, dilation 1)
};

template <int S> struct DwStrip { static constexpr int value = S == 1 ? 10 : 5; };   // outputs per 12 input columns

__device__ __forceinline__ float small_int_to_float(int v) {      // exact for |v| < 2^22, full-rate instructions
    return __fadd_rn(__int_as_float(0x4B400000 + v), -12582912.0f);
}

// a.pad is 1 (the convolution pads with zeros) or 0 (DeepLab: F.pad ran ahead of the QuantMeasure, the padded tensor exists
// as codes -- dfq_pad_q8 -- and every tap is in bounds).
// DIL = 2 (stride 1, no padding: DeepLab's dilated block): the taps of a row are bytes (k, k + 2, k + 4) of the same
// 12-byte string, eight outputs per strip.
template <int S, bool DUAL, int PAD, int DIL = 1>
__global__ void __launch_bounds__(kThreads, 2)
dwq8_kernel(DwQ8Args a) {
    constexpr int P = DIL == 2 ? 8 : DwStrip<S>::value;
    const unsigned cw = (unsigned)a.Cp >> 2;                      // channel words per pixel
    const unsigned idx = blockIdx.x * kThreads + threadIdx.x;
    if (idx >= (unsigned)a.N * (unsigned)a.OH * (unsigned)a.segs * cw) return;
    const unsigned cg = idx % cw;
    unsigned rest = idx / cw;
    const int seg = (int)(rest % (unsigned)a.segs); rest /= (unsigned)a.segs;
    const int oy = (int)(rest % (unsigned)a.OH);
    const int n = (int)(rest / (unsigned)a.OH);
    const int c0 = (int)cg * 4;
    const RangeSlot in = *a.in_slot;
    const float sx = in.scale, cx = __fmaf_rn(128.0f, sx, in.mn);
    const float fa = __fmul_rn(sx, a.s_w), fb = __fmul_rn(sx, a.c_w);
    const float cxsw = __fmul_rn(cx, a.s_w), cxcw = __fmul_rn(cx, a.c_w);
    const QuantConsts q0 = consts_from_slot(*a.out_slot[0], 8, 0, 0.f, 0.f);
    QuantConsts q1 = q0;
    if (DUAL) q1 = consts_from_slot(*a.out_slot[1], 8, 0, 0.f, 0.f);
    const int iy0 = oy * S - PAD;
    const bool rowok[3] = {(unsigned)iy0 < (unsigned)a.H, (unsigned)(iy0 + DIL) < (unsigned)a.H,
                           (unsigned)(iy0 + 2 * DIL) < (unsigned)a.H};
    uint32_t wr[3][4];                                            // [row][channel] = (w0, w1, w2, 0) centred s8
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(a.wrow + (int64_t)r * a.Cp + c0));
        wr[r][0] = rowok[r] ? v.x : 0u; wr[r][1] = rowok[r] ? v.y : 0u;
        wr[r][2] = rowok[r] ? v.z : 0u; wr[r][3] = rowok[r] ? v.w : 0u;
    }
    float b4[4];
    {
        const float4 v = a.bias ? __ldg(reinterpret_cast<const float4 *>(a.bias + c0)) : make_float4(0.f, 0.f, 0.f, 0.f);
        b4[0] = v.x; b4[1] = v.y; b4[2] = v.z; b4[3] = v.w;
    }
    // column sums of the in-bound rows' weights, per tap column j and channel: SUM_inb b for any column mask
    int rs[3][4];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int j = 0; j < 3; ++j)
            rs[j][c] = (int)(int8_t)(wr[0][c] >> (8 * j)) + (int)(int8_t)(wr[1][c] >> (8 * j)) +
                       (int)(int8_t)(wr[2][c] >> (8 * j));
    const int nrows = (int)rowok[0] + (int)rowok[1] + (int)rowok[2];
    const uint32_t ones[3] = {rowok[0] ? 0x00010101u : 0u, rowok[1] ? 0x00010101u : 0u, rowok[2] ? 0x00010101u : 0u};
    // c_x s_w SUM_inb b + n_inb c_x c_w + bias for the four column cases: all three tap columns in bounds, the left one
    // outside (first pixel of a row), the right one outside (last pixel), both (a one-pixel-wide image)
    float cst_in[4], cst_l[4], cst_r[4], cst_lr[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        auto mk = [&](int sb, int ncols) {
            return __fadd_rn(__fmaf_rn(cxsw, (float)sb, __fmul_rn((float)(ncols * nrows), cxcw)), b4[c]);
        };
        cst_in[c] = mk(rs[0][c] + rs[1][c] + rs[2][c], 3);
        cst_l[c] = mk(rs[1][c] + rs[2][c], 2);
        cst_r[c] = mk(rs[0][c] + rs[1][c], 2);
        cst_lr[c] = mk(rs[1][c], 1);
    }
    const int8_t *xn = a.x + (int64_t)n * a.H * a.W * a.Cp + c0;
    const float floor_v = a.relu ? 0.0f : -INFINITY;
    const int64_t orow = ((int64_t)n * a.OH + oy) * a.OW;
    const int left = a.C - c0;
    for (int sp = seg * a.spp; sp < min(a.strips, (seg + 1) * a.spp); ++sp) {
        const int ox0 = sp * P, ixs = ox0 * S - PAD;
        // 12 input columns x 3 rows (zero words outside the image), transposed per 4-column block
        uint32_t t[3][3][4];                                      // [row][block][channel]: 4 adjacent pixels of a channel
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const int8_t *xr = xn + (int64_t)(iy0 + r * DIL) * a.W * a.Cp;
#pragma unroll
            for (int blk = 0; blk < 3; ++blk) {
                uint32_t p[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int ix = ixs + 4 * blk + q;
                    p[q] = (rowok[r] && (unsigned)ix < (unsigned)a.W)
                               ? __ldg(reinterpret_cast<const uint32_t *>(xr + (int64_t)ix * a.Cp)) : 0u;
                }
                const uint32_t a0 = __byte_perm(p[0], p[1], 0x5140), a1 = __byte_perm(p[0], p[1], 0x7362);
                const uint32_t b0 = __byte_perm(p[2], p[3], 0x5140), b1 = __byte_perm(p[2], p[3], 0x7362);
                t[r][blk][0] = __byte_perm(a0, b0, 0x5410); t[r][blk][1] = __byte_perm(a0, b0, 0x7632);
                t[r][blk][2] = __byte_perm(a1, b1, 0x5410); t[r][blk][3] = __byte_perm(a1, b1, 0x7632);
            }
        }
#pragma unroll
        for (int k = 0; k < P; ++k) {
            const int ox = ox0 + k;
            if (ox >= a.OW) break;                                // (uniform across the warp: same strip, same k)
            constexpr int o = 0;  (void)o;
            const int off_b = k * S;                              // byte offset of the window in the 12-byte row string
            const int wi = off_b >> 2, sh = off_b & 3;
            const uint32_t sel = DIL == 2 ? ((uint32_t)sh | ((uint32_t)(sh + 2) << 4) | ((uint32_t)(sh + 4) << 8))
                                          : 0x3210u + 0x1111u * (uint32_t)sh;     // (byte 3 meets a zero weight byte)
            int acc[4], sa[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                acc[c] = 0; sa[c] = 0;
#pragma unroll
                for (int r = 0; r < 3; ++r) {
                    const uint32_t hiw = wi < 2 ? t[r][wi < 2 ? wi + 1 : 2][c] : 0u;
                    const uint32_t win = __byte_perm(t[r][wi][c], hiw, sel);
                    acc[c] = __dp4a((int)win, (int)wr[r][c], acc[c]);
                    sa[c] = __dp4a((int)win, (int)ones[r], sa[c]);
                }
            }
            const int ixl = ox * S - PAD;
            const bool lin = ixl >= 0, rin = ixl + 2 * DIL < a.W;  // left / right tap column in bounds (centre always is)
            float yv[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const float cst = lin ? (rin ? cst_in[c] : cst_r[c]) : (rin ? cst_l[c] : cst_lr[c]);
                yv[c] = fmaxf(__fmaf_rn(fa, small_int_to_float(acc[c]), __fmaf_rn(fb, small_int_to_float(sa[c]), cst)),
                              floor_v);
            }
            const int64_t off = (orow + ox) * a.Cp + c0;
            uint32_t w[1];
            pack_codes<4>(yv, q0, w);
            *reinterpret_cast<uint32_t *>(a.y[0] + off) = keep_left(w[0], left);
            if (DUAL) {
                pack_codes<4>(yv, q1, w);
                *reinterpret_cast<uint32_t *>(a.y[1] + off) = keep_left(w[0], left);
            }
        }
    }
}

// Depthwise 3 x 3 with any zero padding / dilation (DeepLab's dilated block: one layer of 960 channels at 33 x 33): one
// thread per output pixel and 4-channel word, nine word loads, the same integer sums and the same fp32 expression
// as dwq8_kernel (bit-identical where both apply), without its strip transposes.
__global__ void __launch_bounds__(kThreads)
dwq8_generic_kernel(DwQ8Args a) {
    const unsigned cw = (unsigned)a.Cp >> 2;
    const unsigned idx = blockIdx.x * kThreads + threadIdx.x;   // (the launcher refuses >= 2^31 threads)
    if (idx >= (unsigned)a.N * a.OH * a.OW * cw) return;
    const int cg = (int)(idx % cw);
    unsigned rest = idx / cw;
    const int ox = (int)(rest % (unsigned)a.OW); rest /= (unsigned)a.OW;
    const int oy = (int)(rest % (unsigned)a.OH);
    const int n = (int)(rest / (unsigned)a.OH);
    const int c0 = cg * 4;
    const RangeSlot in = *a.in_slot;
    const float sx = in.scale, cx = __fmaf_rn(128.0f, sx, in.mn);
    const float fa = __fmul_rn(sx, a.s_w), fb = __fmul_rn(sx, a.c_w);
    const float cxsw = __fmul_rn(cx, a.s_w), cxcw = __fmul_rn(cx, a.c_w);
    int acc[4] = {0, 0, 0, 0}, sa[4] = {0, 0, 0, 0}, sb[4] = {0, 0, 0, 0}, ninb = 0;
    const int8_t *xn = a.x + (int64_t)n * a.H * a.W * a.Cp + c0;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const uint4 wv = __ldg(reinterpret_cast<const uint4 *>(a.wrow + (int64_t)r * a.Cp + c0));
        const uint32_t wr[4] = {wv.x, wv.y, wv.z, wv.w};
        const int iy = oy * a.stride - a.pad + r * a.dil;
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const int ix = ox * a.stride - a.pad + j * a.dil;
            if (!((unsigned)iy < (unsigned)a.H && (unsigned)ix < (unsigned)a.W)) continue;   // zero padding: tap absent
            const uint32_t w = __ldg(reinterpret_cast<const uint32_t *>(xn + ((int64_t)iy * a.W + ix) * a.Cp));
            ++ninb;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int av = (int)(int8_t)(w >> (8 * c)), bv = (int)(int8_t)(wr[c] >> (8 * j));
                acc[c] += av * bv; sa[c] += av; sb[c] += bv;
            }
        }
    }
    const float4 bq = a.bias ? __ldg(reinterpret_cast<const float4 *>(a.bias + c0)) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float b4[4] = {bq.x, bq.y, bq.z, bq.w};
    const float floor_v = a.relu ? 0.0f : -INFINITY;
    float yv[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const float cst = __fadd_rn(__fmaf_rn(cxsw, (float)sb[c], __fmul_rn((float)ninb, cxcw)), b4[c]);
        yv[c] = fmaxf(__fmaf_rn(fa, small_int_to_float(acc[c]), __fmaf_rn(fb, small_int_to_float(sa[c]), cst)), floor_v);
    }
    const int64_t off = (((int64_t)n * a.OH + oy) * a.OW + ox) * a.Cp + c0;
    const int left = a.C - c0;
    uint32_t w[1];
    pack_codes<4>(yv, consts_from_slot(*a.out_slot[0], 8, 0, 0.f, 0.f), w);
    *reinterpret_cast<uint32_t *>(a.y[0] + off) = keep_left(w[0], left);
    if (a.y[1]) {
        pack_codes<4>(yv, consts_from_slot(*a.out_slot[1], 8, 0, 0.f, 0.f), w);
        *reinterpret_cast<uint32_t *>(a.y[1] + off) = keep_left(w[0], left);
    }
}

// ---------------------------------------------------------------------------------------------------------
// Stem: dense 3 x 3, 3 input channels, stride 2, pad 1, fp32 NCHW image in (quantised on load with the first
// measure's static range), O <= 32 output channels, codes out.  One thread per output pixel, all channels.
struct StemQ8Args {
    const float *x; const float *w; const float *bias; const RangeSlot *in_slot;
    const RangeSlot *out_slot[2]; int8_t *y[2];
    int N, H, W, OH, OW, O, Cp, stride, pad, relu;
    Q8OutMap omap[2];
    const uint32_t *wq; float s_w, c_w;       // integer form: [32][8] words of centred s8 weight codes (taps 4k .. 4k + 3)
    float *yf; unsigned *stats;               // F32 form (dynamic-range plan): fp32 NCHW out + encoded per-sample extrema
};

__global__ void __launch_bounds__(kThreads)
stemq8_kernel(StemQ8Args a) {
    __shared__ __align__(16) float sw[27][32];
    __shared__ float sb[32];
    for (int i = threadIdx.x; i < 27 * 32; i += kThreads) {
        const int t = i >> 5, o = i & 31;
        sw[t][o] = o < a.O ? __ldg(a.w + (int64_t)o * 27 + t) : 0.0f;
    }
    if (threadIdx.x < 32) sb[threadIdx.x] = (a.bias && threadIdx.x < a.O) ? __ldg(a.bias + threadIdx.x) : 0.0f;
    __syncthreads();
    const int64_t p = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    const int64_t total = (int64_t)a.N * a.OH * a.OW;
    if (p >= total) return;
    const int ox = (int)(p % a.OW);
    const int oy = (int)((p / a.OW) % a.OH);
    const int n = (int)(p / ((int64_t)a.OW * a.OH));
    const QuantConsts qi = consts_from_slot(*a.in_slot, 8, 0, 0.f, 0.f);
    float xin[27];
    bool ok[27];
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const int iy = oy * a.stride - a.pad + i, ix = ox * a.stride - a.pad + j;
                const int t = (c * 3 + i) * 3 + j;
                ok[t] = (unsigned)iy < (unsigned)a.H && (unsigned)ix < (unsigned)a.W;
                xin[t] = ok[t] ? __ldg(a.x + (((int64_t)n * 3 + c) * a.H + iy) * a.W + ix) : 0.0f;
            }
    // utils/quantize.py:72-76 on the 27 inputs: reciprocal estimates in straight-line code, one branch to the exact
    // division when any of them is within a few ulp of a tie; taps in the zero padding stay exact zeros
    if (qi.fast) {
        float t27[27], cl[27], q[27];
        bool near_tie = false;
#pragma unroll
        for (int t = 0; t < 27; ++t) {
            t27[t] = __fadd_rn(xin[t], qi.neg_min);
            cl[t] = fminf(fmaxf(__fmul_rn(t27[t], qi.rscale), qi.qmin), qi.qmax);
            q[t] = rint_small(cl[t]);
            near_tie |= fabsf(__fsub_rn(cl[t], q[t])) > qi.amb;
        }
        if (near_tie) {
#pragma unroll
            for (int t = 0; t < 27; ++t)
                if (fabsf(__fsub_rn(cl[t], q[t])) > qi.amb) {
                    const float tc = fminf(fmaxf(t27[t], qi.t_lo), qi.t_hi);
                    q[t] = rint_small(fminf(fmaxf(div_const_rn(tc, qi.scale, qi.rscale), qi.qmin), qi.qmax));
                }
        }
#pragma unroll
        for (int t = 0; t < 27; ++t) xin[t] = ok[t] ? __fadd_rn(__fmul_rn(q[t], qi.scale), qi.add_back) : 0.0f;
    } else {
#pragma unroll
        for (int t = 0; t < 27; ++t) xin[t] = ok[t] ? quant_dequant(xin[t], qi) : 0.0f;
    }
    const bool has1 = a.y[1] != nullptr;
    const QuantConsts q0 = consts_from_slot(*a.out_slot[0], 8, 0, 0.f, 0.f);
    QuantConsts q1 = q0;
    if (has1) q1 = consts_from_slot(*a.out_slot[1], 8, 0, 0.f, 0.f);
    const float floor_v = a.relu ? 0.0f : -INFINITY;
    for (int ob = 0; ob < a.Cp; ob += 16) {
        float acc[16];
#pragma unroll
        for (int t = 0; t < 16; ++t) acc[t] = 0.0f;
#pragma unroll
        for (int t = 0; t < 27; ++t) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 w = *reinterpret_cast<const float4 *>(&sw[t][ob + 4 * q]);
                acc[4 * q] = __fmaf_rn(xin[t], w.x, acc[4 * q]);
                acc[4 * q + 1] = __fmaf_rn(xin[t], w.y, acc[4 * q + 1]);
                acc[4 * q + 2] = __fmaf_rn(xin[t], w.z, acc[4 * q + 2]);
                acc[4 * q + 3] = __fmaf_rn(xin[t], w.w, acc[4 * q + 3]);
            }
        }
        float yv[16];
#pragma unroll
        for (int t = 0; t < 16; ++t) yv[t] = fmaxf(__fadd_rn(acc[t], sb[ob + t]), floor_v);
#pragma unroll
        for (int t2 = 0; t2 < 2; ++t2) {
            if (t2 && !has1) continue;
            const QuantConsts &qc = t2 ? q1 : q0;
            uint32_t w4[4];
            pack_codes<16>(yv, qc, w4);
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4) w4[j4] = keep_left(w4[j4], a.O - ob - 4 * j4);
            *reinterpret_cast<uint4 *>(a.y[t2] + q8_out_pixel(a.omap[t2], p) * a.Cp + ob) = make_uint4(w4[0], w4[1], w4[2], w4[3]);
        }
    }
}

// The stem on exact integer sums (weights on a grid of <= 8 bits, which quantize_targ_layer leaves them on): the 27
// quantised inputs of a pixel become seven words of centred s8 codes (taps in the zero padding: 0, the tap is absent in
// the reference), and an output channel is seven dp4a against its packed weights instead of 27 FMAs on dequantised
// values.  Same decomposition as the depthwise kernel:
//   y = s_x s_w SUM a b + s_x c_w SUM a + c_x s_w SUM_inb b + n_inb c_x c_w + bias.
// F32 (the dynamic-range plan, eval_plan.py): the same sums leave as fp32 NCHW with the per-sample extrema of the result
// for the next QuantMeasure (utils/quantize.py:110-111) instead of codes.
template <bool F32>
__global__ void __launch_bounds__(kThreads, 3)
stemq8_i8_kernel(StemQ8Args a) {
    __shared__ __align__(16) uint32_t swq[32][8];
    __shared__ __align__(16) float scst[32];                     // c_x s_w SUM b + 27 c_x c_w + bias (no tap in the padding)
    __shared__ float sbias[32];
    const RangeSlot in = *a.in_slot;
    const float sx = in.scale, cx = __fmaf_rn(128.0f, sx, in.mn);
    const float fa = __fmul_rn(sx, a.s_w), fb = __fmul_rn(sx, a.c_w);
    const float cxsw = __fmul_rn(cx, a.s_w), cxcw = __fmul_rn(cx, a.c_w);
    swq[threadIdx.x >> 3][threadIdx.x & 7] = __ldg(a.wq + threadIdx.x);            // kThreads == 32 * 8
    __syncthreads();
    if (threadIdx.x < 32) {
        int sb = 0;
#pragma unroll
        for (int k = 0; k < 7; ++k) sb = __dp4a((int)swq[threadIdx.x][k], 0x01010101, sb);
        const float bias = (a.bias && threadIdx.x < a.O) ? __ldg(a.bias + threadIdx.x) : 0.0f;
        sbias[threadIdx.x] = bias;
        scst[threadIdx.x] = __fadd_rn(__fmaf_rn(cxsw, small_int_to_float(sb), __fmul_rn(27.0f, cxcw)), bias);
    }
    __syncthreads();
    const int64_t total = (int64_t)a.N * a.OH * a.OW;
    const QuantConsts qi = consts_from_slot(in, 8, 0, 0.f, 0.f);
    const bool has1 = !F32 && a.y[1] != nullptr;
    QuantConsts q0 = qi, q1 = qi;
    if (!F32) q0 = consts_from_slot(*a.out_slot[0], 8, 0, 0.f, 0.f);
    if (has1) q1 = consts_from_slot(*a.out_slot[1], 8, 0, 0.f, 0.f);
    const float floor_v = a.relu ? 0.0f : -INFINITY;
    for (int64_t base = (int64_t)blockIdx.x * kThreads; base < total; base += (int64_t)gridDim.x * kThreads) {
    const bool valid = base + threadIdx.x < total;                // (the loop bound is block-uniform: F32 ends in shuffles)
    const int64_t p = valid ? base + threadIdx.x : total - 1;
    const int ox = (int)(p % a.OW);
    const int oy = (int)((p / a.OW) % a.OH);
    const int n = (int)(p / ((int64_t)a.OW * a.OH));
    float xin[28];
    bool rok[3], cok[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        rok[i] = (unsigned)(oy * a.stride - a.pad + i) < (unsigned)a.H;
        cok[i] = (unsigned)(ox * a.stride - a.pad + i) < (unsigned)a.W;
    }
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const int iy = oy * a.stride - a.pad + i, ix = ox * a.stride - a.pad + j;
                xin[(c * 3 + i) * 3 + j] = (rok[i] && cok[j]) ? __ldg(a.x + (((int64_t)n * 3 + c) * a.H + iy) * a.W + ix) : 0.0f;
            }
    xin[27] = 0.0f;
    uint32_t aw[7], keep[7];
    pack_codes<28>(xin, qi, aw);                                  // utils/quantize.py:72-74 on the inputs, centred codes
#pragma unroll
    for (int k = 0; k < 7; ++k) {
        keep[k] = 0u;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int t = 4 * k + b;                              // tap (c, i, j) = (t / 9, (t / 3) % 3, t % 3)
            if (t < 27 && rok[(t / 3) % 3] && cok[t % 3]) keep[k] |= 0xffu << (8 * b);
        }
        aw[k] &= keep[k];
    }
    const int nr = (int)rok[0] + (int)rok[1] + (int)rok[2], nc = (int)cok[0] + (int)cok[1] + (int)cok[2];
    const bool interior = nr == 3 && nc == 3;
    int sa = 0;
#pragma unroll
    for (int k = 0; k < 7; ++k) sa = __dp4a((int)aw[k], 0x01010101, sa);
    const float ninb_cxcw = __fmul_rn((float)(3 * nr * nc), cxcw);
    float lo = INFINITY, hi = -INFINITY;
    for (int ob = 0; ob < a.Cp; ob += 16) {
        float yv[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) {
            const uint4 w0 = *reinterpret_cast<const uint4 *>(&swq[ob + c][0]);
            const uint4 w1 = *reinterpret_cast<const uint4 *>(&swq[ob + c][4]);
            const uint32_t wk[7] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z};
            int acc = 0;
#pragma unroll
            for (int k = 0; k < 7; ++k) acc = __dp4a((int)aw[k], (int)wk[k], acc);
            float cst = scst[ob + c];
            if (!interior) {                                      // border pixel: the constants count the in-bound taps only
                int sb = 0;
#pragma unroll
                for (int k = 0; k < 7; ++k) sb = __dp4a((int)(keep[k] & 0x01010101u), (int)wk[k], sb);
                cst = __fadd_rn(__fmaf_rn(cxsw, small_int_to_float(sb), ninb_cxcw), sbias[ob + c]);
            }
            yv[c] = fmaxf(__fmaf_rn(fa, small_int_to_float(acc), __fmaf_rn(fb, small_int_to_float(sa), cst)), floor_v);
        }
        if (F32) {
            if (valid) {
                const int64_t plane = (int64_t)a.OH * a.OW;
                float *yp = a.yf + ((int64_t)n * a.O + ob) * plane + (int64_t)oy * a.OW + ox;
#pragma unroll
                for (int c = 0; c < 16; ++c)
                    if (ob + c < a.O) {
                        yp[c * plane] = yv[c];
                        lo = fminf(lo, yv[c]); hi = fmaxf(hi, yv[c]);
                    }
            }
            continue;
        }
        if (!valid) continue;
#pragma unroll
        for (int t2 = 0; t2 < 2; ++t2) {
            if (t2 && !has1) continue;
            const QuantConsts &qc = t2 ? q1 : q0;
            uint32_t w4[4];
            pack_codes<16>(yv, qc, w4);
#pragma unroll
            for (int j4 = 0; j4 < 4; ++j4) w4[j4] = keep_left(w4[j4], a.O - ob - 4 * j4);
            *reinterpret_cast<uint4 *>(a.y[t2] + q8_out_pixel(a.omap[t2], p) * a.Cp + ob) = make_uint4(w4[0], w4[1], w4[2], w4[3]);
        }
    }
    if (F32 && a.stats) {                                         // per-sample extrema, encoded (common.cuh)
        const int n0 = __shfl_sync(0xffffffffu, n, 0);
        if (__all_sync(0xffffffffu, valid && n == n0)) {
            lo = warp_min(lo); hi = warp_max(hi);
            if ((threadIdx.x & 31) == 0 && hi >= lo) {
                atomicMax(a.stats + 2 * n, enc_ordered(-lo));
                atomicMax(a.stats + 2 * n + 1, enc_ordered(hi));
            }
        } else if (valid && hi >= lo) {
            atomicMax(a.stats + 2 * n, enc_ordered(-lo));
            atomicMax(a.stats + 2 * n + 1, enc_ordered(hi));
        }
    }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Residual add: out = Q_a(x_a) + Q_b(x_b) from the two operands' code tensors (utils/layer_transform.py:18-32),
// requantised to the grid(s) of the consumer(s).  16 elements per thread.
__global__ void __launch_bounds__(kThreads)
addq8_kernel(const int8_t *__restrict__ xa, const int8_t *__restrict__ xb, const RangeSlot *sa, const RangeSlot *sb,
             int8_t *y0, const RangeSlot *s0, int8_t *y1, const RangeSlot *s1, int64_t n16, int C, int Cp, Q8OutMap m0,
             Q8OutMap m1) {
    const int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (i >= n16) return;
    const RangeSlot ra = *sa, rb = *sb;
    const QuantConsts q0 = consts_from_slot(*s0, 8, 0, 0.f, 0.f);
    QuantConsts q1 = q0;
    if (y1) q1 = consts_from_slot(*s1, 8, 0, 0.f, 0.f);
    const uint4 va = __ldg(reinterpret_cast<const uint4 *>(xa) + i), vb = __ldg(reinterpret_cast<const uint4 *>(xb) + i);
    const uint32_t wa[4] = {va.x, va.y, va.z, va.w}, wb[4] = {vb.x, vb.y, vb.z, vb.w};
    const int cbase = (int)((i * 16) % Cp);
    uint32_t o0[4], o1[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        float fa[4], fb[4], s[4];
        dequant4(wa[j], ra.scale, ra.mn, fa);
        dequant4(wb[j], rb.scale, rb.mn, fb);
#pragma unroll
        for (int c = 0; c < 4; ++c) s[c] = __fadd_rn(fa[c], fb[c]);
        const int left = C - cbase - 4 * j;
        const uint32_t mask = left >= 4 ? 0xffffffffu : (left <= 0 ? 0u : (0xffffffffu >> (8 * (4 - left))));
        o0[j] = pack4(s[0], s[1], s[2], s[3], q0) & mask;
        o1[j] = y1 ? (pack4(s[0], s[1], s[2], s[3], q1) & mask) : 0u;
    }
    const int64_t pix = (i * 16) / Cp;
    *reinterpret_cast<uint4 *>(y0 + q8_out_pixel(m0, pix) * Cp + cbase) = make_uint4(o0[0], o0[1], o0[2], o0[3]);
    if (y1) *reinterpret_cast<uint4 *>(y1 + q8_out_pixel(m1, pix) * Cp + cbase) = make_uint4(o1[0], o1[1], o1[2], o1[3]);
}

// Spatial mean of Q(x) (utils/layer_transform.py:79-95 around x.mean(3).mean(2) / torch.mean): codes [N][HW][Cp] ->
// fp32 mean per (n, c) in the order of dfq_qmean_f32 (32 partial sums over k = lane, lane + 32, ..., then the
// shuffle tree), requantised to the consumer's grid -> codes [N][Cp_out], and / or written as fp32 [N][C].
__global__ void __launch_bounds__(kThreads)
meanq8_kernel(const int8_t *__restrict__ x, const RangeSlot *sin, int8_t *y, const RangeSlot *sout, float *yf,
              int C, int Cp, int HW, int Cp_out, int64_t N) {
    // one thread per 4-channel word (consecutive threads read consecutive words of a pixel), pixels in order
    const int cw = Cp >> 2;
    const int64_t idx = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (idx >= N * cw) return;
    const int cg = (int)(idx % cw);
    const int64_t n = idx / cw;
    const RangeSlot in = *sin;
    const uint32_t *p = reinterpret_cast<const uint32_t *>(x + n * (int64_t)HW * Cp) + cg;
    float s[4] = {0.f, 0.f, 0.f, 0.f};
    for (int k = 0; k < HW; ++k) {
        float v[4];
        dequant4(__ldg(p + (int64_t)k * cw), in.scale, in.mn, v);
#pragma unroll
        for (int c = 0; c < 4; ++c) s[c] = __fadd_rn(s[c], v[c]);
    }
    float m[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) m[c] = __fdiv_rn(s[c], (float)HW);
    const int c0 = cg * 4;
    if (yf) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if (c0 + c < C) yf[n * C + c0 + c] = m[c];
    }
    if (y) {
        const QuantConsts qc = consts_from_slot(*sout, 8, 0, 0.f, 0.f);
        uint32_t w[1];
        pack_codes<4>(m, qc, w);
        *reinterpret_cast<uint32_t *>(y + n * Cp_out + c0) = keep_left(w[0], C - c0);
    }
}

}  // namespace
}  // namespace dfq

using namespace dfq;

// ---- host side --------------------------------------------------------------------------------------------------
static int q8_plan(int O, int K, int o_cap, int *n_pass, int *o_per_pass, int *npass, int *slabs, int64_t *bytes) {
    if (O <= 0 || K <= 0) return (int)cudaErrorInvalidValue;
    int cap = o_cap > 0 ? (o_cap + 15) / 16 * 16 : 240;      // 240 real channels + the row of ones = 256 columns
    if (cap > 240) cap = 240;
    if (cap < 16) cap = 16;
    int np = (O + cap - 1) / cap;
    int opp = ((O + np - 1) / np + 15) / 16 * 16;             // equal passes, multiples of 16
    np = (O + opp - 1) / opp;
    const int Kp = (K + 15) / 16 * 16;
    *n_pass = np; *o_per_pass = opp;
    *npass = (opp + 1 + 15) / 16 * 16;                        // + the row of ones
    *slabs = (Kp + kQSlab - 1) / kQSlab;
    *bytes = (int64_t)np * *slabs * *npass * kQSlab;
    return 0;
}

extern "C" int dfq_pw_q8_plan_host(int O, int K, int o_cap, int *n_pass, int *o_per_pass, int *npass, int *slabs,
                                   int64_t *packed_bytes) {
    return q8_plan(O, K, o_cap, n_pass, o_per_pass, npass, slabs, packed_bytes);
}

// codes: [O, K] centred weight codes (code - 128).  packed: [pass][slab][K chunk of 16 (8 per slab)][row][16 B], zero
// padded; row n_real of every pass is the row of ones over the real K (its accumulator column is SUM_k a[p, k]).
extern "C" int dfq_pw_q8_pack_host(const int8_t *codes, int O, int K, int o_cap, int8_t *packed) {
    int np, opp, npass, slabs; int64_t bytes;
    int e = q8_plan(O, K, o_cap, &np, &opp, &npass, &slabs, &bytes);
    if (e) return e;
    for (int64_t i = 0; i < bytes; ++i) packed[i] = 0;
    auto at = [&](int pass, int row, int k) -> int8_t & {
        const int slab = k / kQSlab, kc = (k % kQSlab) / 16, kb = k % 16;
        return packed[((((int64_t)pass * slabs + slab) * 8 + kc) * npass + row) * 16 + kb];
    };
    for (int o = 0; o < O; ++o) {
        const int pass = o / opp, ol = o - pass * opp;
        for (int k = 0; k < K; ++k) at(pass, ol, k) = codes[(int64_t)o * K + k];
    }
    for (int pass = 0; pass < np; ++pass) {
        const int n_real = (O - pass * opp) < opp ? (O - pass * opp) : opp;
        for (int k = 0; k < K; ++k) at(pass, n_real, k) = 1;
    }
    return 0;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
    }
    return fn;
}

extern "C" int dfq_pw_q8(const DfqPwQ8Desc *d, void *stream) {
    if (!d || !d->x || !d->wpk || !d->wr || !d->in_slot || d->G <= 0 || (!d->y0 && !d->yf) || (d->y1 && !d->y0) ||
        (d->y0 && !d->out_slot0) || (d->y1 && !d->out_slot1))
        return (int)cudaErrorInvalidValue;
    PwQ8Args a = {};
    int64_t bytes;
    int e = q8_plan(d->O, d->K, d->o_cap, &a.n_pass, &a.o_per_pass, &a.npass, &a.slabs, &bytes);
    if (e) return e;
    a.wpk = d->wpk; a.wr = d->wr; a.bias = d->bias; a.in_slot = (const RangeSlot *)d->in_slot;
    a.out_slot[0] = (const RangeSlot *)d->out_slot0; a.out_slot[1] = (const RangeSlot *)d->out_slot1;
    a.y[0] = d->y0; a.y[1] = d->y1; a.cp_out = (d->O + 15) / 16 * 16;
    a.omap[0] = q8_make_map(&d->map0); a.omap[1] = q8_make_map(&d->map1);
    a.yf = d->yf; a.yf_g = d->yf_g; a.yf_o = d->yf_o;
    a.G = d->G; a.tiles = (d->G + kQM - 1) / kQM;
    a.K = d->K; a.Kp = (d->K + 15) / 16 * 16; a.O = d->O; a.relu = d->relu; a.s_w = d->s_w; a.c_w = d->c_w;
    if (a.tiles * kQM > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    a.tmem_cols = 32;
    while (a.tmem_cols < 2 * a.npass) a.tmem_cols *= 2;
    const size_t stage_bytes = (size_t)kQASlab + (size_t)a.npass * kQSlab;
    // CTAs per SM: as many as the TMEM allows (each holds tmem_cols of the 512 columns; a CTA that could not get its
    // columns would spin in tcgen05.alloc), at most two; the shared-memory request is sized / padded to enforce it
    const int ctas_tmem = 512 / a.tmem_cols;
    const int want_ctas = ctas_tmem < 2 ? 1 : 2;
    const size_t smem_cap = 227 * 1024 - 2048;                 // dynamic part: the kernel has ~1.2 KB of static shared memory
    const size_t budget = smem_cap / (size_t)want_ctas - 1024;
    int stages = (int)((budget - 1024) / stage_bytes);
    if (stages > kQMaxStages) stages = kQMaxStages;
    const int64_t need = (int64_t)a.slabs * 2 < 6 ? 6 : (int64_t)a.slabs * 2;   // single-slab layers: A tiles of the next tiles prefetched
    if (stages > need) stages = (int)need;
    if (stages < 2) return (int)cudaErrorInvalidValue;
    a.stages = stages;
    size_t smem = (size_t)stages * stage_bytes + 1024;
    {
        const size_t floor_bytes = smem_cap / (size_t)(want_ctas + 1) + 1024;
        if (smem < floor_bytes) smem = floor_bytes;
    }
    static size_t configured = 0;
    if (configured < smem) {
        cudaError_t ce = cudaFuncSetAttribute(pwq8_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_cap);
        if (ce == cudaSuccess)
            ce = cudaFuncSetAttribute(pwq8_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_cap);
        if (ce != cudaSuccess) return (int)ce;
        configured = smem_cap;
    }
    EncodeTiledFn enc = encode_tiled();
    if (!enc) return (int)cudaErrorNotSupported;
    CUtensorMap tm;
    const cuuint64_t gdim[2] = {(cuuint64_t)a.Kp, (cuuint64_t)d->G};
    const cuuint64_t gstr[1] = {(cuuint64_t)a.Kp};
    const cuuint32_t box[2] = {(cuuint32_t)kQSlab, (cuuint32_t)kQM};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, (void *)d->x, gdim, gstr, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return (int)cudaErrorInvalidValue;
    int occ = 0;
    cudaError_t ce = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, pwq8_kernel<false, false>, kQThreads, smem);
    if (ce != cudaSuccess) return (int)ce;
    if (occ < 1) return (int)cudaErrorLaunchOutOfResources;
    if (occ > want_ctas) occ = want_ctas;
    int64_t per_pass = ((int64_t)sm_count() * occ) / a.n_pass;
    if (per_pass < 1) per_pass = 1;
    if (per_pass > a.tiles) per_pass = a.tiles;
    const unsigned grid = (unsigned)(per_pass * a.n_pass);
    if (a.omap[0].Wp || a.omap[1].Wp) pwq8_kernel<false, true><<<grid, kQThreads, smem, (cudaStream_t)stream>>>(tm, a);
    else pwq8_kernel<false, false><<<grid, kQThreads, smem, (cudaStream_t)stream>>>(tm, a);
    g_launches += 1;
    return (int)cudaGetLastError();
}

// ---- dense k x k on the same pipeline (stride 1) ------------------------------------------------------------------
// Packed reduction order: tap-major, then 128-byte channel slabs (channels zero-padded to a multiple of 128 per tap), so
// that one ring stage = one TMA box of one tap.
extern "C" int dfq_conv_q8_plan_host(int O, int C, int taps, int o_cap, int *kq, int64_t *packed_bytes) {
    if (O <= 0 || C <= 0 || taps <= 0 || taps > 25) return (int)cudaErrorInvalidValue;
    const int Cp = (C + 15) / 16 * 16, cslabs = (Cp + kQSlab - 1) / kQSlab;
    *kq = taps * cslabs * kQSlab;
    int np, opp, npass, slabs;
    return q8_plan(O, *kq, o_cap, &np, &opp, &npass, &slabs, packed_bytes);
}

// codes: [O][taps][C padded to 16] centred weight codes
extern "C" int dfq_conv_q8_pack_host(const int8_t *codes, int O, int C, int taps, int o_cap, int8_t *packed) {
    int kq; int64_t bytes;
    int e = dfq_conv_q8_plan_host(O, C, taps, o_cap, &kq, &bytes);
    if (e) return e;
    const int Cp = (C + 15) / 16 * 16, per_tap = kq / taps;
    std::vector<int8_t> lay((size_t)O * kq, 0);
    for (int o = 0; o < O; ++o)
        for (int t = 0; t < taps; ++t)
            for (int c = 0; c < Cp; ++c)
                lay[((size_t)o * taps + t) * per_tap + c] = codes[((size_t)o * taps + t) * Cp + c];
    return dfq_pw_q8_pack_host(lay.data(), O, kq, o_cap, packed);
}

namespace dfq {
int conv_q8_tma(const DfqConvQ8Desc *d, void *stream) {
    PwQ8Args a = {};
    a.taps = d->kh * d->kw;
    a.C = d->C; a.Cp = (d->C + 15) / 16 * 16; a.cslabs = (a.Cp + kQSlab - 1) / kQSlab;
    const int kq = a.taps * a.cslabs * kQSlab;
    int64_t bytes;
    int e = q8_plan(d->O, kq, d->o_cap, &a.n_pass, &a.o_per_pass, &a.npass, &a.slabs, &bytes);
    if (e) return e;
    if (a.slabs != a.taps * a.cslabs || a.taps > 25) return (int)cudaErrorInvalidValue;
    a.wpk = d->wpk; a.wr = d->wr; a.wrt = d->wrt; a.bias = d->bias; a.in_slot = (const RangeSlot *)d->in_slot;
    a.out_slot[0] = (const RangeSlot *)d->out_slot0; a.out_slot[1] = (const RangeSlot *)d->out_slot1;
    a.y[0] = d->y0; a.y[1] = d->y1; a.cp_out = (d->O + 15) / 16 * 16;
    a.K = a.taps * d->C; a.Kp = kq; a.O = d->O; a.relu = d->relu; a.s_w = d->s_w; a.c_w = d->c_w;
    a.N = d->N; a.H = d->H; a.W = d->W; a.OH = d->OH; a.OW = d->OW; a.kh = d->kh; a.kw = d->kw; a.pad = d->pad; a.dil = d->dil;
    // the 128-pixel tile as a BH x BW patch: the shape that wastes the fewest pixels past the image's edges
    int64_t best = -1;
    for (int bw = 128; bw >= 8; bw >>= 1) {
        const int bh = 128 / bw;
        const int64_t t = (int64_t)((d->OW + bw - 1) / bw) * ((d->OH + bh - 1) / bh);
        if (best < 0 || t < best) { best = t; a.BW = bw; a.BH = bh; }
    }
    a.tiles_x = (d->OW + a.BW - 1) / a.BW; a.tiles_y = (d->OH + a.BH - 1) / a.BH;
    a.tiles = (int64_t)d->N * a.tiles_x * a.tiles_y;
    a.G = (int64_t)d->N * d->OH * d->OW;
    if (a.tiles > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    a.tmem_cols = 32;
    while (a.tmem_cols < 2 * a.npass) a.tmem_cols *= 2;
    if (a.tmem_cols > 512) return (int)cudaErrorInvalidValue;
    const size_t stage_bytes = (size_t)kQASlab + (size_t)a.npass * kQSlab;
    const size_t tap_bytes = (size_t)a.taps * a.npass * sizeof(float);
    const int want_ctas = 512 / a.tmem_cols < 2 ? 1 : 2;
    const size_t smem_cap = 227 * 1024 - 2048;
    const size_t budget = smem_cap / (size_t)want_ctas - 1024 - tap_bytes;
    int stages = (int)((budget - 1024) / stage_bytes);
    if (stages > kQMaxStages) stages = kQMaxStages;
    if (stages < 2) return (int)cudaErrorInvalidValue;
    a.stages = stages;
    size_t smem = (size_t)stages * stage_bytes + tap_bytes + 1024;
    {
        const size_t floor_bytes = smem_cap / (size_t)(want_ctas + 1) + 1024;
        if (smem < floor_bytes) smem = floor_bytes;
    }
    static bool configured = false;
    if (!configured) {
        cudaError_t ce = cudaFuncSetAttribute(pwq8_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_cap);
        if (ce != cudaSuccess) return (int)ce;
        configured = true;
    }
    EncodeTiledFn enc = encode_tiled();
    if (!enc) return (int)cudaErrorNotSupported;
    CUtensorMap tm;
    const cuuint64_t gdim[4] = {(cuuint64_t)a.Cp, (cuuint64_t)d->W, (cuuint64_t)d->H, (cuuint64_t)d->N};
    const cuuint64_t gstr[3] = {(cuuint64_t)a.Cp, (cuuint64_t)a.Cp * d->W, (cuuint64_t)a.Cp * d->W * d->H};
    const cuuint32_t box[4] = {(cuuint32_t)kQSlab, (cuuint32_t)a.BW, (cuuint32_t)a.BH, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, (void *)d->x, gdim, gstr, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return (int)cudaErrorInvalidValue;
    int occ = 0;
    cudaError_t ce = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, pwq8_kernel<true, false>, kQThreads, smem);
    if (ce != cudaSuccess) return (int)ce;
    if (occ < 1) return (int)cudaErrorLaunchOutOfResources;
    if (occ > want_ctas) occ = want_ctas;
    int64_t per_pass = ((int64_t)sm_count() * occ) / a.n_pass;
    if (per_pass < 1) per_pass = 1;
    if (per_pass > a.tiles) per_pass = a.tiles;
    pwq8_kernel<true, false><<<(unsigned)(per_pass * a.n_pass), kQThreads, smem, (cudaStream_t)stream>>>(tm, a);
    g_launches += 1;
    return (int)cudaGetLastError();
}
}  // namespace dfq

// Threads per output row: one walks the whole row when the launch fills the GPU anyway (batch 256); at small batches
// (DeepLab: 16 images, 33 .. 257 rows) a row is cut into segments of whole strips until there are about three waves of
// blocks (2 resident per SM).  DFQ_Q8_DW_SEGS=<n> forces a count (A/B runs; 1 = the one-thread-per-row form).
static void dw_row_segments(DwQ8Args &a) {
    const int64_t rows = (int64_t)a.N * a.OH * (a.Cp / 4);
    const int64_t want = (int64_t)sm_count() * 2 * kThreads * 3;
    int64_t segs = (want + rows - 1) / rows;
    const char *e = getenv("DFQ_Q8_DW_SEGS");
    if (e && atoi(e) > 0) segs = atoi(e);
    if (segs > a.strips) segs = a.strips;
    if (segs < 1) segs = 1;
    a.spp = (int)((a.strips + segs - 1) / segs);
    a.segs = (a.strips + a.spp - 1) / a.spp;
}

static bool dw_dil2_off() {                       // DFQ_Q8_DW_DIL2=0: dilated layers on the generic kernel (A/B runs)
    const char *e = getenv("DFQ_Q8_DW_DIL2");
    return e && e[0] == '0';
}

extern "C" int dfq_dw_q8(const DfqDwQ8Desc *d, void *stream) {
    if (!d || !d->x || !d->wrow || !d->in_slot || !d->y0 || !d->out_slot0 || (d->y1 && !d->out_slot1) || d->N <= 0 ||
        (d->stride != 1 && d->stride != 2) || d->pad < 0 || d->dil < 0)
        return (int)cudaErrorInvalidValue;
    DwQ8Args a = {};
    a.x = d->x; a.wrow = d->wrow; a.bias = d->bias; a.in_slot = (const RangeSlot *)d->in_slot;
    a.out_slot[0] = (const RangeSlot *)d->out_slot0; a.out_slot[1] = (const RangeSlot *)d->out_slot1;
    a.y[0] = d->y0; a.y[1] = d->y1;
    a.N = d->N; a.H = d->H; a.W = d->W; a.C = d->C; a.Cp = (d->C + 15) / 16 * 16; a.stride = d->stride; a.relu = d->relu;
    a.s_w = d->s_w; a.c_w = d->c_w;
    a.dil = d->dil > 0 ? d->dil : 1;
    a.pad = d->pad;
    a.OH = (d->H + 2 * a.pad - 2 * a.dil - 1) / d->stride + 1; a.OW = (d->W + 2 * a.pad - 2 * a.dil - 1) / d->stride + 1;
    if (a.OH <= 0 || a.OW <= 0) return (int)cudaErrorInvalidValue;
    const cudaStream_t st = (cudaStream_t)stream;
    if (a.dil == 2 && a.pad == 0 && d->stride == 1 && !dw_dil2_off()) {      // the strip kernel on dilated taps
        a.strips = (a.OW + 7) / 8;
        dw_row_segments(a);
        const int64_t total = (int64_t)a.N * a.OH * a.segs * (a.Cp / 4);
        if (total >= 0x7fffffffll) return (int)cudaErrorInvalidValue;
        const unsigned grid = (unsigned)((total + kThreads - 1) / kThreads);
        if (d->y1) dwq8_kernel<1, true, 0, 2><<<grid, kThreads, 0, st>>>(a);
        else dwq8_kernel<1, false, 0, 2><<<grid, kThreads, 0, st>>>(a);
        g_launches += 1;
        return (int)cudaGetLastError();
    }
    if (a.pad > 1 || a.dil != 1) {
        const int64_t total = (int64_t)a.N * a.OH * a.OW * (a.Cp / 4);
        const int64_t blocks = (total + kThreads - 1) / kThreads;
        if (blocks >= (1ll << 23)) return (int)cudaErrorInvalidValue;
        dwq8_generic_kernel<<<(unsigned)blocks, kThreads, 0, st>>>(a);
        g_launches += 1;
        return (int)cudaGetLastError();
    }
    const int P = d->stride == 1 ? DwStrip<1>::value : DwStrip<2>::value;
    a.strips = (a.OW + P - 1) / P;
    dw_row_segments(a);
    const int64_t total = (int64_t)a.N * a.OH * a.segs * (a.Cp / 4);
    if (total >= 0x7fffffffll) return (int)cudaErrorInvalidValue;
    const unsigned grid = (unsigned)((total + kThreads - 1) / kThreads);
    const int variant = (d->stride == 2 ? 4 : 0) | (d->y1 ? 2 : 0) | (a.pad ? 1 : 0);
    switch (variant) {
    case 0: dwq8_kernel<1, false, 0><<<grid, kThreads, 0, st>>>(a); break;
    case 1: dwq8_kernel<1, false, 1><<<grid, kThreads, 0, st>>>(a); break;
    case 2: dwq8_kernel<1, true, 0><<<grid, kThreads, 0, st>>>(a); break;
    case 3: dwq8_kernel<1, true, 1><<<grid, kThreads, 0, st>>>(a); break;
    case 4: dwq8_kernel<2, false, 0><<<grid, kThreads, 0, st>>>(a); break;
    case 5: dwq8_kernel<2, false, 1><<<grid, kThreads, 0, st>>>(a); break;
    case 6: dwq8_kernel<2, true, 0><<<grid, kThreads, 0, st>>>(a); break;
    default: dwq8_kernel<2, true, 1><<<grid, kThreads, 0, st>>>(a); break;
    }
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_stem_q8(const DfqStemQ8Desc *d, void *stream) {
    if (!d || !d->x || !d->w || !d->in_slot || (d->y1 && !d->out_slot1) || d->N <= 0 || d->O < 1 || d->O > 32 ||
        (d->yf ? (d->y0 || d->y1 || !d->wq) : (!d->y0 || !d->out_slot0)))
        return (int)cudaErrorInvalidValue;
    StemQ8Args a = {};
    a.x = d->x; a.w = d->w; a.bias = d->bias; a.in_slot = (const RangeSlot *)d->in_slot;
    a.out_slot[0] = (const RangeSlot *)d->out_slot0; a.out_slot[1] = (const RangeSlot *)d->out_slot1;
    a.y[0] = d->y0; a.y[1] = d->y1;
    a.N = d->N; a.H = d->H; a.W = d->W; a.O = d->O; a.Cp = (d->O + 15) / 16 * 16; a.stride = d->stride; a.pad = d->pad;
    a.relu = d->relu;
    a.OH = (d->H + 2 * d->pad - 3) / d->stride + 1; a.OW = (d->W + 2 * d->pad - 3) / d->stride + 1;
    a.omap[0] = q8_make_map(&d->map0); a.omap[1] = q8_make_map(&d->map1);
    const int64_t total = (int64_t)a.N * a.OH * a.OW;
    a.wq = reinterpret_cast<const uint32_t *>(d->wq); a.s_w = d->s_w; a.c_w = d->c_w;
    static_assert(kThreads == 256, "stemq8_i8_kernel stages its 32 x 8 weight words with one load per thread");
    if (a.wq) {             // blocks walk the pixels grid-stride: the weight words are staged once per resident block
        const int64_t blocks = (total + kThreads - 1) / kThreads, resident = (int64_t)sm_count() * 3 * 4;
        const unsigned grid = (unsigned)(blocks < resident ? blocks : resident);
        a.yf = d->yf; a.stats = d->out_stats;
        if (a.yf) stemq8_i8_kernel<true><<<grid, kThreads, 0, (cudaStream_t)stream>>>(a);
        else stemq8_i8_kernel<false><<<grid, kThreads, 0, (cudaStream_t)stream>>>(a);
    }
    else
    stemq8_kernel<<<(unsigned)((total + kThreads - 1) / kThreads), kThreads, 0, (cudaStream_t)stream>>>(a);
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_add_q8(const int8_t *xa, const int8_t *xb, const float *slot_a, const float *slot_b, int8_t *y0,
                          const float *out_slot0, int8_t *y1, const float *out_slot1, int64_t pixels, int C,
                          const DfqQ8OutMap *map0, const DfqQ8OutMap *map1, void *stream) {
    if (!xa || !xb || !slot_a || !slot_b || !y0 || !out_slot0 || (y1 && !out_slot1) || pixels <= 0 || C <= 0)
        return (int)cudaErrorInvalidValue;
    const int Cp = (C + 15) / 16 * 16;
    const int64_t n16 = pixels * (Cp / 16);
    addq8_kernel<<<(unsigned)((n16 + kThreads - 1) / kThreads), kThreads, 0, (cudaStream_t)stream>>>(
        xa, xb, (const RangeSlot *)slot_a, (const RangeSlot *)slot_b, y0, (const RangeSlot *)out_slot0, y1,
        (const RangeSlot *)out_slot1, n16, C, Cp, q8_make_map(map0), q8_make_map(map1));
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_mean_q8(const int8_t *x, const float *in_slot, int8_t *y, const float *out_slot, float *yf,
                           int64_t N, int C, int HW, void *stream) {
    if (!x || !in_slot || (!y && !yf) || (y && !out_slot) || N <= 0 || C <= 0 || HW <= 0)
        return (int)cudaErrorInvalidValue;
    const int Cp = (C + 15) / 16 * 16;
    const int64_t total = N * (Cp / 4);
    meanq8_kernel<<<(unsigned)((total + kThreads - 1) / kThreads), kThreads, 0, (cudaStream_t)stream>>>(
        x, (const RangeSlot *)in_slot, y, (const RangeSlot *)out_slot, yf, C, Cp, HW, Cp, N);
    g_launches += 1;
    return (int)cudaGetLastError();
}
