// Kernel (e), static-range form, data-movement ops on code tensors (what DeepLabV3+ adds to MobileNetV2's graph):
//   interpq8_kernel  F.interpolate(Q(x), size, 'bilinear', align_corners) as patched by utils/layer_transform.py:98-110:
//                    the operand's codes are dequantised (they ARE Q(x)), blended with ATen's index / weight arithmetic,
//                    and leave on the consumer's grid -- or as fp32 NCHW (the network's logits);
//   regridq8_kernel  one operand of torch.cat (utils/layer_transform.py:112-125): every operand has its own QuantMeasure,
//                    the concatenation is quantised again by the consuming convolution's: codes on grid A -> values ->
//                    codes on grid B, written at the operand's channel offset of the wider NHWC tensor;
//   padq8_kernel     F.pad ahead of a QuantMeasure (never quantised itself: utils/layer_transform.py ignore_op=['pad']; the
//                    padded zeros are quantised by the consumer's measure, so the border is code(0.0) on its grid);
//   poolq8_kernel    nn.AdaptiveAvgPool2d(1) (not a patched op: it sees the RAW fp32 tensor, which the producing 1 x 1
//                    convolution writes next to its codes) -> codes on the grid of the following convolution.
// Activation format: centred s8 codes, NHWC, channels padded to 16 with zero bytes (see q8.cu).
#include "common.cuh"
#include "q8_pack.cuh"

namespace dfq {
namespace {

// ATen UpSample.h area_pixel_compute_source_index + the two-tap weights (same as move.cu)
__device__ __forceinline__ void q8_interp_index(int o, float scale, int size, int align, int &i0, int &i1, float &l1) {
    const float s = align ? __fmul_rn(scale, (float)o)
                          : fmaxf(__fsub_rn(__fmul_rn(scale, __fadd_rn((float)o, 0.5f)), 0.5f), 0.0f);
    i0 = min((int)s, size - 1);
    i1 = i0 + (i0 < size - 1 ? 1 : 0);
    l1 = fminf(fmaxf(__fsub_rn(s, (float)i0), 0.0f), 1.0f);
}

struct InterpQ8Args {
    const int8_t *x; const RangeSlot *in_slot;
    int8_t *y[2]; const RangeSlot *out_slot[2]; float *yf;
    int N, C, Cp, H, W, OH, OW, align;
    float sy, sx;
};

// the four values of one channel word at output pixel (oy, ox) of image n
__device__ __forceinline__ void interp_word(const InterpQ8Args &a, const int8_t *xn, int cwi, int y0, int y1, float l1y,
                                            int x0, int x1, float l1x, float s, float mn, float *v) {
    const int64_t rs = (int64_t)a.W * a.Cp;
    const uint32_t w00 = __ldg(reinterpret_cast<const uint32_t *>(xn + y0 * rs + (int64_t)x0 * a.Cp) + cwi);
    const uint32_t w01 = __ldg(reinterpret_cast<const uint32_t *>(xn + y0 * rs + (int64_t)x1 * a.Cp) + cwi);
    const uint32_t w10 = __ldg(reinterpret_cast<const uint32_t *>(xn + y1 * rs + (int64_t)x0 * a.Cp) + cwi);
    const uint32_t w11 = __ldg(reinterpret_cast<const uint32_t *>(xn + y1 * rs + (int64_t)x1 * a.Cp) + cwi);
    float a00[4], a01[4], a10[4], a11[4];
    dequant4(w00, s, mn, a00); dequant4(w01, s, mn, a01); dequant4(w10, s, mn, a10); dequant4(w11, s, mn, a11);
    const float l0x = __fsub_rn(1.0f, l1x), l0y = __fsub_rn(1.0f, l1y);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const float top = __fadd_rn(__fmul_rn(l0x, a00[c]), __fmul_rn(l1x, a01[c]));
        const float bot = __fadd_rn(__fmul_rn(l0x, a10[c]), __fmul_rn(l1x, a11[c]));
        v[c] = __fadd_rn(__fmul_rn(l0y, top), __fmul_rn(l1y, bot));
    }
}

// codes out: one thread per (pixel, 16-channel chunk), chunks fastest: four 128-bit corner loads, one 128-bit store per
// grid, the indices and weights computed once per 16 channels
__global__ void __launch_bounds__(kThreads)
interpq8_codes_kernel(InterpQ8Args a) {
    const unsigned chunks = (unsigned)a.Cp >> 4;
    const unsigned idx = blockIdx.x * kThreads + threadIdx.x;   // 32-bit index math (the launcher refuses >= 2^31 threads):
    if (idx >= (unsigned)a.N * a.OH * a.OW * chunks) return;     // the 64-bit divisions were most of this kernel
    const int ch = (int)(idx % chunks);
    unsigned rest = idx / chunks;
    const int ox = (int)(rest % (unsigned)a.OW); rest /= (unsigned)a.OW;
    const int oy = (int)(rest % (unsigned)a.OH);
    const int n = (int)(rest / (unsigned)a.OH);
    int y0, y1, x0, x1; float l1y, l1x;
    q8_interp_index(oy, a.sy, a.H, a.align, y0, y1, l1y);
    q8_interp_index(ox, a.sx, a.W, a.align, x0, x1, l1x);
    const RangeSlot in = *a.in_slot;
    const int8_t *xn = a.x + (int64_t)n * a.H * a.W * a.Cp;
    const int64_t rs = (int64_t)a.W * a.Cp;
    const uint4 q00 = __ldg(reinterpret_cast<const uint4 *>(xn + y0 * rs + (int64_t)x0 * a.Cp) + ch);
    const uint4 q01 = __ldg(reinterpret_cast<const uint4 *>(xn + y0 * rs + (int64_t)x1 * a.Cp) + ch);
    const uint4 q10 = __ldg(reinterpret_cast<const uint4 *>(xn + y1 * rs + (int64_t)x0 * a.Cp) + ch);
    const uint4 q11 = __ldg(reinterpret_cast<const uint4 *>(xn + y1 * rs + (int64_t)x1 * a.Cp) + ch);
    const uint32_t w00[4] = {q00.x, q00.y, q00.z, q00.w}, w01[4] = {q01.x, q01.y, q01.z, q01.w};
    const uint32_t w10[4] = {q10.x, q10.y, q10.z, q10.w}, w11[4] = {q11.x, q11.y, q11.z, q11.w};
    const float l0x = __fsub_rn(1.0f, l1x), l0y = __fsub_rn(1.0f, l1y);
    float v[16];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        float a00[4], a01[4], a10[4], a11[4];
        dequant4(w00[k], in.scale, in.mn, a00); dequant4(w01[k], in.scale, in.mn, a01);
        dequant4(w10[k], in.scale, in.mn, a10); dequant4(w11[k], in.scale, in.mn, a11);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const float top = __fadd_rn(__fmul_rn(l0x, a00[c]), __fmul_rn(l1x, a01[c]));
            const float bot = __fadd_rn(__fmul_rn(l0x, a10[c]), __fmul_rn(l1x, a11[c]));
            v[4 * k + c] = __fadd_rn(__fmul_rn(l0y, top), __fmul_rn(l1y, bot));
        }
    }
    const int64_t off = (((int64_t)n * a.OH + oy) * a.OW + ox) * a.Cp + 16 * ch;
    const int left = a.C - 16 * ch;
    for (int gi = 0; gi < 2; ++gi) {
        if (!a.y[gi]) break;
        uint32_t w[4];
        pack_codes<16>(v, consts_from_slot(*a.out_slot[gi], 8, 0, 0.f, 0.f), w);
        *reinterpret_cast<uint4 *>(a.y[gi] + off) = make_uint4(keep_left(w[0], left), keep_left(w[1], left - 4),
                                                               keep_left(w[2], left - 8), keep_left(w[3], left - 12));
    }
}

// fp32 NCHW out (the logits): one thread per output pixel, pixels fastest (coalesced fp32 stores per channel plane; the
// source image is a few hundred KB of codes and stays in L1 / L2)
__global__ void __launch_bounds__(kThreads)
interpq8_f32_kernel(InterpQ8Args a) {
    const unsigned idx = blockIdx.x * kThreads + threadIdx.x;
    if (idx >= (unsigned)a.N * a.OH * a.OW) return;
    const int ox = (int)(idx % (unsigned)a.OW);
    const int oy = (int)((idx / (unsigned)a.OW) % (unsigned)a.OH);
    const int n = (int)(idx / ((unsigned)a.OW * (unsigned)a.OH));
    int y0, y1, x0, x1; float l1y, l1x;
    q8_interp_index(oy, a.sy, a.H, a.align, y0, y1, l1y);
    q8_interp_index(ox, a.sx, a.W, a.align, x0, x1, l1x);
    const RangeSlot in = *a.in_slot;
    const int8_t *xn = a.x + (int64_t)n * a.H * a.W * a.Cp;
    const int64_t plane = (int64_t)a.OH * a.OW;
    float *yp = a.yf + (int64_t)n * a.C * plane + (int64_t)oy * a.OW + ox;
    for (int cwi = 0; 4 * cwi < a.C; ++cwi) {
        float v[4];
        interp_word(a, xn, cwi, y0, y1, l1y, x0, x1, l1x, in.scale, in.mn, v);
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if (4 * cwi + c < a.C) yp[(int64_t)(4 * cwi + c) * plane] = v[c];
    }
}

struct RegridQ8Args {
    const int8_t *x; const RangeSlot *in_slot;
    int8_t *y[2]; const RangeSlot *out_slot[2];
    int64_t G; int C, cp_in, cp_out, c_off, words_out;    // words_out: 4-channel words this launch writes per pixel
};

__global__ void __launch_bounds__(kThreads)
regridq8_kernel(RegridQ8Args a) {
    const unsigned idx = blockIdx.x * kThreads + threadIdx.x;
    if (idx >= (unsigned)a.G * (unsigned)a.words_out) return;
    const int wi = (int)(idx % (unsigned)a.words_out);
    const int64_t g = idx / (unsigned)a.words_out;
    const int left = a.C - 4 * wi;                               // real channels in this word (<= 0: the padded tail)
    uint32_t w0 = 0u, w1 = 0u;
    if (left > 0) {
        const RangeSlot in = *a.in_slot;
        float v[4];
        dequant4(__ldg(reinterpret_cast<const uint32_t *>(a.x + g * a.cp_in) + wi), in.scale, in.mn, v);
        uint32_t w[1];
        pack_codes<4>(v, consts_from_slot(*a.out_slot[0], 8, 0, 0.f, 0.f), w);
        w0 = keep_left(w[0], left);
        if (a.y[1]) {
            pack_codes<4>(v, consts_from_slot(*a.out_slot[1], 8, 0, 0.f, 0.f), w);
            w1 = keep_left(w[0], left);
        }
    }
    const int64_t off = g * a.cp_out + a.c_off + 4 * wi;
    *reinterpret_cast<uint32_t *>(a.y[0] + off) = w0;
    if (a.y[1]) *reinterpret_cast<uint32_t *>(a.y[1] + off) = w1;
}

// the same on whole 16-channel chunks (operand offset and row pitch multiples of 16, and either C a multiple of 16 or
// this operand the row's last): one 128-bit load and store per thread, the slot constants formed once per 16 values
__global__ void __launch_bounds__(kThreads)
regridq8_v16_kernel(RegridQ8Args a) {
    const unsigned chunks = (unsigned)(a.words_out + 3) >> 2;
    const unsigned idx = blockIdx.x * kThreads + threadIdx.x;
    if (idx >= (unsigned)a.G * chunks) return;
    const int ci = (int)(idx % chunks);
    const int64_t g = idx / chunks;
    const int left = a.C - 16 * ci;
    uint4 o0 = make_uint4(0u, 0u, 0u, 0u), o1 = o0;
    if (left > 0) {
        const RangeSlot in = *a.in_slot;
        const uint4 q = __ldg(reinterpret_cast<const uint4 *>(a.x + g * a.cp_in) + ci);
        const uint32_t qw[4] = {q.x, q.y, q.z, q.w};
        float v[16];
#pragma unroll
        for (int k = 0; k < 4; ++k) dequant4(qw[k], in.scale, in.mn, v + 4 * k);
        uint32_t w[4];
        pack_codes<16>(v, consts_from_slot(*a.out_slot[0], 8, 0, 0.f, 0.f), w);
        o0 = make_uint4(keep_left(w[0], left), keep_left(w[1], left - 4), keep_left(w[2], left - 8), keep_left(w[3], left - 12));
        if (a.y[1]) {
            pack_codes<16>(v, consts_from_slot(*a.out_slot[1], 8, 0, 0.f, 0.f), w);
            o1 = make_uint4(keep_left(w[0], left), keep_left(w[1], left - 4), keep_left(w[2], left - 8), keep_left(w[3], left - 12));
        }
    }
    const int64_t off = g * a.cp_out + a.c_off + 16 * ci;
    *reinterpret_cast<uint4 *>(a.y[0] + off) = o0;
    if (a.y[1]) *reinterpret_cast<uint4 *>(a.y[1] + off) = o1;
}

// mean over the HW rows of x[n] ([HW][C] fp32, rows `ld` floats apart): a block per (n, 32 channels), lanes on channels
// (coalesced 128-byte row segments), warps striding the rows; fp64 accumulation, fixed combination order
__global__ void __launch_bounds__(kThreads)
poolq8_kernel(const float *__restrict__ x, int64_t ld, int N, int HW, int C, int Cp, int8_t *y0, const RangeSlot *s0,
              int8_t *y1, const RangeSlot *s1) {
    __shared__ double part[kWarps][32];
    const int groups = (Cp + 31) >> 5;
    const int n = blockIdx.x / groups, c = (blockIdx.x % groups) * 32 + (threadIdx.x & 31);
    const int warp = threadIdx.x >> 5;
    double acc = 0.0;
    if (c < C) {
        const float *p = x + (int64_t)n * HW * ld + c;
        for (int i = warp; i < HW; i += kWarps) acc += (double)__ldg(p + (int64_t)i * ld);
    }
    part[warp][threadIdx.x & 31] = acc;
    __syncthreads();
    if (warp == 0 && c < Cp) {
        int8_t code0 = 0, code1 = 0;
        if (c < C) {
            double t = 0.0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) t += part[w][threadIdx.x];
            const float m = (float)(t / (double)HW);
            code0 = (int8_t)(pack4(m, m, m, m, consts_from_slot(*s0, 8, 0, 0.f, 0.f)) & 0xffu);
            if (y1) code1 = (int8_t)(pack4(m, m, m, m, consts_from_slot(*s1, 8, 0, 0.f, 0.f)) & 0xffu);
        }
        y0[n * Cp + c] = code0;
        if (y1) y1[n * Cp + c] = code1;
    }
}

// F.pad(x, (l, r, t, b)) AHEAD of a QuantMeasure (DeepLab's fixed_padding): the measure quantises the padded zeros with
// the tensor, so the border holds the code of 0.0 on that grid.  One thread per output pixel and 16-byte channel chunk.
__global__ void __launch_bounds__(kThreads)
padq8_kernel(const int8_t *__restrict__ x, const RangeSlot *slot, int8_t *__restrict__ y, int N, int H, int W, int Cp,
             int C, int pl, int pt, int OH, int OW) {
    const unsigned chunks = (unsigned)Cp >> 4;
    const unsigned idx = blockIdx.x * kThreads + threadIdx.x;
    if (idx >= (unsigned)N * OH * OW * chunks) return;
    const int ch = (int)(idx % chunks);
    unsigned rest = idx / chunks;
    const int ox = (int)(rest % (unsigned)OW); rest /= (unsigned)OW;
    const int oy = (int)(rest % (unsigned)OH);
    const int n = (int)(rest / (unsigned)OH);
    const int iy = oy - pt, ix = ox - pl;
    uint4 v;
    if ((unsigned)iy < (unsigned)H && (unsigned)ix < (unsigned)W) {
        v = __ldg(reinterpret_cast<const uint4 *>(x + (((int64_t)n * H + iy) * W + ix) * Cp) + ch);
    } else {
        const uint32_t w = pack4(0.0f, 0.0f, 0.0f, 0.0f, consts_from_slot(*slot, 8, 0, 0.f, 0.f));
        const int left = C - 16 * ch;                             // the tensor's padding CHANNELS stay s8 zero
        v = make_uint4(keep_left(w, left), keep_left(w, left - 4), keep_left(w, left - 8), keep_left(w, left - 12));
    }
    reinterpret_cast<uint4 *>(y + (((int64_t)n * OH + oy) * OW + ox) * Cp)[ch] = v;
}

// ---- upscaling forms: a thread walks kInterpRows consecutive output rows of one column.  The horizontal blends
// `top` / `bot` of the reference's expression l0y * (l0x * a00 + l1x * a01) + l1y * (l0x * a10 + l1x * a11) depend on the
// input row pair only, so they are formed once per pair (four rows share one when the image grows 4 x, DeepLab's
// 33 -> 129 and 129 -> 513) and an output row costs the vertical blend alone.  Same operations in the same order per
// output value as the one-row kernels: the same bits.
constexpr int kInterpRows = 4;

__global__ void __launch_bounds__(kThreads)
interpq8_codes_rows_kernel(InterpQ8Args a) {
    const unsigned chunks = (unsigned)a.Cp >> 4, groups = ((unsigned)a.OH + kInterpRows - 1) / kInterpRows;
    const unsigned idx = blockIdx.x * kThreads + threadIdx.x;
    if (idx >= (unsigned)a.N * groups * a.OW * chunks) return;
    const int ch = (int)(idx % chunks);
    unsigned rest = idx / chunks;
    const int ox = (int)(rest % (unsigned)a.OW); rest /= (unsigned)a.OW;
    const int g = (int)(rest % groups);
    const int n = (int)(rest / groups);
    int x0, x1; float l1x;
    q8_interp_index(ox, a.sx, a.W, a.align, x0, x1, l1x);
    const float l0x = __fsub_rn(1.0f, l1x);
    const RangeSlot in = *a.in_slot;
    const int8_t *xn = a.x + (int64_t)n * a.H * a.W * a.Cp;
    const int64_t rs = (int64_t)a.W * a.Cp;
    const QuantConsts q0 = consts_from_slot(*a.out_slot[0], 8, 0, 0.f, 0.f);
    const int left = a.C - 16 * ch;
    int cy0 = -1, cy1 = -1;
    float top[16], bot[16];
#pragma unroll 1
    for (int r = 0; r < kInterpRows; ++r) {
        const int oy = g * kInterpRows + r;
        if (oy >= a.OH) break;
        int y0, y1; float l1y;
        q8_interp_index(oy, a.sy, a.H, a.align, y0, y1, l1y);
        if (y0 != cy0 || y1 != cy1) {
            cy0 = y0; cy1 = y1;
            const uint4 q00 = __ldg(reinterpret_cast<const uint4 *>(xn + y0 * rs + (int64_t)x0 * a.Cp) + ch);
            const uint4 q01 = __ldg(reinterpret_cast<const uint4 *>(xn + y0 * rs + (int64_t)x1 * a.Cp) + ch);
            const uint4 q10 = __ldg(reinterpret_cast<const uint4 *>(xn + y1 * rs + (int64_t)x0 * a.Cp) + ch);
            const uint4 q11 = __ldg(reinterpret_cast<const uint4 *>(xn + y1 * rs + (int64_t)x1 * a.Cp) + ch);
            const uint32_t w00[4] = {q00.x, q00.y, q00.z, q00.w}, w01[4] = {q01.x, q01.y, q01.z, q01.w};
            const uint32_t w10[4] = {q10.x, q10.y, q10.z, q10.w}, w11[4] = {q11.x, q11.y, q11.z, q11.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float a00[4], a01[4], a10[4], a11[4];
                dequant4(w00[k], in.scale, in.mn, a00); dequant4(w01[k], in.scale, in.mn, a01);
                dequant4(w10[k], in.scale, in.mn, a10); dequant4(w11[k], in.scale, in.mn, a11);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    top[4 * k + c] = __fadd_rn(__fmul_rn(l0x, a00[c]), __fmul_rn(l1x, a01[c]));
                    bot[4 * k + c] = __fadd_rn(__fmul_rn(l0x, a10[c]), __fmul_rn(l1x, a11[c]));
                }
            }
        }
        const float l0y = __fsub_rn(1.0f, l1y);
        float v[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) v[c] = __fadd_rn(__fmul_rn(l0y, top[c]), __fmul_rn(l1y, bot[c]));
        const int64_t off = (((int64_t)n * a.OH + oy) * a.OW + ox) * a.Cp + 16 * ch;
        uint32_t w[4];
        pack_codes<16>(v, q0, w);
        *reinterpret_cast<uint4 *>(a.y[0] + off) = make_uint4(keep_left(w[0], left), keep_left(w[1], left - 4),
                                                              keep_left(w[2], left - 8), keep_left(w[3], left - 12));
        if (a.y[1]) {
            pack_codes<16>(v, consts_from_slot(*a.out_slot[1], 8, 0, 0.f, 0.f), w);
            *reinterpret_cast<uint4 *>(a.y[1] + off) = make_uint4(keep_left(w[0], left), keep_left(w[1], left - 4),
                                                                  keep_left(w[2], left - 8), keep_left(w[3], left - 12));
        }
    }
}

// fp32 NCHW out, C <= 32: all channel words of the column's row pair blended once, then one vertical blend and one
// store per channel and output row
__global__ void __launch_bounds__(kThreads)
interpq8_f32_rows_kernel(InterpQ8Args a) {
    const unsigned groups = ((unsigned)a.OH + kInterpRows - 1) / kInterpRows;
    const unsigned idx = blockIdx.x * kThreads + threadIdx.x;
    if (idx >= (unsigned)a.N * groups * a.OW) return;
    const int ox = (int)(idx % (unsigned)a.OW);
    const int g = (int)((idx / (unsigned)a.OW) % groups);
    const int n = (int)(idx / ((unsigned)a.OW * groups));
    int x0, x1; float l1x;
    q8_interp_index(ox, a.sx, a.W, a.align, x0, x1, l1x);
    const float l0x = __fsub_rn(1.0f, l1x);
    const RangeSlot in = *a.in_slot;
    const int8_t *xn = a.x + (int64_t)n * a.H * a.W * a.Cp;
    const int64_t rs = (int64_t)a.W * a.Cp, plane = (int64_t)a.OH * a.OW;
    int cy0 = -1, cy1 = -1;
    float top[32], bot[32];
#pragma unroll 1
    for (int r = 0; r < kInterpRows; ++r) {
        const int oy = g * kInterpRows + r;
        if (oy >= a.OH) break;
        int y0, y1; float l1y;
        q8_interp_index(oy, a.sy, a.H, a.align, y0, y1, l1y);
        if (y0 != cy0 || y1 != cy1) {
            cy0 = y0; cy1 = y1;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                if (16 * h >= a.C) break;
                const uint4 q00 = __ldg(reinterpret_cast<const uint4 *>(xn + y0 * rs + (int64_t)x0 * a.Cp) + h);
                const uint4 q01 = __ldg(reinterpret_cast<const uint4 *>(xn + y0 * rs + (int64_t)x1 * a.Cp) + h);
                const uint4 q10 = __ldg(reinterpret_cast<const uint4 *>(xn + y1 * rs + (int64_t)x0 * a.Cp) + h);
                const uint4 q11 = __ldg(reinterpret_cast<const uint4 *>(xn + y1 * rs + (int64_t)x1 * a.Cp) + h);
                const uint32_t w00[4] = {q00.x, q00.y, q00.z, q00.w}, w01[4] = {q01.x, q01.y, q01.z, q01.w};
                const uint32_t w10[4] = {q10.x, q10.y, q10.z, q10.w}, w11[4] = {q11.x, q11.y, q11.z, q11.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    float a00[4], a01[4], a10[4], a11[4];
                    dequant4(w00[k], in.scale, in.mn, a00); dequant4(w01[k], in.scale, in.mn, a01);
                    dequant4(w10[k], in.scale, in.mn, a10); dequant4(w11[k], in.scale, in.mn, a11);
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        top[16 * h + 4 * k + c] = __fadd_rn(__fmul_rn(l0x, a00[c]), __fmul_rn(l1x, a01[c]));
                        bot[16 * h + 4 * k + c] = __fadd_rn(__fmul_rn(l0x, a10[c]), __fmul_rn(l1x, a11[c]));
                    }
                }
            }
        }
        const float l0y = __fsub_rn(1.0f, l1y);
        float *yp = a.yf + (int64_t)n * a.C * plane + (int64_t)oy * a.OW + ox;
#pragma unroll
        for (int c = 0; c < 32; ++c)
            if (c < a.C) yp[(int64_t)c * plane] = __fadd_rn(__fmul_rn(l0y, top[c]), __fmul_rn(l1y, bot[c]));
    }
}

inline float q8_interp_scale(int in, int out, int align) {
    if (align) return out > 1 ? (float)(in - 1) / (float)(out - 1) : 0.0f;
    return out > 0 ? (float)in / (float)out : 0.0f;
}

}  // namespace
}  // namespace dfq

using namespace dfq;

extern "C" int dfq_interp_q8(const int8_t *x, const float *in_slot, int N, int C, int H, int W, int OH, int OW,
                             int align_corners, int8_t *y0, const float *out_slot0, int8_t *y1, const float *out_slot1,
                             float *yf, void *stream) {
    if (!x || !in_slot || N <= 0 || C <= 0 || H <= 0 || W <= 0 || OH <= 0 || OW <= 0 || (!y0 && !yf) || (y0 && !out_slot0) ||
        (y1 && (!y0 || !out_slot1)))
        return (int)cudaErrorInvalidValue;
    InterpQ8Args a = {};
    a.x = x; a.in_slot = (const RangeSlot *)in_slot;
    a.y[0] = y0; a.y[1] = y1; a.out_slot[0] = (const RangeSlot *)out_slot0; a.out_slot[1] = (const RangeSlot *)out_slot1;
    a.yf = yf;
    a.N = N; a.C = C; a.Cp = (C + 15) / 16 * 16; a.H = H; a.W = W; a.OH = OH; a.OW = OW; a.align = align_corners ? 1 : 0;
    a.sy = q8_interp_scale(H, OH, a.align); a.sx = q8_interp_scale(W, OW, a.align);
    const cudaStream_t st = (cudaStream_t)stream;
    // growing at least 2 x vertically: the row-walking forms (DFQ_Q8_INTERP_ROWS=0: the one-row kernels, for A/B runs)
    const char *env = getenv("DFQ_Q8_INTERP_ROWS");
    const bool rows = OH >= 2 * H && !(env && env[0] == '0');
    const int64_t groups = (OH + kInterpRows - 1) / kInterpRows;
    if (y0) {
        const int64_t blocks = ((int64_t)N * (rows ? groups : OH) * OW * (a.Cp / 16) + kThreads - 1) / kThreads;
        if (((int64_t)N * OH * OW * (a.Cp / 16) + kThreads - 1) / kThreads >= (1ll << 23))
            return (int)cudaErrorInvalidValue;                             // 32-bit thread indices in the kernels
        if (rows) interpq8_codes_rows_kernel<<<(unsigned)blocks, kThreads, 0, st>>>(a);
        else interpq8_codes_kernel<<<(unsigned)blocks, kThreads, 0, st>>>(a);
        g_launches += 1;
    }
    if (yf) {
        const bool rows_f = rows && a.Cp <= 32;
        const int64_t blocks = ((int64_t)N * (rows_f ? groups : OH) * OW + kThreads - 1) / kThreads;
        if (((int64_t)N * OH * OW + kThreads - 1) / kThreads >= (1ll << 23))
            return (int)cudaErrorInvalidValue;                             // 32-bit thread indices in the kernels
        if (rows_f) interpq8_f32_rows_kernel<<<(unsigned)blocks, kThreads, 0, st>>>(a);
        else interpq8_f32_kernel<<<(unsigned)blocks, kThreads, 0, st>>>(a);
        g_launches += 1;
    }
    return (int)cudaGetLastError();
}

extern "C" int dfq_regrid_q8(const int8_t *x, const float *in_slot, int64_t G, int C, int8_t *y0, const float *out_slot0,
                             int8_t *y1, const float *out_slot1, int cp_out, int c_off, int fill_tail, void *stream) {
    if (!x || !in_slot || !y0 || !out_slot0 || (y1 && !out_slot1) || G <= 0 || C <= 0 || (c_off & 3) || cp_out <= 0 ||
        (cp_out & 15) || c_off + C > cp_out)
        return (int)cudaErrorInvalidValue;
    RegridQ8Args a = {};
    a.x = x; a.in_slot = (const RangeSlot *)in_slot;
    a.y[0] = y0; a.y[1] = y1; a.out_slot[0] = (const RangeSlot *)out_slot0; a.out_slot[1] = (const RangeSlot *)out_slot1;
    a.G = G; a.C = C; a.cp_in = (C + 15) / 16 * 16; a.cp_out = cp_out; a.c_off = c_off;
    // the last operand also clears the row's padding channels (buffers are recycled between tensors)
    a.words_out = fill_tail ? (cp_out - c_off) / 4 : (C + 3) / 4;
    const int64_t blocks = (G * a.words_out + kThreads - 1) / kThreads;
    if (blocks >= (1ll << 23)) return (int)cudaErrorInvalidValue;      // 32-bit thread indices in the kernels
    const char *env = getenv("DFQ_Q8_REGRID_V16");                     // (=0: the word kernel, for A/B runs)
    const bool v16 = (c_off % 16) == 0 && (cp_out % 16) == 0 && ((C % 16) == 0 || fill_tail) &&
                     ((uintptr_t)y0 % 16) == 0 && (!y1 || ((uintptr_t)y1 % 16) == 0) && ((uintptr_t)x % 16) == 0 &&
                     !(env && env[0] == '0');
    if (v16) {
        const int64_t chunks = (a.words_out + 3) / 4;
        regridq8_v16_kernel<<<(unsigned)((G * chunks + kThreads - 1) / kThreads), kThreads, 0, (cudaStream_t)stream>>>(a);
    } else
    regridq8_kernel<<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(a);
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_pool_q8(const float *x, int64_t ld, int N, int HW, int C, int8_t *y0, const float *out_slot0,
                           int8_t *y1, const float *out_slot1, void *stream) {
    if (!x || !y0 || !out_slot0 || (y1 && !out_slot1) || N <= 0 || HW <= 0 || C <= 0 || ld < C)
        return (int)cudaErrorInvalidValue;
    const int Cp = (C + 15) / 16 * 16;
    poolq8_kernel<<<(unsigned)(N * ((Cp + 31) / 32)), kThreads, 0, (cudaStream_t)stream>>>(
        x, ld, N, HW, C, Cp, y0, (const RangeSlot *)out_slot0, y1, (const RangeSlot *)out_slot1);
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_pad_q8(const int8_t *x, const float *slot, int N, int C, int H, int W, int pl, int pr, int pt, int pb,
                          int8_t *y, void *stream) {
    if (!x || !slot || !y || N <= 0 || C <= 0 || H <= 0 || W <= 0 || pl < 0 || pr < 0 || pt < 0 || pb < 0)
        return (int)cudaErrorInvalidValue;
    const int Cp = (C + 15) / 16 * 16, OH = H + pt + pb, OW = W + pl + pr;
    const int64_t blocks = ((int64_t)N * OH * OW * (Cp / 16) + kThreads - 1) / kThreads;
    if (blocks >= (1ll << 23)) return (int)cudaErrorInvalidValue;      // 32-bit thread indices in the kernels
    padq8_kernel<<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(x, (const RangeSlot *)slot, y, N, H, W, Cp, C, pl,
                                                                         pt, OH, OW);
    g_launches += 1;
    return (int)cudaGetLastError();
}
