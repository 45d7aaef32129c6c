// Code packing helpers of the static-range (u8 activation) kernels, shared by q8.cu and pw_i8.cu: the exact quantiser
// of utils/quantize.py:72-74 evaluated on groups of values, centred s8 code words, and their inverse.
#pragma once
#include "common.cuh"

namespace dfq {
namespace {

// The integer code of utils/quantize.py:72-74 in the low byte of the result (see quant_code, common.cuh)
__device__ __forceinline__ uint32_t code_bits(float x, const QuantConsts &c) {
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
// The exact quantiser (IEEE division) for a value whose reciprocal estimate sits near a tie: rare, kept OUT of line so
// that the unrolled fast paths stay small (the depthwise kernel was 94 KB of code and stalled on instruction fetch).
__device__ __noinline__ uint32_t code_bits_exact(float x, float neg_min, float scale, float rscale, float qmin,
                                                 float qmax, float t_lo, float t_hi) {
    const float t = __fadd_rn(x, neg_min);
    const float tc = fminf(fmaxf(t, t_lo), t_hi);
    return __float_as_uint(__fadd_rn(fminf(fmaxf(div_const_rn(tc, scale, rscale), qmin), qmax), 12582912.0f));
}

// N values -> N centred s8 codes, four per word.  The reciprocal estimates of all N values are formed in straight-line
// code (the values interleave in the pipeline); ONE branch per call goes to the exact division, taken when any of
// them sits within a few ulp of a tie (~5e-4 per value).
template <int N>
__device__ __forceinline__ void pack_codes(const float (&v)[N], const QuantConsts &c, uint32_t (&w)[N / 4]) {
    uint32_t b[N];
    if (c.fast) {                                           // block-uniform
        const float magic = 12582912.0f;
        uint32_t flagged = 0;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            const float cl = fminf(fmaxf(__fmul_rn(__fadd_rn(v[i], c.neg_min), c.rscale), c.qmin), c.qmax);
            const float m = __fadd_rn(cl, magic);
            b[i] = __float_as_uint(m);
            if (fabsf(__fsub_rn(cl, __fadd_rn(m, -magic))) > c.amb) flagged |= 1u << i;
        }
        while (flagged) {                                    // rare: out-of-line exact division, one value at a time
            const int i = __ffs(flagged) - 1;
            flagged &= flagged - 1;
            float x = v[0];
#pragma unroll
            for (int j = 1; j < N; ++j) x = (j == i) ? v[j] : x;
            const uint32_t e = code_bits_exact(x, c.neg_min, c.scale, c.rscale, c.qmin, c.qmax, c.t_lo, c.t_hi);
#pragma unroll
            for (int j = 0; j < N; ++j) b[j] = (j == i) ? e : b[j];
        }
    } else {
#pragma unroll
        for (int i = 0; i < N; ++i) b[i] = code_bits(v[i], c);
    }
#pragma unroll
    for (int j = 0; j < N / 4; ++j) {
        const uint32_t lo = __byte_perm(b[4 * j], b[4 * j + 1], 0x0040), hi = __byte_perm(b[4 * j + 2], b[4 * j + 3], 0x0040);
        w[j] = __byte_perm(lo, hi, 0x5410) ^ 0x80808080u;
    }
}
// four values -> four centred s8 codes in one word
__device__ __forceinline__ uint32_t pack4(float v0, float v1, float v2, float v3, const QuantConsts &c) {
    const float v[4] = {v0, v1, v2, v3};
    uint32_t w[1];
    pack_codes<4>(v, c, w);
    return w[0];
}
// bytes of channels >= `left` (counted from the word's first channel) cleared: padding channels are s8 zero
__device__ __forceinline__ uint32_t keep_left(uint32_t w, int left) {
    return left >= 4 ? w : (left <= 0 ? 0u : (w & (0xffffffffu >> (8 * (4 - left)))));
}
// centred s8 code word -> the four fp32 values the reference's quantiser produced: fl(fl(code * s) + min)
__device__ __forceinline__ void dequant4(uint32_t w, float s, float mn, float *out) {
    const uint32_t u = w ^ 0x80808080u;                                     // codes 0..255
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float q = __fadd_rn(__uint_as_float(__byte_perm(u, 0x4B000000u, 0x7650 + i)), -8388608.0f);
        out[i] = __fadd_rn(__fmul_rn(q, s), mn);
    }
}

// device form of DfqQ8OutMap: Wp == 0 -> row g; else the interior of a padded tensor
struct Q8OutMap { int OW, OHW, pl, pt, Wp, HpWp; };
__device__ __forceinline__ int64_t q8_out_pixel(const Q8OutMap &m, int64_t g) {
    if (m.Wp == 0) return g;
    const uint32_t g32 = (uint32_t)g;                           // pixel counts are < 2^31 (checked by the launchers)
    const uint32_t n = g32 / (uint32_t)m.OHW;
    const uint32_t r = g32 - n * (uint32_t)m.OHW;
    const uint32_t oh = r / (uint32_t)m.OW, ow = r - oh * (uint32_t)m.OW;
    return (int64_t)n * m.HpWp + (int64_t)((oh + m.pt) * (uint32_t)m.Wp + ow + m.pl);
}
inline Q8OutMap q8_make_map(const DfqQ8OutMap *h) {
    Q8OutMap m = {0, 0, 0, 0, 0, 0};
    if (h && (h->pl | h->pr | h->pt | h->pb)) {
        m.OW = h->OW; m.OHW = h->OH * h->OW; m.pl = h->pl; m.pt = h->pt;
        m.Wp = h->OW + h->pl + h->pr; m.HpWp = (h->OH + h->pt + h->pb) * m.Wp;
    }
    return m;
}

}  // namespace
}  // namespace dfq
