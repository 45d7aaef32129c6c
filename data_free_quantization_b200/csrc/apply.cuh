// Elementwise fake-quant pass over [begin, end) of one tensor, shared by quant.cu and act.cu.
#pragma once
#include "common.cuh"

namespace dfq {

// one float4: quantise, store, optional packed code store (cp = address of the first of the four codes, or null)
__device__ __forceinline__ void apply_vec4(const float4 &v, float4 *o, uint8_t *cp, int cb, const QuantConsts &c,
                                           bool clamp_out) {
    float q0, q1, q2, q3;
    float4 r;
    r.x = fake_quant_elem(v.x, c, q0);
    r.y = fake_quant_elem(v.y, c, q1);
    r.z = fake_quant_elem(v.z, c, q2);
    r.w = fake_quant_elem(v.w, c, q3);
    if (clamp_out) {
        r.x = fminf(fmaxf(r.x, c.clip_lo), c.clip_hi); r.y = fminf(fmaxf(r.y, c.clip_lo), c.clip_hi);
        r.z = fminf(fmaxf(r.z, c.clip_lo), c.clip_hi); r.w = fminf(fmaxf(r.w, c.clip_lo), c.clip_hi);
    }
    *o = r;
    if (cp) {
        unsigned c0 = (unsigned)(int)(q0 - c.qmin), c1 = (unsigned)(int)(q1 - c.qmin);
        unsigned c2 = (unsigned)(int)(q2 - c.qmin), c3 = (unsigned)(int)(q3 - c.qmin);
        if ((((uintptr_t)cp) & (4 * cb - 1)) == 0) {   // one packed store per 4 codes
            if (cb == 1)
                *reinterpret_cast<unsigned *>(cp) = c0 | (c1 << 8) | (c2 << 16) | (c3 << 24);
            else
                *reinterpret_cast<uint2 *>(cp) = make_uint2(c0 | (c1 << 16), c2 | (c3 << 16));
        } else if (cb == 1) {
            cp[0] = (uint8_t)c0; cp[1] = (uint8_t)c1; cp[2] = (uint8_t)c2; cp[3] = (uint8_t)c3;
        } else {
            uint16_t *cw = reinterpret_cast<uint16_t *>(cp);
            cw[0] = (uint16_t)c0; cw[1] = (uint16_t)c1; cw[2] = (uint16_t)c2; cw[3] = (uint16_t)c3;
        }
    }
}

// kFixedBytes = 1 / 2: code width known at compile time; 0: run-time `code_bytes` (multi-tensor launches, where one
// instantiation of the loop instead of one per width and per work mode keeps the kernel at 5 blocks per SM).
template <int kFixedBytes>
__device__ __forceinline__ void apply_range_w(const float *x, float *y, void *codes_v, int code_bytes, int64_t begin,
                                              int64_t end, const QuantConsts &c, bool clamp_out, int tid, int nthr) {
    const int cb = kFixedBytes ? kFixedBytes : code_bytes;
    // global state space (LDG / STG instead of generic LD / ST): every tensor the library touches is device memory
    uint8_t *codes = codes_v ? reinterpret_cast<uint8_t *>(__cvta_generic_to_global(codes_v)) : nullptr;
    const float *p = reinterpret_cast<const float *>(__cvta_generic_to_global(x)) + begin;
    float *o = reinterpret_cast<float *>(__cvta_generic_to_global(y)) + begin;
    int64_t n = end - begin;
    bool vec = ((((uintptr_t)p) | ((uintptr_t)o)) & 15) == 0;
    int64_t n4 = vec ? (n >> 2) : 0;
    const float4 *p4 = reinterpret_cast<const float4 *>(p);
    float4 *o4 = reinterpret_cast<float4 *>(o);
    for (int64_t base = 0; base < n4; base += 4 * (int64_t)nthr) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            int64_t k = base + tid + (int64_t)u * nthr;
            if (k < n4) v[u] = p4[k];        // plain load: y may alias x
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            int64_t k = base + tid + (int64_t)u * nthr;
            if (k >= n4) continue;
            apply_vec4(v[u], o4 + k, codes ? codes + (begin + 4 * k) * cb : nullptr, cb, c, clamp_out);
        }
    }
    for (int64_t k = (n4 << 2) + tid; k < n; k += nthr) {
        float q;
        float r = fake_quant_elem(p[k], c, q);
        if (clamp_out) r = fminf(fmaxf(r, c.clip_lo), c.clip_hi);
        o[k] = r;
        if (codes) {
            if (cb == 1) codes[begin + k] = (uint8_t)(int)(q - c.qmin);
            else reinterpret_cast<uint16_t *>(codes)[begin + k] = (uint16_t)(int)(q - c.qmin);
        }
    }
}

template <typename CodeT>
__device__ __forceinline__ void apply_range(const float *x, float *y, CodeT *codes, int64_t begin, int64_t end,
                                            const QuantConsts &c, bool clamp_out, int tid, int nthr) {
    apply_range_w<(int)sizeof(CodeT)>(x, y, codes, (int)sizeof(CodeT), begin, end, c, clamp_out, tid, nthr);
}

}  // namespace dfq
