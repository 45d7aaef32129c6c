// Elementwise fake-quant pass over [begin, end) of one tensor, shared by quant.cu and act.cu.
#pragma once
#include "common.cuh"

namespace dfq {

template <typename CodeT>
__device__ __forceinline__ void apply_range(const float *x, float *y, CodeT *codes, int64_t begin, int64_t end,
                                            const QuantConsts &c, bool clamp_out, int tid, int nthr) {
    const float *p = x + begin;
    float *o = y + begin;
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
            float q0, q1, q2, q3;
            float4 r;
            r.x = fake_quant_elem(v[u].x, c, q0);
            r.y = fake_quant_elem(v[u].y, c, q1);
            r.z = fake_quant_elem(v[u].z, c, q2);
            r.w = fake_quant_elem(v[u].w, c, q3);
            if (clamp_out) {
                r.x = fminf(fmaxf(r.x, c.clip_lo), c.clip_hi); r.y = fminf(fmaxf(r.y, c.clip_lo), c.clip_hi);
                r.z = fminf(fmaxf(r.z, c.clip_lo), c.clip_hi); r.w = fminf(fmaxf(r.w, c.clip_lo), c.clip_hi);
            }
            o4[k] = r;
            if (codes) {
                CodeT *cp = codes + begin + 4 * k;
                unsigned c0 = (unsigned)(int)(q0 - c.qmin), c1 = (unsigned)(int)(q1 - c.qmin);
                unsigned c2 = (unsigned)(int)(q2 - c.qmin), c3 = (unsigned)(int)(q3 - c.qmin);
                if ((((uintptr_t)cp) & (4 * sizeof(CodeT) - 1)) == 0) {   // one packed store per 4 codes
                    if (sizeof(CodeT) == 1)
                        *reinterpret_cast<unsigned *>(cp) = c0 | (c1 << 8) | (c2 << 16) | (c3 << 24);
                    else
                        *reinterpret_cast<uint2 *>(cp) = make_uint2(c0 | (c1 << 16), c2 | (c3 << 16));
                } else {
                    cp[0] = (CodeT)c0; cp[1] = (CodeT)c1; cp[2] = (CodeT)c2; cp[3] = (CodeT)c3;
                }
            }
        }
    }
    for (int64_t k = (n4 << 2) + tid; k < n; k += nthr) {
        float q;
        float r = fake_quant_elem(p[k], c, q);
        if (clamp_out) r = fminf(fmaxf(r, c.clip_lo), c.clip_hi);
        o[k] = r;
        if (codes) codes[begin + k] = (CodeT)(int)(q - c.qmin);
    }
}

}  // namespace dfq
