// Shared device helpers of libdfq_b200 (sm_100a only).
// Bit-exactness rules: every fp32 operation of the reference's elementwise
// chains is issued as an explicitly rounded intrinsic (__fadd_rn/__fmul_rn/...), which nvcc never
// contracts into an FMA; the build also passes -fmad=false.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include "../../include/dfq_b200.h"

namespace dfq {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kChunk = 4096;            // fp32 elements per block of the streaming kernels

extern long long g_launches;            // host-side launch counter (api.cu)
int sm_count();                         // SMs of the current device, cached (api.cu)
// grid of a grid-stride pass: exactly the blocks that are resident at once (occupancy of THAT kernel x SMs), never
// more than there are chunks.  A fixed "8 per SM" over-subscribed kernels that fit 5 or 6: the extra blocks only
// started when a first-wave block had finished its whole share, and ran the last quarter of the data at 1/3 occupancy.
int resident_blocks_per_sm(const void *kernel, int threads, size_t smem);   // cached per kernel (api.cu)
template <typename K>
inline unsigned reduce_grid(K kernel, int64_t virtual_blocks, int threads = kThreads, size_t smem = 0) {
    int64_t g = (int64_t)sm_count() * resident_blocks_per_sm((const void *)kernel, threads, smem);
    return (unsigned)(virtual_blocks < g ? virtual_blocks : g);
}

// torch.min / torch.max propagate NaN (utils/quantize.py:29-37, float(W.min())); fminf / fmaxf drop it.
__device__ __forceinline__ float fmin_nan(float a, float b) {
    float r;
    asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float fmax_nan(float a, float b) {
    float r;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}

// Monotone float -> uint map (larger float => larger code); 0 is below every real value.  NaN (either sign) maps to
// the largest code, so it wins every atomicMax and decodes to NaN again: a range that saw a NaN is NaN.
__device__ __forceinline__ unsigned enc_ordered(float f) {
    unsigned u = __float_as_uint(f);
    if (f != f) return 0xffffffffu;
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float dec_ordered(unsigned e) {
    unsigned u = (e & 0x80000000u) ? (e & 0x7fffffffu) : ~e;
    return __uint_as_float(u);
}

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v = fmin_nan(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v = fmax_nan(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block-wide min/max with one shared-memory round; result valid in every thread.
__device__ __forceinline__ void block_minmax(float &lo, float &hi, float *smem /* 2*kWarps */) {
    lo = warp_min(lo);
    hi = warp_max(hi);
    int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) { smem[w] = lo; smem[kWarps + w] = hi; }
    __syncthreads();
    lo = smem[0];
    hi = smem[kWarps];
#pragma unroll
    for (int i = 1; i < kWarps; ++i) { lo = fmin_nan(lo, smem[i]); hi = fmax_nan(hi, smem[kWarps + i]); }
}

// min/max of [begin,end) of x; vectorised when the address allows it.
__device__ __forceinline__ float4 ld_coherent4(const float4 *p) {      // never the non-coherent (LDG.NC) path
    float4 v;
    asm volatile("ld.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

__device__ __forceinline__ float ld_coherent1(const float *p) {
    float v;
    asm volatile("ld.global.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}

// kNc = false: plain loads, for kernels that write the same memory later in the same launch (fused range + apply).
template <bool kNc = true>
__device__ __forceinline__ void chunk_minmax(const float *__restrict__ x, int64_t begin, int64_t end,
                                             float &lo, float &hi) {
    const float *p = x + begin;
    int64_t n = end - begin;
    if ((((uintptr_t)p) & 15) == 0) {
        const float4 *p4 = reinterpret_cast<const float4 *>(p);
        int64_t n4 = n >> 2;
        float4 v[4];
        int64_t i = threadIdx.x;
        // kChunk/4 = 1024 float4 per block = 4 per thread: issue all loads, then reduce
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            int64_t k = i + (int64_t)u * kThreads;
            v[u] = (k < n4) ? (kNc ? __ldg(p4 + k) : ld_coherent4(p4 + k)) : make_float4(INFINITY, INFINITY, INFINITY, INFINITY);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            int64_t k = i + (int64_t)u * kThreads;
            if (k < n4) {
                lo = fmin_nan(lo, fmin_nan(fmin_nan(v[u].x, v[u].y), fmin_nan(v[u].z, v[u].w)));
                hi = fmax_nan(hi, fmax_nan(fmax_nan(v[u].x, v[u].y), fmax_nan(v[u].z, v[u].w)));
            }
        }
        for (int64_t k = (n4 << 2) + threadIdx.x; k < n; k += kThreads) {
            float t = kNc ? p[k] : ld_coherent1(p + k); lo = fmin_nan(lo, t); hi = fmax_nan(hi, t);
        }
    } else {
        for (int64_t k = threadIdx.x; k < n; k += kThreads) {
            float t = kNc ? p[k] : ld_coherent1(p + k); lo = fmin_nan(lo, t); hi = fmax_nan(hi, t);
        }
    }
}

// The five constants of utils/quantize.py:51-68, computed exactly as the host does
// (double arithmetic for the Python-float path, fp32 for the min/max=None path).
__host__ __device__ inline DfqQuantParams make_quant_params(double mn, double mx, int bits, int flags) {
    double qlo, qhi, sc;
    if (flags & DFQ_Q_SYMMETRIC) {
        qhi = (double)(1ll << (bits - 1)) - 1.0;
        qlo = -(double)(1ll << (bits - 1));
        double a = fabs(mx), b = fabs(mn);
        if (a < b) a = b;
        sc = (flags & DFQ_Q_RANGE_F32) ? (double)((float)a / (float)qhi) : a / qhi;
        mn = 0.0;
    } else {
        qlo = 0.0;
        qhi = (double)(1ll << bits) - 1.0;
        sc = (flags & DFQ_Q_RANGE_F32) ? (double)(((float)mx - (float)mn) / (float)(qhi - qlo))
                                       : (mx - mn) / (qhi - qlo);
    }
    if (1e-8 > sc) sc = 1e-8;              // Python max(scale, 1e-8): a NaN scale stays NaN (utils/quantize.py:68)
    DfqQuantParams p;
    p.neg_min = (float)(-mn);
    p.scale = (float)sc;
    p.qmin = (float)qlo;
    p.qmax = (float)qhi;
    p.add_back = (float)mn;
    return p;
}

// Correctly rounded a / b for a constant divisor b > 0 with rb = RN(1/b) (Markstein): two FMA refinements make
// q faithful, the third makes it the IEEE-rounded quotient.  Valid while a, q*b and the residuals stay clear of
// overflow; underflowing residuals only occur for |a/b| << 0.5 where the rounded code is 0 either way.  The
// caller guarantees both by clamping the numerator first (QuantConsts.t_lo / t_hi).
__device__ __forceinline__ float div_const_rn(float a, float b, float rb) {
    float q = __fmul_rn(a, rb);
    float r = __fmaf_rn(-q, b, a);
    q = __fmaf_rn(r, rb, q);
    r = __fmaf_rn(-q, b, a);
    return __fmaf_rn(r, rb, q);
}

// round-half-even for |c| < 2^22 without the quarter-rate F2F instruction
__device__ __forceinline__ float rint_small(float c) {
    const float magic = 12582912.0f;    // 1.5 * 2^23
    return __fadd_rn(__fadd_rn(c, magic), -magic);
}

struct QuantConsts {
    float neg_min, scale, rscale, qmin, qmax, add_back, clip_lo, clip_hi;
    float t_lo, t_hi;                   // numerator clamp: quotients beyond [qmin-2, qmax+2] clamp to the ends anyway
    float amb;                          // |quotient - code| above this: too close to a tie for the reciprocal estimate
    int fast;                           // 0: exotic scale or > 22-bit grid -> __fdiv_rn / rintf path
};

__device__ __forceinline__ QuantConsts make_consts(const DfqQuantParams &p, float clip_lo, float clip_hi) {
    QuantConsts c;
    c.neg_min = p.neg_min; c.scale = p.scale; c.rscale = __frcp_rn(p.scale);
    c.qmin = p.qmin; c.qmax = p.qmax; c.add_back = p.add_back;
    c.clip_lo = clip_lo; c.clip_hi = clip_hi;
    c.t_lo = __fmul_rn(p.qmin - 2.0f, p.scale);
    c.t_hi = __fmul_rn(p.qmax + 2.0f, p.scale);
    c.amb = 0.5f - (2e-5f + 1e-6f * fmaxf(fabsf(p.qmax), fabsf(p.qmin)));
    c.fast = (p.qmax < 4194304.0f && p.qmin > -4194304.0f && p.scale >= 1e-8f && p.scale < 1e25f);
    return c;
}

// One slot of the device-side range table written by the range pass: the five constants follow from it in fp32.
struct RangeSlot { float mn, mx, scale, pad; };

__device__ __forceinline__ QuantConsts consts_from_slot(const RangeSlot &r, int bits, int flags, float clip_lo,
                                                        float clip_hi) {
    DfqQuantParams p;
    const bool sym = flags & DFQ_Q_SYMMETRIC;
    p.scale = r.scale;
    p.qmin = sym ? -(float)(1 << (bits - 1)) : 0.0f;
    p.qmax = sym ? (float)((1 << (bits - 1)) - 1) : (float)((1ll << bits) - 1);
    p.neg_min = sym ? -0.0f : -r.mn;
    p.add_back = sym ? 0.0f : r.mn;
    return make_consts(p, clip_lo, clip_hi);
}

// utils/quantize.py:72-76, one element.  Returns the dequantised value; q receives the grid value.
__device__ __forceinline__ float fake_quant_elem(float x, const QuantConsts &c, float &q) {
    float t = __fadd_rn(x, c.neg_min);
    float cl;
    if (c.fast) {                                           // block-uniform
        float tc = fminf(fmaxf(t, c.t_lo), c.t_hi);
        float d = div_const_rn(tc, c.scale, c.rscale);
        cl = fminf(fmaxf(d, c.qmin), c.qmax);
        q = rint_small(cl);
    } else {
        float d = __fdiv_rn(t, c.scale);
        cl = fminf(fmaxf(d, c.qmin), c.qmax);
        q = rintf(cl);
    }
    float y = __fadd_rn(__fmul_rn(q, c.scale), c.add_back);
    if (x != x) { q = x; y = x; }                           // torch.clamp propagates NaN
    return y;
}

// The integer code of utils/quantize.py:72-74 as an integer-valued float, for the eval kernels that quantise on load
// (instruction-bound).  t * RN(1/s) is within 2 ulp of the correctly rounded quotient, so both round to the same
// integer unless the estimate sits within a few ulp of a tie; only then (probability ~3e-4 per element) the exact
// Markstein division decides.  Same results as fake_quant_elem for every finite input, 9 instead of 15 operations.
__device__ __forceinline__ float quant_code(float x, const QuantConsts &c) {
    const float t = __fadd_rn(x, c.neg_min);
    if (c.fast) {                                           // block-uniform
        const float cl = fminf(fmaxf(__fmul_rn(t, c.rscale), c.qmin), c.qmax);
        float q = rint_small(cl);
        if (fabsf(__fsub_rn(cl, q)) > c.amb) {
            const float tc = fminf(fmaxf(t, c.t_lo), c.t_hi);
            q = rint_small(fminf(fmaxf(div_const_rn(tc, c.scale, c.rscale), c.qmin), c.qmax));
        }
        return q;
    }
    return rintf(fminf(fmaxf(__fdiv_rn(t, c.scale), c.qmin), c.qmax));
}
__device__ __forceinline__ float quant_dequant(float x, const QuantConsts &c) {     // :72-76, finite inputs
    return __fadd_rn(__fmul_rn(quant_code(x, c), c.scale), c.add_back);
}
// Four elements at once for the fast grids (c.fast checked by the caller, once per kernel): the four reciprocal
// estimates share ONE branch to the exact path, taken when any of them sits near a tie (~1e-3 per call).
__device__ __forceinline__ float4 quant_dequant4_fast(float4 v, const QuantConsts &c) {
    const float x[4] = {v.x, v.y, v.z, v.w};
    float t[4], cl[4], q[4];
    bool near_tie = false;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        t[i] = __fadd_rn(x[i], c.neg_min);
        cl[i] = fminf(fmaxf(__fmul_rn(t[i], c.rscale), c.qmin), c.qmax);
        q[i] = rint_small(cl[i]);
        near_tie |= fabsf(__fsub_rn(cl[i], q[i])) > c.amb;
    }
    if (near_tie) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (fabsf(__fsub_rn(cl[i], q[i])) > c.amb) {
                const float tc = fminf(fmaxf(t[i], c.t_lo), c.t_hi);
                q[i] = rint_small(fminf(fmaxf(div_const_rn(tc, c.scale, c.rscale), c.qmin), c.qmax));
            }
        }
    }
    return make_float4(__fadd_rn(__fmul_rn(q[0], c.scale), c.add_back), __fadd_rn(__fmul_rn(q[1], c.scale), c.add_back),
                       __fadd_rn(__fmul_rn(q[2], c.scale), c.add_back), __fadd_rn(__fmul_rn(q[3], c.scale), c.add_back));
}

__device__ __forceinline__ float quant_or_pass(float v, const QuantConsts &c, bool on) {
    return on ? quant_dequant(v, c) : v;
}

// block-level min/max of `lo/hi` into stats[n] (encoded max(-y), max(y))
__device__ __forceinline__ void stats_flush(unsigned *stats, int n, float lo, float hi, float *sm) {
    block_minmax(lo, hi, sm);
    if (threadIdx.x == 0) {
        atomicMax(stats + 2 * n, enc_ordered(-lo));
        atomicMax(stats + 2 * n + 1, enc_ordered(hi));
    }
}

}  // namespace dfq
