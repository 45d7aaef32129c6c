// Kernel (d): bias correction -- bias_correction.py:111-144 (eps = Q(W) - W) fused with the
// reduction against a per-input-channel vector v (E[x] from the preceding BatchNorm), so eps never
// reaches HBM:  out[o] = sum_{i,k} (Q(W[o,i,k]) - W[o,i,k]) * v[i].   HBM-bound: 4 B/elt for the
// range pass + 4 B/elt for the reduction (outputs are 4*O bytes).
#include "common.cuh"

namespace dfq {

constexpr int kCorrSmallRow = 1024;

__device__ __forceinline__ int find_corr(const DfqCorrDesc *d, int n, int64_t b, bool by_block) {
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        int64_t f = by_block ? d[mid].first_block : d[mid].first_unit;
        if (f <= b) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__global__ void __launch_bounds__(kThreads)
corr_minmax_kernel(const DfqCorrDesc *__restrict__ descs, int n_desc, int64_t total_blocks,
                   unsigned *__restrict__ ranges) {
    __shared__ float sm[2 * kWarps];
    float lo = INFINITY, hi = -INFINITY;
    int cur = -1;
    // grid-stride over 16 KB chunks; one block reduction + atomic pair per layer change
    for (int64_t vb = blockIdx.x; vb < total_blocks; vb += gridDim.x) {
        int t = find_corr(descs, n_desc, vb, true);
        const DfqCorrDesc &d = descs[t];
        if (t != cur) {
            if (cur >= 0) {
                block_minmax(lo, hi, sm);
                if (threadIdx.x == 0) {
                    atomicMax(ranges + 4 * descs[cur].range_slot, enc_ordered(-lo));
                    atomicMax(ranges + 4 * descs[cur].range_slot + 1, enc_ordered(hi));
                }
            }
            lo = INFINITY; hi = -INFINITY; cur = t;
        }
        int64_t n = (int64_t)d.O * d.I2 * d.KK;
        int64_t begin = (vb - d.first_block) * (int64_t)kChunk;
        int64_t end = min(begin + (int64_t)kChunk, n);
        chunk_minmax(d.W, begin, end, lo, hi);
    }
    if (cur >= 0) {
        block_minmax(lo, hi, sm);
        if (threadIdx.x == 0) {
            atomicMax(ranges + 4 * descs[cur].range_slot, enc_ordered(-lo));
            atomicMax(ranges + 4 * descs[cur].range_slot + 1, enc_ordered(hi));
        }
    }
}

// {enc(-min), enc(max)} (encoded != 0) or caller-given {min, max} (given != 0) -> RangeSlot {min, max, scale, 0}
__global__ void corr_decode_kernel(const DfqCorrDesc *__restrict__ descs, unsigned *ranges, int64_t slots, int given) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= slots) return;
    float mn, mx;
    if (given) {
        mn = reinterpret_cast<float *>(ranges)[4 * i];
        mx = reinterpret_cast<float *>(ranges)[4 * i + 1];
    } else {
        mn = -dec_ordered(ranges[4 * i]);
        mx = dec_ordered(ranges[4 * i + 1]);
    }
    DfqQuantParams p = make_quant_params((double)mn, (double)mx, descs[i].bits, descs[i].flags);   // :128
    RangeSlot r = {mn, mx, p.scale, 0.0f};
    reinterpret_cast<RangeSlot *>(ranges)[i] = r;
}

__global__ void corr_zero_kernel(const DfqCorrDesc *__restrict__ descs) {
    const DfqCorrDesc &d = descs[blockIdx.x];
    for (int o = threadIdx.x; o < d.O; o += blockDim.x) d.out[o] = 0.0f;
    if (d.fold)
        for (int c = threadIdx.x; c < d.fold_mod; c += blockDim.x) d.fold[c] = 0.0f;
}

// partial sum over elements [b, e) of one row (indices relative to the row start)
__device__ __forceinline__ double row_partial(const float *__restrict__ row, const float *__restrict__ v,
                                              int64_t b, int64_t e, int KK, float inv_I,
                                              const QuantConsts &c, int tid, int nthr,
                                              float *fold, int fold_mod, int64_t row_chan0) {
    // a thread owns <= 32 elements of a chunk: fp32 partial (error ~1e-7 of the partial), widened once
    float acc = 0.0f;
    const float *p = row + b;
    int64_t n = e - b;
    bool vec = ((((uintptr_t)p) & 15) == 0);
    int64_t n4 = vec ? (n >> 2) : 0;
    const float4 *p4 = reinterpret_cast<const float4 *>(p);
    // pointwise / linear weights (KK == 1): v is read with the same 128-bit pattern as W
    const bool vfast = v && KK == 1 && (((uintptr_t)(v + b)) & 15) == 0;
    const float4 *v4 = reinterpret_cast<const float4 *>(v + b);
    for (int64_t base = 0; base < n4; base += 4 * (int64_t)nthr) {
        float4 w[4], vv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {                      // 4 independent 128-bit loads in flight per thread
            int64_t k = base + tid + (int64_t)u * nthr;
            if (k < n4) {
                w[u] = __ldg(p4 + k);
                if (vfast) vv[u] = __ldg(v4 + k);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            int64_t k = base + tid + (int64_t)u * nthr;
            if (k >= n4) continue;
            float xs[4] = {w[u].x, w[u].y, w[u].z, w[u].w};
            float vs[4] = {vv[u].x, vv[u].y, vv[u].z, vv[u].w};
            int64_t idx = b + 4 * k;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float q;
                float dq = fake_quant_elem(xs[j], c, q);
                float eps = __fsub_rn(dq, xs[j]);                   // bias_correction.py:130
                int64_t ch = (KK == 1) ? (idx + j) : (idx + j) / KK;
                float vi = vfast ? vs[j] : (v ? v[ch] : inv_I);
                acc = __fmaf_rn(eps, vi, acc);
                if (fold) atomicAdd(fold + (row_chan0 + ch) % fold_mod, eps);
            }
        }
    }
    for (int64_t k = (n4 << 2) + tid; k < n; k += nthr) {
        float x = p[k], q;
        float eps = __fsub_rn(fake_quant_elem(x, c, q), x);
        int64_t ch = (KK == 1) ? (b + k) : (b + k) / KK;
        float vi = v ? v[ch] : inv_I;
        acc = __fmaf_rn(eps, vi, acc);
        if (fold) atomicAdd(fold + (row_chan0 + ch) % fold_mod, eps);
    }
    return (double)acc;
}

__global__ void __launch_bounds__(kThreads)
corr_reduce_kernel(const DfqCorrDesc *__restrict__ descs, int n_desc, const RangeSlot *__restrict__ ranges) {
    __shared__ double smd[kWarps];
    int t = find_corr(descs, n_desc, blockIdx.x, false);
    const DfqCorrDesc &d = descs[t];
    const int64_t row_len = (int64_t)d.I2 * d.KK;
    QuantConsts c = consts_from_slot(ranges[d.range_slot], d.bits, d.flags, 0.f, 0.f);
    const float inv_I = 1.0f / (float)d.I2;
    int64_t j = blockIdx.x - d.first_unit;
    if (row_len <= kCorrSmallRow) {                       // one warp per row
        int64_t o = j * kWarps + (threadIdx.x >> 5);
        if (o >= d.O) return;
        const float *vrow = (d.v && d.groups > 1) ? d.v + (o / (d.O / d.groups)) * d.I2 : d.v;
        double acc = row_partial(d.W + o * row_len, vrow, 0, row_len, d.KK, inv_I, c, threadIdx.x & 31, 32,
                                 d.fold, d.fold_mod, o * d.I2);
        acc = warp_sum(acc);
        if ((threadIdx.x & 31) == 0) d.out[o] = (float)acc;
    } else {                                              // one block per (row, chunk)
        int64_t cpr = (row_len + kChunk - 1) / kChunk;
        int64_t o = j / cpr, b = (j % cpr) * kChunk;
        int64_t e = min(b + (int64_t)kChunk, row_len);
        const float *vrow = (d.v && d.groups > 1) ? d.v + (o / (d.O / d.groups)) * d.I2 : d.v;
        double acc = row_partial(d.W + o * row_len, vrow, b, e, d.KK, inv_I, c, threadIdx.x, kThreads,
                                 d.fold, d.fold_mod, o * d.I2);
        acc = warp_sum(acc);
        if ((threadIdx.x & 31) == 0) smd[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) s += smd[w];
            if (cpr == 1) d.out[o] = (float)s; else atomicAdd(d.out + o, (float)s);
        }
    }
}

// ---- streaming reduction for long pointwise / linear rows (KK == 1, v given, no fold) --------------------------
// eps . v over one FULL, 16-byte aligned 16 KB chunk: four 128-bit loads of W (evict-first) and of v (L2-resident,
// re-read by every row) in flight per thread, no index arithmetic per element.  kInRange: the range was reduced
// from this very tensor, so (x - min)/scale already lies in [0, qmax] and the overflow guard of the numerator is
// dead code.  NaN needs no fix-up: eps = dq - NaN is NaN whatever dq is.
template <bool kInRange>
__device__ __forceinline__ float chunk_eps_dot_pw(const float4 *__restrict__ w4, const float4 *__restrict__ v4,
                                                  const QuantConsts &c, int tid) {
    float4 w[4], vv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) { w[u] = __ldcs(w4 + tid + u * kThreads); vv[u] = __ldg(v4 + tid + u * kThreads); }
    float acc0 = 0.0f, acc1 = 0.0f;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const float xs[4] = {w[u].x, w[u].y, w[u].z, w[u].w};
        const float vs[4] = {vv[u].x, vv[u].y, vv[u].z, vv[u].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float t = __fadd_rn(xs[j], c.neg_min);
            if (!kInRange) t = fminf(fmaxf(t, c.t_lo), c.t_hi);
            float d = div_const_rn(t, c.scale, c.rscale);
            // kInRange: 0 <= t <= max - min, so d lies in [0, qmax (1 + 2^-23)] and rounds into the grid unclamped
            float q = rint_small(kInRange ? d : fminf(fmaxf(d, c.qmin), c.qmax));
            float eps = __fsub_rn(__fadd_rn(__fmul_rn(q, c.scale), c.add_back), xs[j]);   // bias_correction.py:130
            if (j & 1) acc1 = __fmaf_rn(eps, vs[j], acc1); else acc0 = __fmaf_rn(eps, vs[j], acc0);
        }
    }
    return __fadd_rn(acc0, acc1);
}

__device__ __forceinline__ int find_stream(const DfqCorrDesc *d, int n, int64_t b) {
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (d[mid].first_stream <= b) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// Persistent: block b owns the contiguous chunk range [b*U/G, (b+1)*U/G) of the stream-chunk space (row-major
// (layer, row, chunk)).  Consecutive chunks of one row accumulate in registers (fp32 inside a chunk, fp64 across
// chunks) and are flushed with one block reduction + one atomicAdd when the row changes: a 1 M-element row costs
// about two atomics instead of 256.
template <bool kInRange>
__global__ void __launch_bounds__(kThreads, 4)
corr_stream_kernel(const DfqCorrDesc *__restrict__ descs, int n_desc, int64_t total,
                   const RangeSlot *__restrict__ ranges) {
    __shared__ double smd[kWarps];
    const int64_t u0 = total * blockIdx.x / gridDim.x, u1 = total * (blockIdx.x + 1) / gridDim.x;
    double carry = 0.0;
    float *open = nullptr;              // &out[row] of the row being accumulated
    int64_t u = u0;
    while (u < u1) {                    // one iteration per (layer, row) segment
        const int t = find_stream(descs, n_desc, u);
        const float *W = descs[t].W, *v = descs[t].v;
        const int64_t row_len = descs[t].I2;
        const int64_t cpr = (row_len + kChunk - 1) / kChunk;
        const int64_t j = u - descs[t].first_stream;
        const int64_t o = j / cpr;
        int64_t ck = j - o * cpr;
        const int64_t seg_end = min(u1, u + (cpr - ck));
        const QuantConsts c = consts_from_slot(ranges[descs[t].range_slot], descs[t].bits, descs[t].flags, 0.f, 0.f);
        const float *row = W + o * row_len;
        open = descs[t].out + o;
        for (; u < seg_end; ++u, ++ck) {
            const int64_t b = ck * kChunk;
            if (c.fast && b + kChunk <= row_len) {
                carry += (double)chunk_eps_dot_pw<kInRange>(reinterpret_cast<const float4 *>(row + b),
                                                            reinterpret_cast<const float4 *>(v + b), c, threadIdx.x);
            } else {                    // ragged last chunk of the row, or an exotic scale
                float acc = 0.0f;
                for (int64_t k = b + threadIdx.x; k < min(b + (int64_t)kChunk, row_len); k += kThreads) {
                    float x = row[k], q;
                    acc = __fmaf_rn(__fsub_rn(fake_quant_elem(x, c, q), x), v[k], acc);
                }
                carry += (double)acc;
            }
        }
        // row (or this block's share of it) finished: block-uniform flush
        double acc = warp_sum(carry);
        __syncthreads();
        if ((threadIdx.x & 31) == 0) smd[threadIdx.x >> 5] = acc;
        __syncthreads();
        if (threadIdx.x == 0) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) s += smd[w];
            atomicAdd(open, (float)s);
        }
        carry = 0.0;
    }
}

__global__ void corr_finalize_kernel(const DfqCorrDesc *__restrict__ descs) {
    const DfqCorrDesc &d = descs[blockIdx.x];
    if (!d.bias) return;
    for (int o = threadIdx.x; o < d.O; o += blockDim.x)
        d.bias[o] = __fadd_rn(d.bias[o], __fmul_rn(d.alpha, d.out[o]));
}

}  // namespace dfq

using namespace dfq;

// a layer takes the streaming reduction when its rows are long, contiguous pointwise / linear rows with 16-byte
// aligned chunk starts and nothing but the eps.v sum is wanted
static bool corr_streams(const DfqCorrDesc &d) {
    return d.KK == 1 && d.I2 > kCorrSmallRow && d.v && d.groups <= 1 && !d.fold && (d.I2 % 4) == 0 &&
           ((((uintptr_t)d.W) | ((uintptr_t)d.v)) & 15) == 0;
}

extern "C" int dfq_corr_plan_host(DfqCorrDesc *d, int n, int64_t *total_units, int64_t *total_blocks,
                                  int64_t *total_stream) {
    int64_t u = 0, b = 0, s = 0;
    for (int i = 0; i < n; ++i) {
        if (d[i].O <= 0 || d[i].I2 <= 0 || d[i].KK <= 0 || !d[i].out) return (int)cudaErrorInvalidValue;
        int64_t row_len = (int64_t)d[i].I2 * d[i].KK, numel = row_len * d[i].O;
        d[i].first_unit = u;
        d[i].first_block = b;
        d[i].first_stream = s;
        d[i].range_slot = i;
        if (corr_streams(d[i]))
            s += (int64_t)d[i].O * ((row_len + kChunk - 1) / kChunk);
        else
            u += (row_len <= kCorrSmallRow) ? (d[i].O + kWarps - 1) / kWarps
                                            : (int64_t)d[i].O * ((row_len + kChunk - 1) / kChunk);
        b += (numel + kChunk - 1) / kChunk;
    }
    *total_units = u;
    *total_blocks = b;
    *total_stream = s;
    return (u > 0x7fffffffll || b > 0x7fffffffll) ? (int)cudaErrorInvalidValue : 0;
}

extern "C" int dfq_bias_correct(const DfqCorrDesc *descs_dev, int n_desc, int64_t total_units, int64_t total_blocks,
                                int64_t total_stream, float *ranges_dev, int range_given, void *stream) {
    if (n_desc <= 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    if (!range_given) {
        cudaError_t e = cudaMemsetAsync(ranges_dev, 0, sizeof(float) * 4 * n_desc, st);
        if (e != cudaSuccess) return (int)e;
        corr_minmax_kernel<<<reduce_grid(corr_minmax_kernel, total_blocks), kThreads, 0, st>>>(descs_dev, n_desc, total_blocks,
                                                                           (unsigned *)ranges_dev);
        g_launches += 1;
    }
    corr_decode_kernel<<<(n_desc + 255) / 256, 256, 0, st>>>(descs_dev, (unsigned *)ranges_dev, n_desc, range_given);
    g_launches += 1;
    corr_zero_kernel<<<n_desc, 256, 0, st>>>(descs_dev);
    g_launches += 2;
    if (total_units > 0) {
        corr_reduce_kernel<<<(unsigned)total_units, kThreads, 0, st>>>(descs_dev, n_desc, (const RangeSlot *)ranges_dev);
        g_launches += 1;
    }
    if (total_stream > 0) {
        int64_t g = (int64_t)sm_count() * 4;              // 4 resident blocks per SM (launch bounds)
        unsigned grid = (unsigned)(total_stream < g ? total_stream : g);
        if (range_given)
            corr_stream_kernel<false><<<grid, kThreads, 0, st>>>(descs_dev, n_desc, total_stream, (const RangeSlot *)ranges_dev);
        else
            corr_stream_kernel<true><<<grid, kThreads, 0, st>>>(descs_dev, n_desc, total_stream, (const RangeSlot *)ranges_dev);
        g_launches += 1;
    }
    corr_finalize_kernel<<<n_desc, 256, 0, st>>>(descs_dev);
    return (int)cudaGetLastError();
}
