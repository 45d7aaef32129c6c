// Kernel (c): bias absorption (bias_absorption.py:9-80) with an optional fused weight clip
// (clip_weight.py:4-33), plus BN folding (utils/layer_transform.py:240-285) and the stand-alone clip.
// One pass over W2: a warp owns one output row, forms c = max(0, beta - N*gamma) on the fly, reduces
// (sum_k W2[o,i,k]) * c[i] and, when asked, writes the clamped row back.  8 B/elt with clip, 4 B/elt
// without.  The bias updates then run in relation order per element (fp32 adds do not commute).
#include "common.cuh"

namespace dfq {

__device__ __forceinline__ float absorb_c(float beta, float gamma, float N) {
    float v = __fsub_rn(beta, __fmul_rn(N, gamma));       // :48
    return (v < 0.0f) ? 0.0f : v;                          // clamp_(0), NaN propagates (:49)
}

__global__ void __launch_bounds__(kThreads)
absorb_wc_kernel(const DfqAbsorbPair *__restrict__ pairs, int n_pairs, int total_rows, float N,
                 int do_clip, float lo_c, float hi_c) {
    int row = blockIdx.x * kWarps + (threadIdx.x >> 5);
    if (row >= total_rows) return;
    int lo = 0, hi = n_pairs - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (pairs[mid].first_row <= row) lo = mid; else hi = mid - 1;
    }
    const DfqAbsorbPair &p = pairs[lo];
    const int o = row - p.first_row, lane = threadIdx.x & 31;
    const int G = p.C1 / p.I2, Og = p.O2 / G, g = o / Og, KK = p.KK;
    float *w = p.W2 + (int64_t)o * p.I2 * KK;
    double acc = 0.0;
    if (KK == 1) {
        // pointwise / linear rows: eight independent loads of W, beta and gamma in flight, then the SAME accumulation
        // order as the plain loop (a 1280-long row was 40 dependent L2 round trips per lane: 41 us for the model)
        constexpr int U = 8;
        for (int base = lane; base < p.I2; base += 32 * U) {
            float v[U], bb[U], bw[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int it = base + 32 * u;
                if (it < p.I2) { v[u] = w[it]; bb[u] = p.bn_b[g * p.I2 + it]; bw[u] = p.bn_w[g * p.I2 + it]; }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int it = base + 32 * u;
                if (it < p.I2) {
                    const float c = absorb_c(bb[u], bw[u], N);
                    const float rs = __fadd_rn(0.0f, v[u]);
                    if (do_clip) w[it] = fminf(fmaxf(v[u], lo_c), hi_c);
                    acc += (double)rs * (double)c;
                }
            }
        }
    } else {
        for (int it = lane; it < p.I2; it += 32) {
            int ci = g * p.I2 + it;
            float c = absorb_c(p.bn_b[ci], p.bn_w[ci], N);
            float rs = 0.0f;
            for (int k = 0; k < KK; ++k) {                     // torch.sum(W2.view(O, I, -1), -1), :52,:65
                float v = w[it * KK + k];
                rs = __fadd_rn(rs, v);
                if (do_clip) w[it * KK + k] = fminf(fmaxf(v, lo_c), hi_c);
            }
            acc += (double)rs * (double)c;                     // matmul(., c), :64-67
        }
    }
    acc = warp_sum(acc);
    if (lane == 0) p.wc[o] = (float)acc;
}

// b1 -= c ; fake_bias -= c ; b2 += wc  (:78-80), relations in list order for every element.
__global__ void absorb_bias_kernel(const DfqAbsorbPair *__restrict__ pairs, int n_pairs, float N) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    // The c of a pair depends only on ITS BatchNorm, which no other pair's update touches: the two loads of pair r + 1
    // are issued before pair r's read-modify-writes (those stay in list order).  37 dependent round trips became ~19.
    float nb = 0.0f, nw = 0.0f;
    bool have = false;
    for (int r = 0; r < n_pairs; ++r) {
        const DfqAbsorbPair &p = pairs[r];
        float bb = nb, bw = nw;
        if (!have && i < p.C1) { bb = p.bn_b[i]; bw = p.bn_w[i]; }
        have = false;
        if (r + 1 < n_pairs && pairs[r + 1].bn_b != p.bn_b && i < pairs[r + 1].C1) {
            nb = pairs[r + 1].bn_b[i]; nw = pairs[r + 1].bn_w[i];
            have = true;
        }
        if (i < p.C1) {
            float c = absorb_c(bb, bw, N);
            p.b1[i] = __fadd_rn(p.b1[i], -c);
            p.bn_b[i] = __fadd_rn(bb, -c);
            if (p.c) p.c[i] = c;
        }
        if (i < p.O2) p.b2[i] = __fadd_rn(p.b2[i], p.wc[i]);
    }
}

__global__ void __launch_bounds__(kThreads) clip_kernel(float *w, int64_t n, float lo, float hi) {
    int64_t begin = (int64_t)blockIdx.x * kChunk, end = min(begin + (int64_t)kChunk, n);
    float *p = w + begin;
    int64_t m = end - begin;
    if ((((uintptr_t)p) & 15) == 0) {
        float4 *p4 = reinterpret_cast<float4 *>(p);
        int64_t m4 = m >> 2;
        for (int64_t k = threadIdx.x; k < m4; k += kThreads) {
            float4 v = p4[k];
            v.x = fminf(fmaxf(v.x, lo), hi); v.y = fminf(fmaxf(v.y, lo), hi);
            v.z = fminf(fmaxf(v.z, lo), hi); v.w = fminf(fmaxf(v.w, lo), hi);
            p4[k] = v;
        }
        for (int64_t k = (m4 << 2) + threadIdx.x; k < m; k += kThreads) p[k] = fminf(fmaxf(p[k], lo), hi);
    } else {
        for (int64_t k = threadIdx.x; k < m; k += kThreads) p[k] = fminf(fmaxf(p[k], lo), hi);
    }
}

// utils/layer_transform.py:255-274 for one (conv, BN) pair; a warp owns one output channel.
__global__ void __launch_bounds__(kThreads)
merge_bn_kernel(float *W, float *b, int64_t O, int64_t K, const float *gamma, const float *beta,
                const float *mean, const float *var, float eps, float *fake_w, float *fake_b) {
    int64_t o = (int64_t)blockIdx.x * kWarps + (threadIdx.x >> 5);
    if (o >= O) return;
    int lane = threadIdx.x & 31;
    float sd = __fsqrt_rn(__fadd_rn(var[o], eps));
    float f = __fdiv_rn(gamma[o], sd);
    float *row = W + o * K;
    for (int64_t k = lane; k < K; k += 32) row[k] = __fmul_rn(row[k], f);      // :260
    if (lane == 0) {
        float sh = __fsub_rn(beta[o], __fdiv_rn(__fmul_rn(gamma[o], mean[o]), sd));
        b[o] = __fadd_rn(__fmul_rn(b[o], f), sh);                                // :269-270
        if (fake_w) fake_w[o] = fabsf(gamma[o]);                                 // :273
        if (fake_b) fake_b[o] = beta[o];                                         // :274
    }
}

// all (conv, BN) pairs of a model: a warp owns one output channel of one pair
__global__ void __launch_bounds__(kThreads)
merge_bn_multi_kernel(const DfqMergeDesc *__restrict__ descs, int n_desc, int total_rows, int reset_bn) {
    int row = blockIdx.x * kWarps + (threadIdx.x >> 5);
    if (row >= total_rows) return;
    int lo = 0, hi = n_desc - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (descs[mid].first_row <= row) lo = mid; else hi = mid - 1;
    }
    const DfqMergeDesc &d = descs[lo];
    const int64_t o = row - d.first_row;
    const int lane = threadIdx.x & 31;
    const float g = d.gamma[o], be = d.beta[o], mu = d.mean[o], va = d.var[o];
    float sd = __fsqrt_rn(__fadd_rn(va, d.eps));
    float f = __fdiv_rn(g, sd);
    float *rowp = d.W + o * d.K;
    for (int64_t k = lane; k < d.K; k += 32) rowp[k] = __fmul_rn(rowp[k], f);        // :260
    __syncwarp();
    if (lane == 0) {
        float sh = __fsub_rn(be, __fdiv_rn(__fmul_rn(g, mu), sd));
        d.b[o] = __fadd_rn(__fmul_rn(d.b[o], f), sh);                               // :269-270
        if (d.fake_weight) d.fake_weight[o] = fabsf(g);                              // :273
        if (d.fake_bias) d.fake_bias[o] = be;                                        // :274
        if (reset_bn) { d.gamma[o] = 1.0f; d.var[o] = 1.0f; d.beta[o] = 0.0f; d.mean[o] = 0.0f; }   // :277-280
    }
}

__global__ void __launch_bounds__(kThreads)
clip_multi_kernel(const DfqClipDesc *__restrict__ descs, int n_desc, float lo_c, float hi_c) {
    int lo = 0, hi = n_desc - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (descs[mid].first_block <= (int64_t)blockIdx.x) lo = mid; else hi = mid - 1;
    }
    const DfqClipDesc &d = descs[lo];
    int64_t begin = ((int64_t)blockIdx.x - d.first_block) * kChunk, end = min(begin + (int64_t)kChunk, d.n);
    float *p = d.w + begin;
    int64_t m = end - begin;
    if ((((uintptr_t)p) & 15) == 0) {
        float4 *p4 = reinterpret_cast<float4 *>(p);
        int64_t m4 = m >> 2;
        for (int64_t k = threadIdx.x; k < m4; k += kThreads) {
            float4 v = p4[k];
            v.x = fminf(fmaxf(v.x, lo_c), hi_c); v.y = fminf(fmaxf(v.y, lo_c), hi_c);
            v.z = fminf(fmaxf(v.z, lo_c), hi_c); v.w = fminf(fmaxf(v.w, lo_c), hi_c);
            p4[k] = v;
        }
        for (int64_t k = (m4 << 2) + threadIdx.x; k < m; k += kThreads) p[k] = fminf(fmaxf(p[k], lo_c), hi_c);
    } else {
        for (int64_t k = threadIdx.x; k < m; k += kThreads) p[k] = fminf(fmaxf(p[k], lo_c), hi_c);
    }
}

}  // namespace dfq

using namespace dfq;

extern "C" int dfq_merge_plan_host(DfqMergeDesc *d, int n, int *total_rows) {
    int64_t r = 0;
    for (int i = 0; i < n; ++i) {
        if (d[i].O <= 0 || d[i].K <= 0) return (int)cudaErrorInvalidValue;
        d[i].first_row = (int32_t)r;
        r += d[i].O;
    }
    if (r > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    *total_rows = (int)r;
    return 0;
}

extern "C" int dfq_merge_bn_multi(const DfqMergeDesc *descs_dev, int n_desc, int total_rows, int reset_bn,
                                  void *stream) {
    if (n_desc <= 0 || total_rows <= 0) return 0;
    merge_bn_multi_kernel<<<(total_rows + kWarps - 1) / kWarps, kThreads, 0, (cudaStream_t)stream>>>(
        descs_dev, n_desc, total_rows, reset_bn);
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_clip_plan_host(DfqClipDesc *d, int n, int64_t *total_blocks) {
    int64_t b = 0;
    for (int i = 0; i < n; ++i) {
        if (d[i].n < 0) return (int)cudaErrorInvalidValue;
        d[i].first_block = b;
        b += (d[i].n + kChunk - 1) / kChunk;
    }
    *total_blocks = b;
    return b > 0x7fffffffll ? (int)cudaErrorInvalidValue : 0;
}

extern "C" int dfq_clip_multi_f32(const DfqClipDesc *descs_dev, int n_desc, int64_t total_blocks, float lo, float hi,
                                  void *stream) {
    if (n_desc <= 0 || total_blocks <= 0) return 0;
    clip_multi_kernel<<<(unsigned)total_blocks, kThreads, 0, (cudaStream_t)stream>>>(descs_dev, n_desc, lo, hi);
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_absorb_plan_host(DfqAbsorbPair *p, int n, int *total_rows) {
    int r = 0;
    for (int i = 0; i < n; ++i) {
        if (p[i].I2 <= 0 || p[i].C1 % p[i].I2 || p[i].O2 % (p[i].C1 / p[i].I2)) return (int)cudaErrorInvalidValue;
        p[i].first_row = r;
        r += p[i].O2;
    }
    *total_rows = r;
    return 0;
}

extern "C" int dfq_absorb_clip(const DfqAbsorbPair *pairs_dev, int n_pairs, int total_rows, int max_dim, float N,
                               int do_clip, float clip_lo, float clip_hi, void *stream) {
    if (n_pairs <= 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    absorb_wc_kernel<<<(total_rows + kWarps - 1) / kWarps, kThreads, 0, st>>>(pairs_dev, n_pairs, total_rows, N,
                                                                              do_clip, clip_lo, clip_hi);
    absorb_bias_kernel<<<(max_dim + 255) / 256, 256, 0, st>>>(pairs_dev, n_pairs, N);
    g_launches += 2;
    return (int)cudaGetLastError();
}

extern "C" int dfq_clip_f32(float *w, int64_t n, float lo, float hi, void *stream) {
    if (n <= 0) return 0;
    int64_t blocks = (n + kChunk - 1) / kChunk;
    if (blocks > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    clip_kernel<<<(unsigned)blocks, kThreads, 0, (cudaStream_t)stream>>>(w, n, lo, hi);
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_merge_bn(float *W, float *b, int64_t O, int64_t K, const float *gamma, const float *beta,
                            const float *mean, const float *var, float eps, float *fake_weight,
                            float *fake_bias, void *stream) {
    if (O <= 0) return 0;
    merge_bn_kernel<<<(unsigned)((O + kWarps - 1) / kWarps), kThreads, 0, (cudaStream_t)stream>>>(
        W, b, O, K, gamma, beta, mean, var, eps, fake_weight, fake_bias);
    g_launches += 1;
    return (int)cudaGetLastError();
}
