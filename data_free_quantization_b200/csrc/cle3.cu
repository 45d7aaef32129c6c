// Kernel (b), resident form: cross-layer equalisation (Cross_layer_equal.py:11-59, 63-115) with every chain's
// weights held in the SHARED MEMORY of a group of SMs for the whole convergence loop.
//
// MobileNetV2 is 3.47 M weights = 13.9 MB (DeepLab 5.8 M = 23 MB); the chip has 148 x 227 KB = 33.6 MB of shared
// memory.  The streaming form (cle2.cu) re-reads every tensor from L2 twice per iteration behind one GRID barrier
// per chain position (264 barriers of ~10 us for MobileNetV2: latency-bound at 0.12 of the HBM roofline).  Here
//   * the 16 chains are independent (a chain = T0 -> T1 -> ... -> Tm, pair k = (Tk, Tk+1)), so each chain gets its
//     own group of CTAs (sized by its element count);
//   * a CTA owns whole ROWS of every tensor of its chain, loaded once; per iteration each tensor is visited in ONE
//     phase: columns *= 1/s_{k-1} (published by the owners of Tk-1's rows), row ranges, s_k = f(r1, r2) with r2 the
//     column ranges of Tk+1 left by the previous iteration, rows *= s_k, |W - W_old| summed on the fly (W_old is the
//     value just read: no second copy), column-range partials of the result for the next iteration;
//   * only the per-channel vectors cross SMs (1/s, column-range partials, metric partials: a few KB through L2).
//     There is NO barrier between the phases of an iteration: the owner of a channel publishes {iteration, 1/s_k} as
//     one 64-bit word as soon as it has it (c3_publish), and a consumer polls exactly the words its columns need
//     (c3_consume) -- the wait and the data fetch are the same L2 round trip, and a CTA whose producers are done never
//     waits for the slowest CTA of the group; the column partials it needs next are prefetched before it polls;
//   * one grid-wide barrier per ITERATION carries the convergence metric (it also bounds how far a producer can run
//     ahead, so the iteration tags never skip); every CTA evaluates Cross_layer_equal.py:107-115 identically from the
//     same partials in the same order, so the loop stops exactly where the reference stops.
// MobileNetV2 on B200: 44 iterations in 0.93 ms (19.7 us per iteration; cle2.cu 2.78 ms), 14.4 MB of DRAM traffic.
// Same arithmetic and the same order of roundings per element as cle.cu / cle2.cu (bit-identical results).
#include <algorithm>
#include <vector>
#include "common.cuh"

namespace dfq {

constexpr int kC3Threads = 512;
constexpr int kC3Warps = kC3Threads / 32;
constexpr int kC3MaxLen = 32;          // tensors per chain
constexpr int kC3BarStride = 32;       // one 128-byte line per barrier counter
constexpr int kC3StaticSmem = 4096;    // static __shared__ of the kernel (upper bound used by the planner)

struct Cle3Args {
    const DfqCle3Tensor *tensors; const DfqCle3Chain *chains; int n_chains, n_tensors;
    unsigned *bars;                  // [n_chains + 1] counters, kC3BarStride apart; only the last (grid-wide) is used
    unsigned long long *s_tab;       // [total_C]: {iteration tag, 1/s} of every pair's channels, one 64-bit word each
    unsigned *cp;                    // column-range partials, encoded max(-v) | max(v)
    double *dp;                      // [2 parities][dp_doubles] metric partials
    int total_C, dp_doubles;
    int fvec_floats, cp_words, max_rows;
    double smin, smax; float smin_f, smax_f, inv_smin_f, inv_smax_f;
    int is_signed; float eps; double threshold; int count; int max_iters;
    int *res_i; double *res_d;
    unsigned long long *dbg;         // optional [grid][kC3MaxLen + 1][8] globaltimer stamps of iteration 2 (tuning aid)
};

// Cross_layer_equal.py:34-43 for one channel (identical to cle2_scale)
__device__ __forceinline__ void c3_scale(const Cle3Args &a, float r1, float r2, float &s, float &sinv) {
    float inv = __fdiv_rn(1.0f, __fadd_rn(r1, a.eps));
    float rt = __fsqrt_rn(__fadd_rn(__fmul_rn(r1, r2), a.eps));
    float st = __fmul_rn(inv, rt);
    if ((double)st < a.smax) {                 // Python min(): keeps the tensor when s < smax
        if ((double)st > a.smin) { s = st; sinv = __fdiv_rn(1.0f, st); }
        else { s = a.smin_f; sinv = a.inv_smin_f; }
    } else if (a.smax > a.smin) { s = a.smax_f; sinv = a.inv_smax_f; }   // also NaN
    else { s = a.smin_f; sinv = a.inv_smin_f; }
}

__device__ __forceinline__ unsigned long long c3_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#define C3_STAMP(slot) do { if (stamp) dbgp[(slot)] = c3_now(); } while (0)

__device__ __forceinline__ void c3_arrive(unsigned *bar) {     // one thread, after a __syncthreads
    // release at gpu scope: cumulative over everything the CTA wrote before the barrier (no separate membar)
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(bar) : "memory");
}
__device__ __forceinline__ void c3_wait(const unsigned *bar, unsigned target) {   // one thread, before a __syncthreads
    // relaxed polls, ONE acquire fence at the end: an acquire load invalidates the L1 on every poll
    const long long t0 = clock64();
    for (;;) {
        unsigned v;
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar) : "memory");
        if (v >= target) break;
        if (clock64() - t0 > (1ll << 33)) __trap();            // ~4 s: a lost peer must not hang the GPU
    }
    asm volatile("fence.acq_rel.gpu;" ::: "memory");
}

// Flag-in-data hand-off of the column factors: the owner of a channel publishes {iteration tag, 1/s} as ONE 64-bit
// word (a single-copy-atomic store), a consumer polls exactly the words it needs until the tag is this iteration's.
// No counter, no fence, no separate vector load: the barrier wait and the data fetch are the same L2 round trip.
// A producer cannot run a whole iteration ahead of a consumer (one grid barrier per iteration), so tags never skip.
__device__ __forceinline__ void c3_publish(unsigned long long *slot, unsigned tag, float v) {
    const unsigned long long w = ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v);
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(slot), "l"(w) : "memory");
}
__device__ __forceinline__ float c3_consume(const unsigned long long *slot, unsigned tag) {
    const long long t0 = clock64();
    for (;;) {
        unsigned long long w;
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(slot) : "memory");
        if ((unsigned)(w >> 32) == tag) return __uint_as_float((unsigned)w);
        if (clock64() - t0 > (1ll << 33)) __trap();            // ~4 s: a lost peer must not hang the GPU
    }
}

// Rescale pass over this CTA's rows of one tensor (in shared memory): optional column factors (fvec), optional row
// factors (s_rows), metric, column-range partials into cp_sm (encoded, atomicMax).  Thread t keeps a fixed position j
// inside the row, so the column factor and the column extrema live in registers; consecutive threads touch
// consecutive words (conflict-free).
__device__ __forceinline__ void c3_rescale(float *__restrict__ w, int nrows, int L, int KK, bool dw, bool cs,
                                           const float *__restrict__ fvec, bool rs, const float *__restrict__ s_rows,
                                           unsigned *__restrict__ cp_sm, int cp_half, bool sgn, bool want_cp,
                                           bool want_diff, double &part) {
    // (all pointers are shared-memory addresses derived from the kernel's dynamic array and never null -- `cs` / `rs`
    // say whether the factors apply -- so the loop compiles to LDS / STS without generic-address conversions)
    // The metric sum_e |W - W_old| is summed in fp32 per thread (a few dozen terms) and in double across threads and
    // CTAs: the reference's own torch.mean is an fp32 reduction, and the loop's tests (:81, :110) have margins of
    // 7 % and 1e-9 absolute against sums of ~1e-7 .. 1e-3.
    float acc_s = 0.0f;
    int R, j0, rsub, step;
    if (L <= kC3Threads) { R = kC3Threads / L; j0 = threadIdx.x % L; rsub = threadIdx.x / L; step = L; }
    else { R = 1; j0 = threadIdx.x; rsub = 0; step = kC3Threads; }
    if (rsub >= R) { part = 0.0; return; }
    for (int j = j0; j < L; j += step) {
        const float fc = (cs && !dw) ? fvec[j] : 1.0f;
        float lo = INFINITY, hi = -INFINITY;
        float *p = w + rsub * L + j;
        const int stride = R * L;
        for (int r = rsub; r < nrows; r += R, p += stride) {
            const float old = *p;
            float v = old;
            if (cs) v = __fmul_rn(v, dw ? fvec[r] : fc);
            if (rs) v = __fmul_rn(v, s_rows[r]);
            if (cs || rs) *p = v;
            if (want_diff) acc_s = __fadd_rn(acc_s, fabsf(__fsub_rn(v, old)));
            if (want_cp) {
                const float m = sgn ? fabsf(v) : v;
                if (dw) {                                       // the row is its own column
                    atomicMax(cp_sm + r, enc_ordered(-m)); atomicMax(cp_sm + cp_half + r, enc_ordered(m));
                } else { lo = fminf(lo, m); hi = fmaxf(hi, m); }
            }
        }
        if (want_cp && !dw && hi >= lo) {
            const int c = (KK == 1) ? j : j / KK;
            atomicMax(cp_sm + c, enc_ordered(-lo)); atomicMax(cp_sm + cp_half + c, enc_ordered(hi));
        }
    }
    part = (double)acc_s;
}

// The same pass for ungrouped 1 x 1 tensors whose rows are a multiple of 4 long: a thread owns FOUR adjacent columns
// (128-bit LDS / STS), rows strided as above; ~9 instructions per element instead of ~50.
__device__ __forceinline__ void c3_rescale4(float *__restrict__ w, int nrows, int L, bool cs,
                                            const float *__restrict__ fvec, bool rs, const float *__restrict__ s_rows,
                                            unsigned *__restrict__ cp_sm, int cp_half, bool sgn, bool want_cp,
                                            double &part) {
    const int L4 = L >> 2;
    float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    int R, j0, rsub, step;
    if (L4 <= kC3Threads) { R = kC3Threads / L4; j0 = threadIdx.x % L4; rsub = threadIdx.x / L4; step = L4; }
    else { R = 1; j0 = threadIdx.x; rsub = 0; step = kC3Threads; }
    part = 0.0;
    if (rsub >= R) return;
    const float4 *f4 = reinterpret_cast<const float4 *>(fvec);
    for (int j = j0; j < L4; j += step) {
        const float4 f = cs ? f4[j] : make_float4(1.0f, 1.0f, 1.0f, 1.0f);
        float lo[4] = {INFINITY, INFINITY, INFINITY, INFINITY}, hi[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        float4 *p = reinterpret_cast<float4 *>(w) + rsub * L4 + j;
        const int stride = R * L4;
        for (int r = rsub; r < nrows; r += R, p += stride) {
            const float4 old = *p;
            float v[4] = {old.x, old.y, old.z, old.w};
            const float o[4] = {old.x, old.y, old.z, old.w};
            const float ff[4] = {f.x, f.y, f.z, f.w};
            const float sr = rs ? s_rows[r] : 1.0f;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (cs) v[q] = __fmul_rn(v[q], ff[q]);
                if (rs) v[q] = __fmul_rn(v[q], sr);
                acc[q] = __fadd_rn(acc[q], fabsf(__fsub_rn(v[q], o[q])));
                const float m = sgn ? fabsf(v[q]) : v[q];
                lo[q] = fminf(lo[q], m); hi[q] = fmaxf(hi[q], m);
            }
            if (cs || rs) *p = make_float4(v[0], v[1], v[2], v[3]);
        }
        if (want_cp && hi[0] >= lo[0]) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                atomicMax(cp_sm + 4 * j + q, enc_ordered(-lo[q]));
                atomicMax(cp_sm + cp_half + 4 * j + q, enc_ordered(hi[q]));
            }
        }
    }
    part = ((double)acc[0] + (double)acc[1]) + ((double)acc[2] + (double)acc[3]);
}

__global__ void __launch_bounds__(kC3Threads, 1)
cle3_kernel(Cle3Args a) {
    extern __shared__ __align__(16) float sm[];
    __shared__ int s_off[kC3MaxLen], s_row0[kC3MaxLen], s_nrows[kC3MaxLen], s_vec[kC3MaxLen];
    __shared__ int s_ws;
    __shared__ DfqCle3Tensor s_T[kC3MaxLen];       // this chain's tensor records: no global metadata reads in the loop
    __shared__ double s_red[kC3Warps];
    __shared__ double st_diff;
    __shared__ int st_stall, st_iters;

    int c = 0;
    while (c + 1 < a.n_chains && (int)blockIdx.x >= a.chains[c].cta0 + a.chains[c].ncta) ++c;
    const DfqCle3Chain ch = a.chains[c];
    const int g = blockIdx.x - ch.cta0, G = ch.ncta;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool sgn = a.is_signed;
    unsigned *allbar = a.bars + (size_t)a.n_chains * kC3BarStride;

    for (int i = tid; i < ch.len; i += kC3Threads) s_T[i] = a.tensors[ch.t0 + i];
    __syncthreads();
    if (tid == 0) {
        int off = 0;
        for (int i = 0; i < ch.len; ++i) {
            const DfqCle3Tensor &T = s_T[i];
            const int r0 = (int)((long long)g * T.O / G), r1 = (int)((long long)(g + 1) * T.O / G);
            s_row0[i] = r0; s_nrows[i] = r1 - r0; s_off[i] = off;
            off += ((r1 - r0) * T.L + 3) & ~3;          // 16-byte aligned regions (128-bit accesses)
        }
        for (int i = 0; i < ch.len; ++i) {
            s_vec[i] = off;
            if (s_T[i].down >= 0) off += 4 * s_nrows[i];
        }
        s_ws = (off + 3) & ~3;
    }
    __syncthreads();
    float *fvec = sm + s_ws;
    unsigned *cp_sm = reinterpret_cast<unsigned *>(fvec + a.fvec_floats);
    const int cp_half = a.cp_words / 2;
    unsigned *r2_sm = cp_sm + a.cp_words;                      // [2][max_rows]
    unsigned *rr_sm = r2_sm + 2 * a.max_rows;                  // [2][max_rows] row ranges met across warps
    float *s_sm = reinterpret_cast<float *>(rr_sm + 2 * a.max_rows);   // [max_rows]
    double *dp_sm = reinterpret_cast<double *>(s_sm + a.max_rows);     // [dp_doubles + n_tensors] (8-byte aligned)
    double *dec_n = dp_sm + a.dp_doubles + a.n_tensors;               // [n_tensors] element counts
    int *dec_off = reinterpret_cast<int *>(dec_n + a.n_tensors);      // [n_tensors] first partial | [n_tensors] partials
    for (int t = tid; t < a.n_tensors; t += kC3Threads) {
        const DfqCle3Tensor &T = a.tensors[t];
        dec_n[t] = (double)T.O * (double)T.L; dec_off[t] = T.dp_off; dec_off[a.n_tensors + t] = T.G;
    }

    // ---- load: this CTA's rows of every tensor of its chain, the per-channel vectors of the pairs it owns
    for (int i = 0; i < ch.len; ++i) {
        const DfqCle3Tensor &T = s_T[i];
        const int n = s_nrows[i] * T.L;
        const float *__restrict__ src = T.W + (size_t)s_row0[i] * T.L;
        float *dst = sm + s_off[i];
        for (int e = tid; e < n; e += kC3Threads) dst[e] = src[e];
        if (T.down >= 0) {
            float *v = sm + s_vec[i];
            const int nr = s_nrows[i], r0 = s_row0[i];
            for (int r = tid; r < nr; r += kC3Threads) {
                v[r] = T.S[r0 + r]; v[nr + r] = T.b1[r0 + r];
                v[2 * nr + r] = T.bn_w ? T.bn_w[r0 + r] : 0.0f;
                v[3 * nr + r] = T.bn_b ? T.bn_b[r0 + r] : 0.0f;
            }
        }
    }
    for (int i = tid; i < a.cp_words; i += kC3Threads) cp_sm[i] = 0u;
    for (int i = tid; i < 4 * a.max_rows; i += kC3Threads) r2_sm[i] = 0u;      // r2_sm and rr_sm
    __syncthreads();

    // column-range partials of tensor i (just written to cp_sm) -> global, cp_sm cleared for the next user
    auto flush_cp = [&](const DfqCle3Tensor &T, int i) {
        __syncthreads();
        if (T.dw) {
            unsigned *dst = a.cp + T.cp_off + s_row0[i];
            for (int r = tid; r < s_nrows[i]; r += kC3Threads) {
                dst[r] = cp_sm[r]; dst[T.C + r] = cp_sm[cp_half + r];
                cp_sm[r] = 0u; cp_sm[cp_half + r] = 0u;
            }
        } else {
            unsigned *dst = a.cp + T.cp_off + (size_t)g * 2 * T.C;
            for (int cc = tid; cc < T.C; cc += kC3Threads) {
                dst[cc] = cp_sm[cc]; dst[T.C + cc] = cp_sm[cp_half + cc];
                cp_sm[cc] = 0u; cp_sm[cp_half + cc] = 0u;
            }
        }
    };

    // ---- prologue: column ranges of the initial state (old_graph = deepcopy(graph), :83, is the shared copy itself)
    {
        double dummy = 0.0;
        for (int i = 0; i < ch.len; ++i) {
            const DfqCle3Tensor &T = s_T[i];
            if (T.up < 0) continue;
            c3_rescale(sm + s_off[i], s_nrows[i], T.L, T.KK, T.dw != 0, false, fvec, false, s_sm, cp_sm, cp_half, sgn,
                       true, false, dummy);
            flush_cp(T, i);
        }
    }
    unsigned all_epoch = 1;
    __syncthreads();
    if (tid == 0) { c3_arrive(allbar); c3_wait(allbar, all_epoch * gridDim.x); }
    __syncthreads();

    double diff = 1e8;                                // :67
    int stall = 0, iters = 0;                         // :68
    while (diff > a.threshold && stall < a.count && iters < a.max_iters) {     // :81
        const int par = iters & 1;
        const bool stamp = a.dbg && iters == 2 && tid == 0;
        unsigned long long *dbgp = a.dbg ? a.dbg + (size_t)blockIdx.x * (kC3MaxLen + 1) * 8 : nullptr;
        for (int i = 0; i < ch.len; ++i, dbgp += 8) {
            const DfqCle3Tensor &T = s_T[i];
            C3_STAMP(0);
            float *w = sm + s_off[i];
            const int nrows = s_nrows[i], row0 = s_row0[i], L = T.L;
            const bool up = T.up >= 0, down = T.down >= 0;
            const bool vec4 = (L & 3) == 0 && !T.dw && T.KK == 1 && L >= 16;     // 128-bit shared accesses
            // column ranges of T(i+1) for my rows: left by the previous iteration, so their loads are ISSUED before the
            // factors of this phase are polled (up to four per thread ride in registers behind the polling loop; r2_sm /
            // rr_sm were cleared behind the previous phase's use)
            unsigned pre_lo[4], pre_hi[4];
            int n_items = 0;
            const unsigned *cpn = nullptr;
            bool next_dw = false;
            int next_C = 0;
            if (down) {
                const DfqCle3Tensor &Tn = s_T[i + 1];
                cpn = a.cp + Tn.cp_off; next_dw = Tn.dw != 0; next_C = Tn.C;
                n_items = next_dw ? nrows : nrows * G;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int it = tid + u * kC3Threads;
                    pre_lo[u] = pre_hi[u] = 0u;
                    if (it < n_items) {
                        const int gg = next_dw ? 0 : it / nrows, r = next_dw ? it : it - gg * nrows;
                        const unsigned *p = cpn + (size_t)gg * 2 * next_C + row0 + r;
                        pre_lo[u] = __ldcg(p); pre_hi[u] = __ldcg(p + next_C);
                    }
                }
            }
            C3_STAMP(1);
            // the column factors 1/s_{i-1}, each polled from the word its owner publishes (tag = this iteration)
            if (up) {
                const unsigned long long *sinv = a.s_tab + T.up_off;
                const unsigned tag = (unsigned)iters + 1u;
                if (T.dw) { for (int r = tid; r < nrows; r += kC3Threads) fvec[r] = c3_consume(sinv + row0 + r, tag); }
                else if (T.KK == 1) { for (int j = tid; j < L; j += kC3Threads) fvec[j] = c3_consume(sinv + j, tag); }
                else { for (int j = tid; j < L; j += kC3Threads) fvec[j] = c3_consume(sinv + j / T.KK, tag); }
            }
            if (down) {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int it = tid + u * kC3Threads;
                    if (it < n_items) {
                        const int r = next_dw ? it : it % nrows;
                        atomicMax(r2_sm + r, pre_lo[u]); atomicMax(r2_sm + a.max_rows + r, pre_hi[u]);
                    }
                }
                for (int it = tid + 4 * kC3Threads; it < n_items; it += kC3Threads) {       // (groups wider than 4 x 512 items)
                    const int gg = next_dw ? 0 : it / nrows, r = next_dw ? it : it - gg * nrows;
                    const unsigned *p = cpn + (size_t)gg * 2 * next_C + row0 + r;
                    atomicMax(r2_sm + r, __ldcg(p)); atomicMax(r2_sm + a.max_rows + r, __ldcg(p + next_C));
                }
            }
            __syncthreads();
            C3_STAMP(2);
            // row ranges of the column-scaled rows -> s, 1/s, per-channel vectors (:38-57, relation.py:26-30)
            if (down) {
                float *vec = sm + s_vec[i];
                auto finish = [&](int r, float r1) {
                    const float lo2 = -dec_ordered(r2_sm[r]), hi2 = dec_ordered(r2_sm[a.max_rows + r]);
                    const float r2v = sgn ? hi2 : __fsub_rn(hi2, lo2);
                    float s, sinv;
                    c3_scale(a, r1, r2v, s, sinv);
                    s_sm[r] = s;
                    c3_publish(a.s_tab + T.s_off + row0 + r, (unsigned)iters + 1u, sinv);
                    vec[r] = __fmul_rn(vec[r], s);
                    vec[nrows + r] = __fmul_rn(vec[nrows + r], s);
                    vec[2 * nrows + r] = __fmul_rn(vec[2 * nrows + r], s);
                    vec[3 * nrows + r] = __fmul_rn(vec[3 * nrows + r], s);
                    r2_sm[r] = 0u; r2_sm[a.max_rows + r] = 0u;           // ready for the next phase's prefetch
                };
                if (vec4) {
                    // P lanes per row (a whole warp when the CTA owns few rows, 8 lanes otherwise), 128-bit loads of
                    // the row and of the column factors, log2(P) shuffle steps, the group's first lane finishes
                    const int P = nrows <= 16 ? 32 : 8;
                    const int gid = tid / P, lp = tid % P, groups = kC3Threads / P, L4 = L >> 2;
                    const float4 *f4 = reinterpret_cast<const float4 *>(fvec);
                    for (int rb = 0; rb < nrows; rb += groups) {          // (uniform trip count: shuffles inside)
                        const int r = rb + gid;
                        float lo = INFINITY, hi = -INFINITY;
                        if (r < nrows) {
                            const float4 *row = reinterpret_cast<const float4 *>(w + r * L);
                            for (int j = lp; j < L4; j += P) {
                                float4 v = row[j];
                                if (up) {
                                    const float4 f = f4[j];
                                    v.x = __fmul_rn(v.x, f.x); v.y = __fmul_rn(v.y, f.y);
                                    v.z = __fmul_rn(v.z, f.z); v.w = __fmul_rn(v.w, f.w);
                                }
                                if (sgn) { v.x = fabsf(v.x); v.y = fabsf(v.y); v.z = fabsf(v.z); v.w = fabsf(v.w); }
                                lo = fminf(fminf(lo, fminf(v.x, v.y)), fminf(v.z, v.w));
                                hi = fmaxf(fmaxf(hi, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
                            }
                        }
                        for (int o = P >> 1; o; o >>= 1) {
                            lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
                            hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
                        }
                        if (r < nrows && lp == 0) finish(r, sgn ? hi : __fsub_rn(hi, lo));
                    }
                } else if (L >= 32) {
                    // a warp takes 128-element segments of rows (all 16 warps busy however few rows the CTA owns);
                    // segment extrema meet in rr_sm, then one thread per row finishes (division, sqrt) in parallel
                    const int segs = (L + 127) >> 7;
                    for (int it = warp; it < nrows * segs; it += kC3Warps) {
                        const int r = it / segs, j0 = (it - r * segs) << 7;
                        const float *row = w + r * L;
                        float lo = INFINITY, hi = -INFINITY;
                        const float fr = (up && T.dw) ? fvec[r] : 1.0f;
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int j = j0 + lane + 32 * u;
                            if (j < L) {
                                float v = row[j];
                                if (up) v = __fmul_rn(v, T.dw ? fr : fvec[j]);
                                if (sgn) v = fabsf(v);
                                lo = fminf(lo, v); hi = fmaxf(hi, v);
                            }
                        }
                        lo = warp_min(lo); hi = warp_max(hi);
                        if (lane == 0) { atomicMax(rr_sm + r, enc_ordered(-lo)); atomicMax(rr_sm + a.max_rows + r, enc_ordered(hi)); }
                    }
                    __syncthreads();
                    for (int r = tid; r < nrows; r += kC3Threads) {
                        const float lo = -dec_ordered(rr_sm[r]), hi = dec_ordered(rr_sm[a.max_rows + r]);
                        rr_sm[r] = 0u; rr_sm[a.max_rows + r] = 0u;
                        finish(r, sgn ? hi : __fsub_rn(hi, lo));
                    }
                } else {                              // thread per row (depthwise 3x3, tiny pointwise)
                    for (int r = tid; r < nrows; r += kC3Threads) {
                        const float *row = w + r * L;
                        float lo = INFINITY, hi = -INFINITY;
                        const float fr = (up && T.dw) ? fvec[r] : 1.0f;
                        for (int j = 0; j < L; ++j) {
                            float v = row[j];
                            if (up) v = __fmul_rn(v, T.dw ? fr : fvec[j]);
                            if (sgn) v = fabsf(v);
                            lo = fminf(lo, v); hi = fmaxf(hi, v);
                        }
                        finish(r, sgn ? hi : __fsub_rn(hi, lo));
                    }
                }
                __syncthreads();
                C3_STAMP(3);
            }
            C3_STAMP(4);
            // rescale pass behind the barrier signal
            double part = 0.0;
            if (vec4) c3_rescale4(w, nrows, L, up, fvec, down, s_sm, cp_sm, cp_half, sgn, up, part);
            else c3_rescale(w, nrows, L, T.KK, T.dw != 0, up, fvec, down, s_sm, cp_sm, cp_half, sgn, up, true, part);
            C3_STAMP(5);
            if (up) flush_cp(T, i);
            part = warp_sum(part);
            if (lane == 0) s_red[warp] = part;
            __syncthreads();
            if (tid == 0) {
                double t = 0.0;
#pragma unroll
                for (int q = 0; q < kC3Warps; ++q) t += s_red[q];
                a.dp[(size_t)par * a.dp_doubles + T.dp_off + g] = t;
            }
            C3_STAMP(6);
        }
        if (a.dbg) dbgp = a.dbg + ((size_t)blockIdx.x * (kC3MaxLen + 1) + kC3MaxLen) * 8;
        C3_STAMP(0);
        // ---- one grid-wide barrier per iteration: metric partials and next column ranges become visible
        ++all_epoch;
        __syncthreads();
        if (tid == 0) { c3_arrive(allbar); c3_wait(allbar, all_epoch * gridDim.x); }
        __syncthreads();
        C3_STAMP(1);
        // diff_tmp = np.sum([float(torch.mean(torch.abs(W - W_old))) ...]), :107-108 -- same value in every CTA
        // all partials in one round trip, then one thread per tensor, then one thread over the tensors in order:
        // every CTA does the same additions in the same order
        for (int q = tid; q < a.dp_doubles; q += kC3Threads) dp_sm[q] = __ldcg(a.dp + (size_t)par * a.dp_doubles + q);
        __syncthreads();
        for (int t = tid; t < a.n_tensors; t += kC3Threads) {
            const int q0 = dec_off[t], q1 = q0 + dec_off[a.n_tensors + t];
            double s = 0.0;
            for (int q = q0; q < q1; ++q) s += dp_sm[q];
            dp_sm[a.dp_doubles + t] = (double)(float)(s / dec_n[t]);
        }
        __syncthreads();
        if (tid == 0) {
            double tot = 0.0;
            for (int t = 0; t < a.n_tensors; ++t) tot += dp_sm[a.dp_doubles + t];
            if (fabs(diff - tot) > 1e-9) { stall = 0; diff = tot; } else { stall += 1; }   // :110-115
            iters += 1;
            st_diff = diff; st_stall = stall; st_iters = iters;
            if (blockIdx.x == 0) a.res_d[iters] = tot;
        }
        __syncthreads();
        diff = st_diff; stall = st_stall; iters = st_iters;
        __syncthreads();
        C3_STAMP(2);
    }

    // ---- write back
    for (int i = 0; i < ch.len; ++i) {
        const DfqCle3Tensor &T = s_T[i];
        const int n = s_nrows[i] * T.L;
        float *__restrict__ dst = T.W + (size_t)s_row0[i] * T.L;
        const float *src = sm + s_off[i];
        for (int e = tid; e < n; e += kC3Threads) dst[e] = src[e];
        if (T.down >= 0) {
            const float *v = sm + s_vec[i];
            const int nr = s_nrows[i], r0 = s_row0[i];
            for (int r = tid; r < nr; r += kC3Threads) {
                T.S[r0 + r] = v[r]; T.b1[r0 + r] = v[nr + r];
                if (T.bn_w) T.bn_w[r0 + r] = v[2 * nr + r];
                if (T.bn_b) T.bn_b[r0 + r] = v[3 * nr + r];
            }
        }
    }
    if (blockIdx.x == 0 && tid == 0) {
        a.res_i[0] = iters;
        a.res_i[1] = stall;
        a.res_d[0] = diff;
    }
}

}  // namespace dfq

using namespace dfq;

static const int kC3Unsupported = (int)cudaErrorNotSupported;

static int c3_smem_limit() {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) != cudaSuccess || v <= 0)
        return 232448;                                           // 227 KB (sm_100)
    return v;
}

// Groups the tensors into chains, deals the SMs to the chains in proportion to their size and sizes the shared
// memory.  Returns cudaErrorNotSupported when the relations are not simple chains of ungrouped / depthwise layers
// or do not fit into the chip's shared memory: the caller then uses the streaming form (dfq_cle2_run).
extern "C" int dfq_cle3_plan_host(const DfqClePair *pairs, int n_pairs, const DfqCleLayer *layers, int n_layers,
                                  int n_sm, int smem_limit, DfqCle3Tensor *tensors, DfqCle3Chain *chains,
                                  DfqCle3Plan *plan) {
    if (n_pairs <= 0 || n_layers <= 0) return (int)cudaErrorInvalidValue;
    if (n_sm <= 0) n_sm = sm_count();
    if (smem_limit <= 0) smem_limit = c3_smem_limit();
    std::vector<int> chan_off(n_pairs + 1, 0);
    for (int i = 0; i < n_pairs; ++i) {
        chan_off[i + 1] = chan_off[i] + pairs[i].C1;
        for (int j = 0; j < i; ++j)
            if ((pairs[i].bn_w && pairs[i].bn_w == pairs[j].bn_w) || (pairs[i].bn_b && pairs[i].bn_b == pairs[j].bn_b) ||
                pairs[i].S == pairs[j].S || pairs[i].b1 == pairs[j].b1)
                return kC3Unsupported;
    }
    struct Tmp { int up, down, O, L, KK, C, dw; };
    std::vector<Tmp> tt(n_layers);
    for (int l = 0; l < n_layers; ++l) {
        Tmp &T = tt[l];
        T.up = T.down = -1; T.KK = 1; T.C = 0; T.dw = 0; T.O = T.L = 0;
        for (int i = 0; i < n_pairs; ++i) {
            if (pairs[i].W2 == layers[l].W) { if (T.up >= 0) return kC3Unsupported; T.up = i; }
            if (pairs[i].W1 == layers[l].W) { if (T.down >= 0) return kC3Unsupported; T.down = i; }
        }
        if (T.up < 0 && T.down < 0) return kC3Unsupported;
        if (T.down >= 0) { T.O = pairs[T.down].C1; T.L = pairs[T.down].K1; }
        if (T.up >= 0) {
            const DfqClePair &u = pairs[T.up];
            if (T.down >= 0 && (T.O != u.O2 || T.L != u.I2 * u.KK || u.stage >= pairs[T.down].stage))
                return kC3Unsupported;
            T.O = u.O2; T.L = u.I2 * u.KK; T.KK = u.KK; T.C = u.C1;
            if (u.C1 == u.I2) T.dw = 0;                            // one group: the column is the input channel
            else if (u.I2 == 1 && u.O2 == u.C1) T.dw = 1;          // depthwise: the row is its own column
            else return kC3Unsupported;
        }
        if ((int64_t)T.O * T.L != layers[l].n || T.O < 1 || T.L < 1) return kC3Unsupported;
    }
    // chains in order of their heads; tensors of a chain contiguous in the output table
    std::vector<int> order;
    std::vector<std::pair<int, int>> spans;                       // (t0, len)
    for (int l = 0; l < n_layers; ++l) {
        if (tt[l].up >= 0) continue;
        const int t0 = (int)order.size();
        int cur = l;
        for (;;) {
            order.push_back(cur);
            if (tt[cur].down < 0) break;
            int nxt = -1;
            for (int m = 0; m < n_layers; ++m) if (layers[m].W == pairs[tt[cur].down].W2) nxt = m;
            if (nxt < 0) return kC3Unsupported;
            cur = nxt;
        }
        if ((int)order.size() - t0 > kC3MaxLen) return kC3Unsupported;
        spans.push_back({t0, (int)order.size() - t0});
    }
    if ((int)order.size() != n_layers) return kC3Unsupported;    // a cycle
    const int n_chains = (int)spans.size();
    if (n_chains > n_sm) return kC3Unsupported;

    auto rows_of = [&](int O, int g, int G) { return (int)((long long)(g + 1) * O / G) - (int)((long long)g * O / G); };
    // workspace (floats) and data (floats) of CTA g of a chain with G CTAs
    std::vector<int> Gs(n_chains, 1);
    int fvec = 0, cpw = 0, maxrows = 0;
    auto workspace = [&]() {
        fvec = 4; cpw = 4; maxrows = 4;
        for (int c = 0; c < n_chains; ++c)
            for (int i = 0; i < spans[c].second; ++i) {
                const Tmp &T = tt[order[spans[c].first + i]];
                const int mr = (T.O + Gs[c] - 1) / Gs[c];
                if (T.down >= 0) maxrows = std::max(maxrows, mr);
                if (T.up >= 0) {
                    fvec = std::max(fvec, T.dw ? mr : T.L);
                    cpw = std::max(cpw, 2 * (T.dw ? mr : T.C));
                }
            }
        fvec = (fvec + 3) & ~3; cpw = (cpw + 7) & ~7; maxrows = (maxrows + 3) & ~3;
        int dpn = 0;
        for (int c = 0; c < n_chains; ++c) dpn += Gs[c] * spans[c].second;
        return fvec + cpw + 5 * maxrows + 2 * (dpn + 3 * n_layers) + 8;
    };
    auto data_floats = [&](int c, int G) {
        long long worst = 0;
        for (int g = 0; g < G; ++g) {
            long long f = 0;
            for (int i = 0; i < spans[c].second; ++i) {
                const Tmp &T = tt[order[spans[c].first + i]];
                const int nr = rows_of(T.O, g, G);
                f += (((long long)nr * T.L + 3) & ~3LL) + (T.down >= 0 ? 4 * nr : 0);
            }
            worst = std::max(worst, f);
        }
        return worst;
    };
    std::vector<long long> elems(n_chains, 0);
    std::vector<int> maxG(n_chains, 1);
    for (int c = 0; c < n_chains; ++c)
        for (int i = 0; i < spans[c].second; ++i) {
            const Tmp &T = tt[order[spans[c].first + i]];
            elems[c] += (long long)T.O * T.L;
            maxG[c] = std::max(maxG[c], T.O);
        }
    const long long budget = (smem_limit - kC3StaticSmem) / 4;
    // smallest group per chain that fits, then the spare SMs go to whichever chain carries most per CTA
    for (int round = 0; round < 4096; ++round) {
        const int ws = workspace();
        bool grown = false;
        int used = 0;
        for (int c = 0; c < n_chains; ++c) {
            while (data_floats(c, Gs[c]) + ws > budget) {
                if (Gs[c] >= maxG[c]) return kC3Unsupported;
                ++Gs[c]; grown = true;
            }
            used += Gs[c];
        }
        if (used > n_sm) return kC3Unsupported;
        if (grown) continue;
        if (used == n_sm) break;
        int best = -1;
        double load = 0.0;
        for (int c = 0; c < n_chains; ++c) {
            const double l = (double)elems[c] / Gs[c];
            if (Gs[c] < maxG[c] && l > load) { load = l; best = c; }
        }
        if (best < 0 || load < 2048.0) break;                   // nothing left worth splitting
        ++Gs[best];
    }
    const int ws = workspace();
    long long need = 0;
    int cta = 0, cp_words = 0, dp = 0;
    for (int c = 0; c < n_chains; ++c) {
        need = std::max(need, data_floats(c, Gs[c]));
        chains[c].t0 = spans[c].first; chains[c].len = spans[c].second; chains[c].cta0 = cta; chains[c].ncta = Gs[c];
        cta += Gs[c];
        for (int i = 0; i < spans[c].second; ++i) {
            const int l = order[spans[c].first + i];
            const Tmp &T = tt[l];
            DfqCle3Tensor &D = tensors[spans[c].first + i];
            D.W = layers[l].W;
            D.S = D.b1 = D.bn_w = D.bn_b = nullptr;
            if (T.down >= 0) {
                const DfqClePair &p = pairs[T.down];
                D.S = p.S; D.b1 = p.b1; D.bn_w = p.bn_w; D.bn_b = p.bn_b;
            }
            D.O = T.O; D.L = T.L; D.KK = T.KK; D.C = T.C; D.dw = T.dw; D.up = T.up; D.down = T.down;
            D.s_off = T.down >= 0 ? chan_off[T.down] : 0;
            D.up_off = T.up >= 0 ? chan_off[T.up] : 0;
            D.cp_off = cp_words;
            if (T.up >= 0) cp_words += 2 * T.C * (T.dw ? 1 : Gs[c]);
            D.dp_off = dp; dp += Gs[c];
            D.G = Gs[c]; D.layer = l;
        }
    }
    plan->n_chains = n_chains; plan->n_tensors = n_layers; plan->grid = cta;
    plan->smem_bytes = (int)((need + ws) * 4 + 16);
    plan->total_C = chan_off[n_pairs]; plan->cp_words = cp_words; plan->dp_doubles = dp;
    plan->fvec_floats = fvec; plan->cp_sm_words = cpw; plan->max_rows = maxrows;
    if (plan->smem_bytes + kC3StaticSmem > smem_limit) return kC3Unsupported;
    return 0;
}

static int64_t c3_align(int64_t v) { return (v + 255) & ~(int64_t)255; }

extern "C" int64_t dfq_cle3_scratch_bytes(const DfqCle3Plan *p) {
    return c3_align((int64_t)(p->n_chains + 1) * kC3BarStride * 4) + c3_align((int64_t)2 * p->total_C * 4) +
           c3_align((int64_t)p->cp_words * 4) + c3_align((int64_t)2 * p->dp_doubles * 8);
}

extern "C" int dfq_cle3_run(const DfqCle3Tensor *tensors_dev, const DfqCle3Chain *chains_dev, const DfqCle3Plan *p,
                            void *scratch_dev, double smin, double smax, int is_signed, float eps, double threshold,
                            int count, int max_iters, void *result_dev, void *debug_dev, void *stream) {
    if (!p || !scratch_dev || p->grid < 1) return (int)cudaErrorInvalidValue;
    static int configured = 0;
    cudaError_t e;
    if (configured < p->smem_bytes) {
        e = cudaFuncSetAttribute(cle3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, p->smem_bytes);
        if (e != cudaSuccess) return (int)e;
        configured = p->smem_bytes;
    }
    Cle3Args a;
    a.tensors = tensors_dev; a.chains = chains_dev; a.n_chains = p->n_chains; a.n_tensors = p->n_tensors;
    char *s = (char *)scratch_dev;
    a.bars = (unsigned *)s; s += c3_align((int64_t)(p->n_chains + 1) * kC3BarStride * 4);
    a.s_tab = (unsigned long long *)s; s += c3_align((int64_t)2 * p->total_C * 4);
    a.cp = (unsigned *)s; s += c3_align((int64_t)p->cp_words * 4);
    a.dp = (double *)s;
    a.total_C = p->total_C; a.dp_doubles = p->dp_doubles;
    a.fvec_floats = p->fvec_floats; a.cp_words = p->cp_sm_words; a.max_rows = p->max_rows;
    a.smin = smin; a.smax = smax;
    a.smin_f = (float)smin; a.smax_f = (float)smax;
    a.inv_smin_f = (float)(1.0 / smin); a.inv_smax_f = (float)(1.0 / smax);
    a.is_signed = is_signed; a.eps = eps; a.threshold = threshold; a.count = count; a.max_iters = max_iters;
    a.res_i = (int *)result_dev;
    a.res_d = (double *)((char *)result_dev + 8);
    a.dbg = (unsigned long long *)debug_dev;
    e = cudaMemsetAsync(scratch_dev, 0, (size_t)dfq_cle3_scratch_bytes(p), (cudaStream_t)stream);
    if (e != cudaSuccess) return (int)e;
    void *args[] = {&a};
    e = cudaLaunchCooperativeKernel((void *)cle3_kernel, dim3(p->grid), dim3(kC3Threads), args, (size_t)p->smem_bytes,
                                    (cudaStream_t)stream);
    g_launches += 1;
    return (int)e;
}
