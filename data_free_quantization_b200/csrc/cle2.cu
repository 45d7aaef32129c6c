// Kernel (b), streaming form: cross-layer equalisation (Cross_layer_equal.py:11-59, 63-115) as row-tile passes over
// whole tensors instead of per-channel gather units.  Same arithmetic, same order of roundings, so the results are
// bit-identical to the reference's sequential loop (and to csrc/cle.cu, which stays as the general fallback).
//
// For a chain T0 -> T1 -> ... -> Tm (pair k = (Tk, Tk+1)) one outer iteration of the reference does, per pair k in
// order:  r1 = range of every row of Tk,  r2 = range of every input-channel "column" of Tk+1,  s = f(r1, r2),
// Tk rows *= s,  Tk+1 columns *= 1/s.  Only the ROW ranges chain through the iteration (rows of Tk must have seen
// the column scaling of pair k-1); the column ranges of Tk+1 are those left by the previous iteration.  So every
// tensor is streamed exactly twice per iteration, coalesced, by all SMs:
//   phase k   "rows" pass over Tk : columns *= 1/s_{k-1} (none for a chain head), row ranges -> r1[k]
//   phase k+1 "cols" pass over Tk : rows *= s_k (s_k from r1[k], r2[k]); |W - W_old| summed, W_old <- W; the
//                                    column ranges of the result -> r2[k-1] of the NEXT iteration (atomics)
//   a chain tail Tm has one "cols" pass in phase m: columns *= 1/s_{m-1}, metric, W_old, next column ranges.
// One grid barrier per phase (chain length, 6 for MobileNetV2) instead of one per pair stage plus the metric pass,
// and no strided column gathers.  The convergence test (:107-115) is evaluated on the device as in cle.cu.
#include <algorithm>
#include <cooperative_groups.h>
#include "common.cuh"

namespace cg = cooperative_groups;

namespace dfq {

constexpr int kC2MaxPhases = 64;
constexpr int kC2Rows = 1024;        // most rows of a flat "cols" tile
constexpr int kC2ColRows = 32;       // rows of a 2-D tile
constexpr int kC2Strip = 64;         // columns of a 2-D tile
constexpr int kC2Tile = 2048;        // elements per tile (a column-owner tile has at least 4 rows)

// One self-contained work item: everything a block needs arrives with ONE struct load (the planned DfqCleTask /
// DfqCleTensor / DfqClePair tables are expanded into these records by cle2_expand_kernel before the loop starts), so
// a task costs two dependent memory round trips (record, data) instead of five (task, tensor, pair, offsets, data).
struct __align__(16) Cle2Fat {
    float *W, *W_old;                // tile base: row0 (and the column strip of a 2-D tile) applied
    float *s_tab;                    // down pair: s of row0.. (1/s at + total_C); rows pass writes, cols pass reads
    const float *sinv_up;            // up pair: 1/s by channel, or NULL
    unsigned *r2_up;                 // up pair: encoded column ranges of parity 0 (max at + total_C, parity 1 at + 2 total_C)
    unsigned *r2_down;               // down pair: same table at row0
    float *S, *b1, *bn_w, *bn_b;     // down pair's per-channel vectors at row0
    double *acc;                     // metric slot of parity 0 (parity 1 at + n_tensors)
    int row0, nrows, L, KK, I2, Og, O, kind, strip, up, down, pad;
};

struct Cle2Args {
    const DfqClePair *pairs; int n_pairs;
    const DfqCleTensor *tensors; int n_tensors;
    const DfqCleTask *tasks; int n_tasks;
    Cle2Fat *fat;
    int phase_begin[kC2MaxPhases + 1]; int n_phases;
    int n_init;                      // tasks [n_tasks, n_tasks + n_init) are the prologue's "cols" tasks of all tensors
    const int *chan_off;             // [n_pairs + 1] prefix of C1
    float *r1;                       // [2][total_C]: s | 1/s of every pair's channels, written by the rows pass
    unsigned *r2;                    // [2 parities][2][total_C] encoded max(-v), max(v)
    int total_C;
    double smin, smax; float smin_f, smax_f, inv_smin_f, inv_smax_f;
    int is_signed; float eps; double threshold; int count; int max_iters;
    double *acc;                     // [2][n_tensors]
    int *res_i; double *res_d;
    unsigned long long *dbg;         // optional [blocks][phases][5] globaltimer stamps of iteration 2 (tuning aid)
};

__global__ void cle2_expand_kernel(Cle2Args a) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= a.n_tasks + a.n_init) return;
    const DfqCleTask k = a.tasks[t];
    const DfqCleTensor T = a.tensors[k.tensor];
    Cle2Fat f;
    f.kind = k.kind & 0xff; f.strip = k.kind >> 8;
    f.row0 = k.row0; f.nrows = k.nrows; f.L = T.row_len; f.KK = T.KK; f.I2 = T.I2; f.Og = T.Og; f.O = T.O;
    f.up = T.up; f.down = T.down; f.pad = 0;
    const bool tile2d = f.kind == 1 && T.row_len >= 128 && T.Og >= T.O;
    const int64_t base = (int64_t)k.row0 * T.row_len + (tile2d ? f.strip * kC2Strip : 0);
    f.W = T.W + base; f.W_old = T.W_old + base;
    f.acc = a.acc + T.layer;
    f.s_tab = nullptr; f.r2_down = nullptr; f.S = f.b1 = f.bn_w = f.bn_b = nullptr;
    f.sinv_up = nullptr; f.r2_up = nullptr;
    if (T.down >= 0) {
        const DfqClePair p = a.pairs[T.down];
        const int off = a.chan_off[T.down] + k.row0;
        f.s_tab = a.r1 + off; f.r2_down = a.r2 + off;
        f.S = p.S + k.row0; f.b1 = p.b1 + k.row0;
        f.bn_w = p.bn_w ? p.bn_w + k.row0 : nullptr; f.bn_b = p.bn_b ? p.bn_b + k.row0 : nullptr;
    }
    if (T.up >= 0) {
        f.sinv_up = a.r1 + a.total_C + a.chan_off[T.up];
        f.r2_up = a.r2 + a.chan_off[T.up];
    }
    a.fat[t] = f;
}

// Cross_layer_equal.py:34-43 for one channel
__device__ __forceinline__ void cle2_scale(const Cle2Args &a, float r1, float r2, float &s, float &sinv) {
    float inv = __fdiv_rn(1.0f, __fadd_rn(r1, a.eps));
    float rt = __fsqrt_rn(__fadd_rn(__fmul_rn(r1, r2), a.eps));
    float st = __fmul_rn(inv, rt);
    if ((double)st < a.smax) {                 // Python min(): keeps the tensor when s < smax
        if ((double)st > a.smin) { s = st; sinv = __fdiv_rn(1.0f, st); }
        else { s = a.smin_f; sinv = a.inv_smin_f; }
    } else if (a.smax > a.smin) { s = a.smax_f; sinv = a.inv_smax_f; }   // also NaN
    else { s = a.smin_f; sinv = a.inv_smin_f; }
}

// What the row owner needs besides the row range, fetched while the row itself is still in flight.
struct Cle2RowPre { unsigned lo, hi; float S, b1, bw, bb; };

__device__ __forceinline__ Cle2RowPre cle2_prefetch_row(const Cle2Args &a, const Cle2Fat &k, int r, int par) {
    Cle2RowPre q;
    const unsigned *r2 = k.r2_down + (size_t)par * 2 * a.total_C;
    q.lo = r2[r]; q.hi = r2[a.total_C + r];
    q.S = k.S[r]; q.b1 = k.b1[r];
    q.bw = k.bn_w ? k.bn_w[r] : 0.0f; q.bb = k.bn_b ? k.bn_b[r] : 0.0f;
    return q;
}

// s and 1/s of row r from its range and the column range left by the previous iteration; the per-channel vectors
// follow (relation.py:26-30, Cross_layer_equal.py:48-55).  Once per channel and iteration.
__device__ __forceinline__ void cle2_finish_row(const Cle2Args &a, const Cle2Fat &k, int r, float r1,
                                                const Cle2RowPre &q) {
    const float lo2 = -dec_ordered(q.lo), hi2 = dec_ordered(q.hi);
    const float r2v = a.is_signed ? hi2 : __fsub_rn(hi2, lo2);
    float s, sinv;
    cle2_scale(a, r1, r2v, s, sinv);
    k.s_tab[r] = s;
    k.s_tab[a.total_C + r] = sinv;
    k.S[r] = __fmul_rn(q.S, s);
    k.b1[r] = __fmul_rn(q.b1, s);
    if (k.bn_w) k.bn_w[r] = __fmul_rn(q.bw, s);
    if (k.bn_b) k.bn_b[r] = __fmul_rn(q.bb, s);
}

__device__ __forceinline__ int cle2_cidx(const Cle2Fat &T, int o, int j) {
    const int i = (T.KK == 1) ? j : j / T.KK;
    return (T.Og >= T.O) ? i : (o / T.Og) * T.I2 + i;       // one group: the column is the input channel
}

// ---- "rows" pass: optional column scaling by 1/s of the up pair, row ranges, s of the down pair
__device__ void cle2_rows_pass(const Cle2Args &a, const Cle2Fat &k, int par) {
    const bool cs = k.up >= 0;
    const bool sgn = a.is_signed;
    const float *__restrict__ sinv_up = k.sinv_up;
    const int L = k.L;
    const bool one_group = k.Og >= k.O;
    if (L > 9) {                                             // warp per row
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        // lane i owns the epilogue (division, sqrt, vector updates) of this warp's i-th row, r = warp + 8 i
        const int my_r = warp + kWarps * lane;
        Cle2RowPre pre = {};
        if (my_r < k.nrows) pre = cle2_prefetch_row(a, k, my_r, par);
        float my_range = 0.0f;
        int cnt = 0;
        if (L <= 256) {                                      // 4 rows of the warp in flight together
            for (int rb = warp; rb < k.nrows; rb += 4 * kWarps) {
                float w[4][8], f[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int j = lane + 32 * u;
                    f[u] = (cs && j < L && one_group) ? __ldg(sinv_up + cle2_cidx(k, 0, j)) : 1.0f;
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int r = rb + q * kWarps;
                    const float *row = k.W + (int64_t)r * L;
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int j = lane + 32 * u;
                        w[q][u] = (r < k.nrows && j < L) ? row[j] : 0.0f;
                    }
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int r = rb + q * kWarps;
                    if (r >= k.nrows) break;                 // warp-uniform
                    float *row = k.W + (int64_t)r * L;
                    float lo = INFINITY, hi = -INFINITY;
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int j = lane + 32 * u;
                        if (j >= L) break;
                        float v = w[q][u];
                        if (cs) {
                            const float fj = one_group ? f[u] : __ldg(sinv_up + cle2_cidx(k, k.row0 + r, j));
                            v = __fmul_rn(v, fj); row[j] = v;
                        }
                        if (sgn) v = fabsf(v);
                        lo = fminf(lo, v); hi = fmaxf(hi, v);
                    }
                    lo = warp_min(lo); hi = warp_max(hi);
                    if (lane == cnt) my_range = sgn ? hi : __fsub_rn(hi, lo);
                    ++cnt;
                }
            }
        } else {
            for (int r = warp; r < k.nrows; r += kWarps) {
                float *__restrict__ row = k.W + (int64_t)r * L;
                float lo = INFINITY, hi = -INFINITY;
                for (int jb = lane; jb < L; jb += 32 * 16) {
                    float w[16], f[16];
#pragma unroll
                    for (int u = 0; u < 16; ++u) {
                        const int j = jb + 32 * u;
                        w[u] = (j < L) ? row[j] : 0.0f;
                        f[u] = (cs && j < L) ? __ldg(sinv_up + cle2_cidx(k, k.row0 + r, j)) : 1.0f;
                    }
#pragma unroll
                    for (int u = 0; u < 16; ++u) {
                        const int j = jb + 32 * u;
                        if (j >= L) break;
                        float v = w[u];
                        if (cs) { v = __fmul_rn(v, f[u]); row[j] = v; }
                        if (sgn) v = fabsf(v);
                        lo = fminf(lo, v); hi = fmaxf(hi, v);
                    }
                }
                lo = warp_min(lo); hi = warp_max(hi);
                if (lane == cnt) my_range = sgn ? hi : __fsub_rn(hi, lo);
                ++cnt;
            }
        }
        if (my_r < k.nrows) cle2_finish_row(a, k, my_r, my_range, pre);
    } else if (L == 9) {                                     // thread per row, depthwise 3x3: all 9 loads in flight
        for (int r = threadIdx.x; r < k.nrows; r += kThreads) {
            const int o = k.row0 + r;
            float *__restrict__ row = k.W + (int64_t)r * 9;
            float w[9];
#pragma unroll
            for (int j = 0; j < 9; ++j) w[j] = row[j];
            const float f = cs ? __ldg(sinv_up + cle2_cidx(k, o, 0)) : 1.0f;    // KK == 9: one channel per row
            const Cle2RowPre pre = cle2_prefetch_row(a, k, r, par);
            float lo = INFINITY, hi = -INFINITY;
#pragma unroll
            for (int j = 0; j < 9; ++j) {
                float v = w[j];
                if (cs) {
                    const float fj = (k.KK == 9) ? f : __ldg(sinv_up + cle2_cidx(k, o, j));
                    v = __fmul_rn(v, fj); row[j] = v;
                }
                if (sgn) v = fabsf(v);
                lo = fminf(lo, v); hi = fmaxf(hi, v);
            }
            cle2_finish_row(a, k, r, sgn ? hi : __fsub_rn(hi, lo), pre);
        }
    } else {                                                 // thread per row (rows shorter than 9)
        for (int r = threadIdx.x; r < k.nrows; r += kThreads) {
            const int o = k.row0 + r;
            float *row = k.W + (int64_t)r * L;
            const Cle2RowPre pre = cle2_prefetch_row(a, k, r, par);
            float lo = INFINITY, hi = -INFINITY;
            for (int j = 0; j < L; ++j) {
                float v = row[j];
                if (cs) { v = __fmul_rn(v, __ldg(sinv_up + cle2_cidx(k, o, j))); row[j] = v; }
                if (sgn) v = fabsf(v);
                lo = fminf(lo, v); hi = fmaxf(hi, v);
            }
            cle2_finish_row(a, k, r, sgn ? hi : __fsub_rn(hi, lo), pre);
        }
    }
}

// ---- "cols" pass: optional row scaling (s of the down pair) or column scaling (tail), metric, W_old, next column ranges.
// init: prologue -- no scaling, no metric, W_old <- W, column ranges into parity `par_next`.
__device__ void cle2_cols_pass(const Cle2Args &a, const Cle2Fat &k, int par, int par_next, bool init, double *smd) {
    const bool rs = !init && k.down >= 0;
    const bool cs = !init && k.down < 0 && k.up >= 0;
    const bool cr = k.up >= 0;                               // column ranges wanted (the tensor is some pair's W2)
    const bool sgn = a.is_signed;
    const int L = k.L;
    const float *__restrict__ s_rows = k.s_tab;
    const float *__restrict__ sinv_up = k.sinv_up;
    unsigned *r2lo = nullptr, *r2hi = nullptr;
    if (cr) {
        r2lo = k.r2_up + (size_t)par_next * 2 * a.total_C;
        r2hi = r2lo + a.total_C;
    }
    double part = 0.0;
    const bool one_group = k.Og >= k.O;
    constexpr int U = 8;                                     // loads in flight per thread before any store
    if (L >= 128 && one_group) {                             // 2-D tile: 32 rows x 64 columns, 8 rows per thread
        const int j = k.strip * kC2Strip + (threadIdx.x & (kC2Strip - 1));
        const int rg = threadIdx.x / kC2Strip;               // 4 row groups
        if (j < L) {
            const int c = (k.KK == 1) ? j : j / k.KK;
            float *__restrict__ wp = k.W + (threadIdx.x & (kC2Strip - 1));
            float *__restrict__ op = k.W_old + (threadIdx.x & (kC2Strip - 1));
            float w[U], wo[U], sr[U];
            const float cf = cs ? __ldg(sinv_up + c) : 1.0f;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int r = rg + 4 * u;
                const bool ok = r < k.nrows;
                w[u] = ok ? wp[(int64_t)r * L] : 0.0f;
                wo[u] = (ok && !init) ? op[(int64_t)r * L] : 0.0f;
                sr[u] = (ok && rs) ? s_rows[r] : 1.0f;
            }
            float lo = INFINITY, hi = -INFINITY;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int r = rg + 4 * u;
                if (r >= k.nrows) break;
                float v = w[u];
                if (cs) v = __fmul_rn(v, cf);
                if (rs) v = __fmul_rn(v, sr[u]);
                if (cs || rs) wp[(int64_t)r * L] = v;
                if (!init) part += (double)fabsf(__fsub_rn(v, wo[u]));
                op[(int64_t)r * L] = v;
                const float m = sgn ? fabsf(v) : v;
                lo = fminf(lo, m); hi = fmaxf(hi, m);
            }
            if (cr && hi >= lo) { atomicMax(r2lo + c, enc_ordered(-lo)); atomicMax(r2hi + c, enc_ordered(hi)); }
        }
    } else if (L == 9 && k.KK == 9 && k.Og == 1 && k.I2 == 1) {     // depthwise 3x3: the row is its own "column"
        for (int r = threadIdx.x; r < k.nrows; r += kThreads) {
            const int o = k.row0 + r;
            float *__restrict__ wp = k.W + (int64_t)r * 9;
            float *__restrict__ op = k.W_old + (int64_t)r * 9;
            float w[9], wo[9];
#pragma unroll
            for (int j = 0; j < 9; ++j) { w[j] = wp[j]; wo[j] = init ? 0.0f : op[j]; }
            float f = 1.0f;
            if (rs) f = s_rows[r];
            if (cs) f = __ldg(sinv_up + o);
            float lo = INFINITY, hi = -INFINITY;
#pragma unroll
            for (int j = 0; j < 9; ++j) {
                float v = w[j];
                if (cs || rs) { v = __fmul_rn(v, f); wp[j] = v; }
                if (!init) part += (double)fabsf(__fsub_rn(v, wo[j]));
                op[j] = v;
                const float m = sgn ? fabsf(v) : v;
                lo = fminf(lo, m); hi = fmaxf(hi, m);
            }
            if (cr) { r2lo[o] = enc_ordered(-lo); r2hi[o] = enc_ordered(hi); }     // sole owner of the channel
        }
    } else {                                                 // short rows or grouped: flat, per-element column atomics
        const int n = k.nrows * L;
        const float inv_L = 1.0f / (float)L;
        float *__restrict__ wp = k.W;
        float *__restrict__ op = k.W_old;
        for (int eb = threadIdx.x; eb < n; eb += U * kThreads) {
            float w[U], wo[U], fs[U];
            int cc[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int e = eb + u * kThreads;
                const bool ok = e < n;
                w[u] = ok ? wp[e] : 0.0f;
                wo[u] = (ok && !init) ? op[e] : 0.0f;
                const int r = (int)(((float)e + 0.5f) * inv_L), j = e - r * L;
                cc[u] = (ok && (cs || cr)) ? cle2_cidx(k, k.row0 + r, j) : 0;
                fs[u] = 1.0f;
                if (ok && rs) fs[u] = s_rows[r];
                if (ok && cs) fs[u] = __ldg(sinv_up + cc[u]);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int e = eb + u * kThreads;
                if (e >= n) break;
                float v = w[u];
                if (cs || rs) { v = __fmul_rn(v, fs[u]); wp[e] = v; }
                if (!init) part += (double)fabsf(__fsub_rn(v, wo[u]));
                op[e] = v;
                if (cr) {
                    const float m = sgn ? fabsf(v) : v;
                    atomicMax(r2lo + cc[u], enc_ordered(-m)); atomicMax(r2hi + cc[u], enc_ordered(m));
                }
            }
        }
    }
    if (!init) {                                             // block-uniform
        part = warp_sum(part);
        __syncthreads();
        if ((threadIdx.x & 31) == 0) smd[threadIdx.x >> 5] = part;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) t += smd[w];
            atomicAdd(k.acc + (size_t)par * a.n_tensors, t);
        }
    }
}

__global__ void __launch_bounds__(kThreads)
cle2_kernel(Cle2Args a) {
    cg::grid_group grid = cg::this_grid();
    __shared__ double smd[kWarps];
    __shared__ double st_diff;
    __shared__ int st_stall, st_iters;
    const int gtid = blockIdx.x * kThreads + threadIdx.x, gsize = gridDim.x * kThreads;

    for (int i = gtid; i < 4 * a.total_C; i += gsize) a.r2[i] = 0u;        // both parities
    for (int i = gtid; i < 2 * a.n_tensors; i += gsize) a.acc[i] = 0.0;
    grid.sync();
    for (int t = a.n_tasks + blockIdx.x; t < a.n_tasks + a.n_init; t += gridDim.x)   // old_graph = deepcopy, :83
        cle2_cols_pass(a, a.fat[t], 0, 0, true, smd);
    grid.sync();

    double diff = 1e8;                                // :67
    int stall = 0, iters = 0;                         // :68
    while (diff > a.threshold && stall < a.count && iters < a.max_iters) {     // :81
        const int par = iters & 1, par_next = par ^ 1;
        double *acc = a.acc + par * a.n_tensors;
        {   // the next iteration's column ranges start empty (last read one iteration ago)
            unsigned *z = a.r2 + (size_t)par_next * 2 * a.total_C;
            for (int i = gtid; i < 2 * a.total_C; i += gsize) z[i] = 0u;
        }
        for (int ph = 0; ph < a.n_phases; ++ph) {
            unsigned long long t0 = 0, t1 = 0;
            const bool stamp = a.dbg && iters == 2 && threadIdx.x == 0;
            if (stamp) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
            // (phase 0 holds no "cols" pass, so the zeroing above is ordered before every atomic by its barrier)
            unsigned long long worst = 0, worst_id = 0;
            for (int t = a.phase_begin[ph] + blockIdx.x; t < a.phase_begin[ph + 1]; t += gridDim.x) {
                const Cle2Fat k = a.fat[t];
                unsigned long long ta = 0, tb = 0;
                if (stamp) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(ta));
                if (k.kind == 0) cle2_rows_pass(a, k, par);
                else cle2_cols_pass(a, k, par, par_next, false, smd);
                if (a.dbg && iters == 2) {                   // block-uniform
                    __syncthreads();
                    if (stamp) {
                        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(tb));
                        if (tb - ta > worst) { worst = tb - ta; worst_id = t; }
                    }
                }
            }
            if (stamp) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
            grid.sync();
            if (stamp) {
                unsigned long long t2;
                asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t2));
                unsigned long long *d = a.dbg + ((size_t)blockIdx.x * a.n_phases + ph) * 5;
                d[0] = t0; d[1] = t1; d[2] = t2; d[3] = worst; d[4] = worst_id;
            }
        }
        if (threadIdx.x == 0) {
            // diff_tmp = np.sum([float(torch.mean(torch.abs(W - W_old))) ...]), :107-108
            double tot = 0.0;
            for (int l = 0; l < a.n_tensors; ++l)
                tot += (double)(float)(acc[l] / ((double)a.tensors[l].O * (double)a.tensors[l].row_len));
            if (fabs(diff - tot) > 1e-9) { stall = 0; diff = tot; } else { stall += 1; }   // :110-115
            iters += 1;
            st_diff = diff; st_stall = stall; st_iters = iters;
            if (blockIdx.x == 0) {
                a.res_d[iters] = tot;
                double *other = a.acc + (iters & 1) * a.n_tensors;    // next read after >= 1 barrier
                for (int l = 0; l < a.n_tensors; ++l) other[l] = 0.0;
            }
        }
        __syncthreads();
        diff = st_diff; stall = st_stall; iters = st_iters;
        __syncthreads();
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        a.res_i[0] = iters;
        a.res_i[1] = stall;
        a.res_d[0] = diff;
    }
}

}  // namespace dfq

using namespace dfq;

static const int kUnsupported = (int)cudaErrorNotSupported;

// Builds the tensor table, the per-phase task list (followed by the prologue's tasks) and the channel prefix.
// Returns cudaErrorNotSupported when the relations are not simple chains (a tensor that is W1 or W2 of two
// relations, shared BatchNorms, a row-before-column order, > 4096 channels): the caller then uses dfq_cle_run.
extern "C" int dfq_cle2_plan_host(const DfqClePair *pairs, int n_pairs, const DfqCleLayer *layers, int n_layers,
                                  DfqCleTensor *tensors, DfqCleTask *tasks, int max_tasks, int *n_tasks, int *n_init,
                                  int32_t *phase_begin, int *n_phases, int32_t *chan_off) {
    if (n_pairs <= 0 || n_layers <= 0) return (int)cudaErrorInvalidValue;
    chan_off[0] = 0;
    for (int i = 0; i < n_pairs; ++i) {
        chan_off[i + 1] = chan_off[i] + pairs[i].C1;
        for (int j = 0; j < i; ++j)
            if ((pairs[i].bn_w && pairs[i].bn_w == pairs[j].bn_w) || (pairs[i].bn_b && pairs[i].bn_b == pairs[j].bn_b) ||
                pairs[i].S == pairs[j].S || pairs[i].b1 == pairs[j].b1)
                return kUnsupported;
    }
    int phases = 0;
    int *phaseA = new int[n_layers];
    for (int l = 0; l < n_layers; ++l) {
        DfqCleTensor &T = tensors[l];
        T.W = layers[l].W; T.W_old = layers[l].W_old; T.up = -1; T.down = -1; T.layer = l;
        for (int i = 0; i < n_pairs; ++i) {
            if (pairs[i].W2 == T.W) { if (T.up >= 0) { delete[] phaseA; return kUnsupported; } T.up = i; }
            if (pairs[i].W1 == T.W) { if (T.down >= 0) { delete[] phaseA; return kUnsupported; } T.down = i; }
        }
        if (T.up < 0 && T.down < 0) { delete[] phaseA; return kUnsupported; }
        if (T.down >= 0) { T.O = pairs[T.down].C1; T.row_len = pairs[T.down].K1; }
        if (T.up >= 0) {
            const DfqClePair &u = pairs[T.up];
            const int G = (u.C1 != u.I2) ? u.C1 / u.I2 : 1;
            if (T.down >= 0 && (T.O != u.O2 || T.row_len != u.I2 * u.KK || pairs[T.up].stage >= pairs[T.down].stage)) {
                delete[] phaseA; return kUnsupported;
            }
            T.O = u.O2; T.row_len = u.I2 * u.KK; T.KK = u.KK; T.I2 = u.I2; T.Og = u.O2 / G;
        } else {
            T.KK = 1; T.I2 = T.row_len; T.Og = T.O;
        }
        if ((int64_t)T.O * T.row_len != layers[l].n) { delete[] phaseA; return kUnsupported; }
        phaseA[l] = T.down >= 0 ? pairs[T.down].stage : pairs[T.up].stage + 1;
        const int last = T.down >= 0 ? phaseA[l] + 1 : phaseA[l];
        if (last + 1 > phases) phases = last + 1;
    }
    if (phases > kC2MaxPhases) { delete[] phaseA; return kUnsupported; }
    auto col_rows = [](const DfqCleTensor &T) {
        if (T.row_len >= 128 && T.Og >= T.O) return kC2ColRows;     // 2-D tiles: 32 rows x 64-column strips
        int r = kC2Tile / T.row_len;
        if (r < 1) r = 1;
        return r > kC2Rows ? kC2Rows : r;
    };
    auto row_rows = [](const DfqCleTensor &T) {
        int r = kC2Tile / T.row_len;
        if (T.row_len > 9) { r = r / kWarps * kWarps; if (r < kWarps) r = kWarps; if (r > 256) r = 256; }
        return r < 1 ? 1 : r;
    };
    int nt = 0;
    bool overflow = false;
    auto emit = [&](int tensor, int kind, int rows_per) {
        const DfqCleTensor &T = tensors[tensor];
        const int strips = (kind == 1 && T.row_len >= 128 && T.Og >= T.O) ? (T.row_len + kC2Strip - 1) / kC2Strip : 1;
        for (int r0 = 0; r0 < T.O; r0 += rows_per)
            for (int st = 0; st < strips; ++st) {
                if (nt >= max_tasks) { overflow = true; return; }
                int nr = T.O - r0;
                tasks[nt].tensor = tensor; tasks[nt].row0 = r0; tasks[nt].nrows = nr < rows_per ? nr : rows_per;
                tasks[nt].kind = kind | (st << 8);
                ++nt;
            }
    };
    for (int ph = 0; ph < phases; ++ph) {
        phase_begin[ph] = nt;
        for (int l = 0; l < n_layers; ++l) {
            const DfqCleTensor &T = tensors[l];
            if (T.down >= 0 && phaseA[l] == ph) emit(l, 0, row_rows(T));
            if (T.down >= 0 && phaseA[l] + 1 == ph) emit(l, 1, col_rows(T));
            if (T.down < 0 && phaseA[l] == ph) emit(l, 1, col_rows(T));
        }
    }
    phase_begin[phases] = nt;
    // longest first inside a phase: blocks take tasks b, b + G, ..., so the long row passes start at once and the
    // short tiles fill in behind them
    auto cost = [&](const DfqCleTask &t) {
        const DfqCleTensor &T = tensors[t.tensor];
        if ((t.kind & 0xff) == 0) return 2 * (int64_t)t.nrows * T.row_len + 4096;
        const bool tile2d = T.row_len >= 128 && T.Og >= T.O;
        return tile2d ? (int64_t)t.nrows * kC2Strip : (int64_t)t.nrows * T.row_len;
    };
    for (int ph = 0; ph < phases; ++ph)
        std::stable_sort(tasks + phase_begin[ph], tasks + phase_begin[ph + 1],
                         [&](const DfqCleTask &x, const DfqCleTask &y) { return cost(x) > cost(y); });
    *n_tasks = nt;
    for (int l = 0; l < n_layers; ++l) emit(l, 1, col_rows(tensors[l]));
    *n_init = nt - *n_tasks;
    *n_phases = phases;
    delete[] phaseA;
    return overflow ? (int)cudaErrorInvalidValue : 0;
}

extern "C" int64_t dfq_cle2_fat_bytes(int n_tasks_total) { return (int64_t)n_tasks_total * (int64_t)sizeof(Cle2Fat); }

extern "C" int dfq_cle2_run(const DfqClePair *pairs_dev, int n_pairs, const DfqCleTensor *tensors_dev, int n_tensors,
                            const DfqCleTask *tasks_dev, int n_tasks, int n_init, const int32_t *phase_begin_host,
                            int n_phases, const int32_t *chan_off_dev, int total_C, float *r1_dev, unsigned *r2_dev,
                            double smin, double smax, int is_signed, float eps, double threshold, int count,
                            int max_iters, double *acc_dev, void *result_dev, int blocks_per_sm, void *fat_dev, void *debug_dev, void *stream) {
    if (n_phases > kC2MaxPhases || n_phases < 1) return (int)cudaErrorInvalidValue;
    Cle2Args a;
    a.pairs = pairs_dev; a.n_pairs = n_pairs; a.tensors = tensors_dev; a.n_tensors = n_tensors;
    a.tasks = tasks_dev; a.n_tasks = n_tasks; a.n_init = n_init; a.n_phases = n_phases;
    for (int i = 0; i <= n_phases; ++i) a.phase_begin[i] = phase_begin_host[i];
    a.chan_off = chan_off_dev; a.total_C = total_C; a.r1 = r1_dev; a.r2 = r2_dev;
    a.smin = smin; a.smax = smax;
    a.smin_f = (float)smin; a.smax_f = (float)smax;
    a.inv_smin_f = (float)(1.0 / smin); a.inv_smax_f = (float)(1.0 / smax);
    a.is_signed = is_signed; a.eps = eps; a.threshold = threshold; a.count = count; a.max_iters = max_iters;
    a.acc = acc_dev;
    a.res_i = (int *)result_dev;
    a.res_d = (double *)((char *)result_dev + 8);
    a.dbg = (unsigned long long *)debug_dev;
    a.fat = (Cle2Fat *)fat_dev;
    if (!fat_dev) return (int)cudaErrorInvalidValue;
    cle2_expand_kernel<<<(n_tasks + n_init + 127) / 128, 128, 0, (cudaStream_t)stream>>>(a);
    g_launches += 1;
    int occ = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, cle2_kernel, kThreads, 0);
    if (e != cudaSuccess) return (int)e;
    if (occ < 1) return (int)cudaErrorLaunchOutOfResources;
    if (blocks_per_sm < 1) blocks_per_sm = 2;
    const int per_sm = occ < blocks_per_sm ? occ : blocks_per_sm;
    void *args[] = {&a};
    e = cudaLaunchCooperativeKernel((void *)cle2_kernel, dim3(sm_count() * per_sm), dim3(kThreads), args, 0,
                                    (cudaStream_t)stream);
    g_launches += 1;
    return (int)e;
}
