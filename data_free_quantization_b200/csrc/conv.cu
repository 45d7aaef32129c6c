// Kernel (e), CUDA-core part: the quantised eval forward of utils/quantize.py:225-256 / 336-363 without tensor
// cores -- depthwise 3x3, small-Cin dense kxk (first layer), a reference fp32 pointwise/linear path, the
// quantised residual add and mean (utils/layer_transform.py:18-32, 79-94).  Every kernel
//   * fake-quantises its input ON LOAD with the consumer's QuantMeasure range (RangeSlot on the device), exactly
//     the fp32 recipe of utils/quantize.py:72-76, so no quantised copy of the activation is ever written;
//   * fuses bias (+ identity BatchNorm) + ReLU into the epilogue;
//   * optionally reduces the per-sample min/max of what it writes (the next QuantMeasure's update_stat
//     statistics, utils/quantize.py:110-111) into an encoded [N,2] table with one atomic pair per block.
// Accumulation is fp32 in a fixed order (not the reference's ATen order): parity is the 1e-3 logits tolerance.
#include "common.cuh"
#include <stdlib.h>

namespace dfq {

struct ConvArgs {
    const float *x; float *y; const float *w; const float *bias; const RangeSlot *slot; unsigned *stats;
    int N, C, H, W, O, OH, OW, stride, pad, dil, relu, bits, kh, kw;
    int cb, rows;          // depthwise: channels per block, output rows per block
    int strips;            // row strips per plane
};

// ---------------------------------------------------------------------------------------------------------
// Depthwise 3x3 (groups == C == O), stride 1 or 2, zero padding `pad`, dilation `dil`.
// A block owns `cb` consecutive channels x `rows` output rows of one image; the needed input rows are
// quantised once into shared memory with the zero padding materialised (real 0.0, utils/quantize.py quantises
// before F.conv2d pads), then every thread produces outputs from shared memory.
__global__ void __launch_bounds__(kThreads)
dwconv3x3_kernel(ConvArgs a) {
    extern __shared__ float smem[];
    __shared__ float red[2 * kWarps];
    const int tid = threadIdx.x;
    int b = blockIdx.x;
    const int strip = b % a.strips; b /= a.strips;
    const int cgroups = (a.C + a.cb - 1) / a.cb;
    const int cg = b % cgroups;
    const int n = b / cgroups;
    const int c0 = cg * a.cb, nc = min(a.cb, a.C - c0);
    const int r0 = strip * a.rows, nr = min(a.rows, a.OH - r0);
    const int in_r0 = r0 * a.stride - a.pad;                       // first input row (may be negative)
    const int in_rows = (nr - 1) * a.stride + 2 * a.dil + 1;
    const int pw = a.W + 2 * a.pad;                                 // padded row length in smem
    const bool on = a.slot != nullptr;
    QuantConsts qc;
    if (on) qc = consts_from_slot(*a.slot, a.bits, 0, 0.f, 0.f);
    // ---- stage: smem[c][r][pad + col]
    const int plane = in_rows * pw;
    for (int idx = tid; idx < nc * plane; idx += kThreads) {
        int c = idx / plane, rem = idx - c * plane;
        int r = rem / pw, col = rem - r * pw - a.pad;
        int gr = in_r0 + r;
        float v = 0.0f;
        if (gr >= 0 && gr < a.H && col >= 0 && col < a.W)
            v = quant_or_pass(__ldg(a.x + (((int64_t)n * a.C + c0 + c) * a.H + gr) * a.W + col), qc, on);
        smem[idx] = v;
    }
    __syncthreads();
    float lo = INFINITY, hi = -INFINITY;
    const int outs = nr * a.OW;
    for (int idx = tid; idx < nc * outs; idx += kThreads) {
        int c = idx / outs, rem = idx - c * outs;
        int orow = rem / a.OW, ocol = rem - orow * a.OW;
        const float *w = a.w + (int64_t)(c0 + c) * 9;
        const float *s = smem + c * plane + (orow * a.stride) * pw + ocol * a.stride;
        float acc = 0.0f;
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) acc = __fmaf_rn(s[i * a.dil * pw + j * a.dil], __ldg(w + i * 3 + j), acc);
        if (a.bias) acc = __fadd_rn(acc, __ldg(a.bias + c0 + c));
        if (a.relu) acc = fmaxf(acc, 0.0f);
        a.y[(((int64_t)n * a.O + c0 + c) * a.OH + r0 + orow) * a.OW + ocol] = acc;
        lo = fminf(lo, acc); hi = fmaxf(hi, acc);
    }
    if (a.stats) stats_flush(a.stats, n, lo, hi, red);
}

// ---------------------------------------------------------------------------------------------------------
// Depthwise 3x3 WITHOUT padding, stride S, dilation D, any plane size (odd widths, no alignment): DeepLab's backbone
// pads with F.pad first (modeling/segmentation/backbone/mobilenet.py:16-22, 62), so its 17 depthwise layers are
// valid convolutions of 259/131/67/35/37-wide planes.  A block owns `cb` channels x `rows` output rows; their input
// is one contiguous run per channel, copied flat (no index arithmetic per element), quantised once, into shared
// memory.  A thread then owns 4 vertically adjacent outputs of ONE column: lanes walk adjacent columns, so shared
// loads are conflict-free scalars (3 S + 2 D + 1 input rows x 3 taps = 18 loads for 36 FMAs at S = D = 1), global
// stores are coalesced whatever the row length, and the 9 weights come from shared memory as broadcasts.  Same
// accumulation order as dwconv3x3_kernel (bit-identical results).
//
// kWarp (small planes, a whole plane per channel): every WARP stages, synchronises (__syncwarp) and computes its own
// channel, so the warps of a block drift apart and one warp's loads overlap another's arithmetic -- the block form
// ran load phase / barrier / compute phase in lock-step on 33 x 33 planes (1.1-1.3 TB/s).
template <int S, int D, bool kWarp>
__global__ void __launch_bounds__(kThreads)
dwconv3x3_valid_kernel(ConvArgs a) {
    extern __shared__ float smem[];
    __shared__ float red[2 * kWarps];
    constexpr int R = 3 * S + 2 * D + 1;                          // input rows under 4 stacked outputs
    const int tid = threadIdx.x;
    int b = blockIdx.x;
    const int strip = b % a.strips; b /= a.strips;
    const int cgroups = (a.C + a.cb - 1) / a.cb;
    const int cg = b % cgroups;
    const int n = b / cgroups;
    const int c0 = cg * a.cb, nc = min(a.cb, a.C - c0);
    const int r0 = strip * a.rows, nr = min(a.rows, a.OH - r0);
    const int in_rows = (nr - 1) * S + 2 * D + 1;
    const int plane = in_rows * a.W;
    // allocation stride per channel: the run is stored at the 16-byte phase of its source, so that 128-bit global
    // loads pair with 128-bit shared stores
    const int plane_s = (((a.rows - 1) * S + 2 * D + 1) * a.W + 3 + 3) & ~3;
    float *wsm = smem + a.cb * plane_s;                           // [cb][10]: 9 taps + bias
    const bool on = a.slot != nullptr;
    QuantConsts qc;
    if (on) qc = consts_from_slot(*a.slot, a.bits, 0, 0.f, 0.f);
    const bool fast4 = on && qc.fast;
    // the threads that share a channel set: the block, or one warp (channel c0 + warp)
    const int gt = kWarp ? (tid & 31) : tid, gs = kWarp ? 32 : kThreads;
    const int cbeg = kWarp ? (tid >> 5) : 0, cend = kWarp ? min((tid >> 5) + 1, nc) : nc;
    for (int c = cbeg; c < cend; ++c) {
        const float *p = a.x + (((int64_t)n * a.C + c0 + c) * a.H + (int64_t)r0 * S) * a.W;
        const int m = (int)((((uintptr_t)p) >> 2) & 3);
        float *q = smem + c * plane_s + m;                        // q - m and p - m are 16-byte aligned
        const int j_end = m + plane;
        // groups of 4 floats in the shifted coordinate; two 128-bit loads in flight per thread
        for (int j0 = 4 * gt; j0 < j_end; j0 += 8 * gs) {
            const int j1 = j0 + 4 * gs;
            const bool full0 = j0 >= m && j0 + 4 <= j_end, full1 = j1 >= m && j1 + 4 <= j_end;
            float4 t0 = make_float4(0.f, 0.f, 0.f, 0.f), t1 = t0;
            if (full0) t0 = __ldg(reinterpret_cast<const float4 *>(p - m + j0));
            if (full1) t1 = __ldg(reinterpret_cast<const float4 *>(p - m + j1));
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int j = u ? j1 : j0;
                const bool full = u ? full1 : full0;
                const float4 t = u ? t1 : t0;
                if (full) {
                    float4 r;
                    if (fast4) r = quant_dequant4_fast(t, qc);
                    else r = make_float4(quant_or_pass(t.x, qc, on), quant_or_pass(t.y, qc, on),
                                         quant_or_pass(t.z, qc, on), quant_or_pass(t.w, qc, on));
                    *reinterpret_cast<float4 *>(q - m + j) = r;
                } else if (j < j_end) {
                    for (int k = max(j, m); k < min(j + 4, j_end); ++k) q[k - m] = quant_or_pass(__ldg(p - m + k), qc, on);
                }
            }
        }
    }
    for (int i = cbeg * 10 + gt; i < cend * 10; i += gs) {
        const int c = i / 10, t = i - c * 10;
        wsm[i] = t < 9 ? __ldg(a.w + (int64_t)(c0 + c) * 9 + t) : (a.bias ? __ldg(a.bias + c0 + c) : 0.0f);
    }
    if (kWarp) __syncwarp(); else __syncthreads();
    float lo = INFINITY, hi = -INFINITY;
    const int ngroups = (nr + 3) >> 2, per_c = ngroups * a.OW;
    const float inv_per_c = 1.0f / (float)per_c, inv_OW = 1.0f / (float)a.OW;
    for (int it = cbeg * per_c + gt; it < cend * per_c; it += gs) {
        int c = (int)(((float)it + 0.5f) * inv_per_c), rem = it - c * per_c;
        if (rem < 0) { --c; rem += per_c; } else if (rem >= per_c) { ++c; rem -= per_c; }
        int rg = (int)(((float)rem + 0.5f) * inv_OW), col = rem - rg * a.OW;
        if (col < 0) { --rg; col += a.OW; } else if (col >= a.OW) { ++rg; col -= a.OW; }
        const int orow = rg << 2, nv = min(4, nr - orow);
        const int mc = (int)(((uintptr_t)(a.x + (((int64_t)n * a.C + c0 + c) * a.H + (int64_t)r0 * S) * a.W) >> 2) & 3);
        const float *sp = smem + c * plane_s + mc + orow * S * a.W + col * S;
        const float *w = wsm + c * 10;
        float v[R][3];
        if (nv == 4) {
#pragma unroll
            for (int r = 0; r < R; ++r)
#pragma unroll
                for (int k = 0; k < 3; ++k) v[r][k] = sp[r * a.W + k * D];
        } else {
            const int last = (nv - 1) * S + 2 * D;                // last input row this item may touch
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float *row = sp + min(r, last) * a.W;
#pragma unroll
                for (int k = 0; k < 3; ++k) v[r][k] = row[k * D];
            }
        }
        float wr[9];
#pragma unroll
        for (int t = 0; t < 9; ++t) wr[t] = w[t];
        const float bias = w[9];
        float *yp = a.y + (((int64_t)n * a.O + c0 + c) * a.OH + r0 + orow) * a.OW + col;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float acc = 0.0f;
#pragma unroll
            for (int j = 0; j < 3; ++j)
#pragma unroll
                for (int k = 0; k < 3; ++k) acc = __fmaf_rn(v[i * S + j * D][k], wr[j * 3 + k], acc);
            if (a.bias) acc = __fadd_rn(acc, bias);
            if (a.relu) acc = fmaxf(acc, 0.0f);
            if (i < nv) {
                yp[(int64_t)i * a.OW] = acc;
                lo = fminf(lo, acc); hi = fmaxf(hi, acc);
            }
        }
    }
    if (a.stats) stats_flush(a.stats, n, lo, hi, red);
}

// ---------------------------------------------------------------------------------------------------------
// Depthwise 3x3, pad 1, dilation 1, stride 1 or 2 (every depthwise layer of MobileNetV2): the streaming form.
// A block owns either a strip of output rows of ONE channel plane or `cb` whole planes of consecutive channels, so
// its input is one contiguous run of floats: it is read with 128-bit loads, quantised once and written to a
// zero-initialised shared tile whose rows are [3 unused | left pad | W values | right pad ...] (the interior
// starts on a 16-byte boundary).  A thread then produces 4 adjacent outputs of one row from 3 x (1 scalar + 1 or
// 2 vector) shared loads with the 9 weights in registers, and stores them as one 128-bit word.
// Index decoding uses float reciprocals (exact here: every quotient is < 2^12 and the numerators are offset by
// one half).
struct DwArgs {
    const float *x; float *y; const float *w; const float *bias; const RangeSlot *slot; unsigned *stats;
    int C, H, W, OH, OW, stride, relu, bits;
    int cb, rows, strips, RS, G4;
    float inv_W, inv_W4, inv_G4;
};

__device__ __forceinline__ int fdiv_small(int x, float inv) { return (int)(((float)x + 0.5f) * inv); }

template <bool kVec>
__global__ void __launch_bounds__(kThreads)
dwconv3x3_stream_kernel(DwArgs a) {
    extern __shared__ __align__(16) float smem[];
    __shared__ float red[2 * kWarps];
    const int tid = threadIdx.x;
    int b = blockIdx.x;
    const int strip = b % a.strips; b /= a.strips;
    const int cgroups = (a.C + a.cb - 1) / a.cb;
    const int cg = b % cgroups;
    const int n = b / cgroups;
    const int c0 = cg * a.cb, nc = min(a.cb, a.C - c0);
    const int r0 = strip * a.rows, nr = min(a.rows, a.OH - r0);
    const int in_r0 = r0 * a.stride - 1;
    const int in_rows = (nr - 1) * a.stride + 3;
    const int gr0 = max(in_r0, 0), gr1 = min(in_r0 + in_rows, a.H);     // rows that exist
    const int plane = in_rows * a.RS;
    const bool on = a.slot != nullptr;
    QuantConsts qc;
    if (on) qc = consts_from_slot(*a.slot, a.bits, 0, 0.f, 0.f);
    // ---- zero the tile (pads, halo rows beyond the image, slack)
    const int tile4 = (nc * plane + 4 + 3) >> 2;
    for (int i = tid; i < tile4; i += kThreads) reinterpret_cast<float4 *>(smem)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
    // ---- stage: one contiguous run of nc * Lc floats
    const int Lc = (gr1 - gr0) * a.W;
    const float *src = a.x + (((int64_t)n * a.C + c0) * a.H + gr0) * a.W;
    const int row_shift = gr0 - in_r0;
    if (kVec) {                                   // W % 4 == 0: a 128-bit word never leaves its row
        const int W4 = a.W >> 2, Lc4 = Lc >> 2;
        const float inv_Lc4 = 1.0f / (float)Lc4;
        const float4 *src4 = reinterpret_cast<const float4 *>(src);
        for (int i = tid; i < nc * Lc4; i += kThreads) {
            float4 v = __ldg(src4 + i);
            int c = 0, rem = i;
            if (nc > 1) { c = fdiv_small(i, inv_Lc4); rem = i - c * Lc4; }
            const int rr = fdiv_small(rem, a.inv_W4), c4 = rem - rr * W4;
            float4 o;
            o.x = quant_or_pass(v.x, qc, on); o.y = quant_or_pass(v.y, qc, on);
            o.z = quant_or_pass(v.z, qc, on); o.w = quant_or_pass(v.w, qc, on);
            *reinterpret_cast<float4 *>(smem + c * plane + (rr + row_shift) * a.RS + 4 + 4 * c4) = o;
        }
    } else {
        const float inv_Lc = 1.0f / (float)Lc;
        for (int i = tid; i < nc * Lc; i += kThreads) {
            float v = __ldg(src + i);
            int c = 0, rem = i;
            if (nc > 1) { c = fdiv_small(i, inv_Lc); rem = i - c * Lc; }
            const int rr = fdiv_small(rem, a.inv_W), col = rem - rr * a.W;
            smem[c * plane + (rr + row_shift) * a.RS + 4 + col] = quant_or_pass(v, qc, on);
        }
    }
    __syncthreads();
    // ---- compute: item = (channel, output row, group of 4 output columns)
    float lo = INFINITY, hi = -INFINITY;
    const int per_c = nr * a.G4;
    const float inv_per_c = 1.0f / (float)per_c;
    float wreg[9], bv = 0.0f;
#pragma unroll
    for (int t = 0; t < 9; ++t) wreg[t] = __ldg(a.w + (int64_t)c0 * 9 + t);
    if (a.bias) bv = __ldg(a.bias + c0);
    for (int it = tid; it < nc * per_c; it += kThreads) {
        int c = 0, rem = it;
        if (nc > 1) { c = fdiv_small(it, inv_per_c); rem = it - c * per_c; }
        const int orow = fdiv_small(rem, a.inv_G4), ox0 = (rem - orow * a.G4) * 4;
        if (nc > 1) {                                         // one channel per block keeps its 9 weights in registers
            const float *wp = a.w + (int64_t)(c0 + c) * 9;
#pragma unroll
            for (int t = 0; t < 9; ++t) wreg[t] = __ldg(wp + t);
            bv = a.bias ? __ldg(a.bias + c0 + c) : 0.0f;
        }
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        const float *srow = smem + c * plane + (orow * a.stride) * a.RS;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const float w0 = wreg[3 * i], w1 = wreg[3 * i + 1], w2 = wreg[3 * i + 2];
            const float *s = srow + i * a.RS;
            if (a.stride == 1) {
                const float L = s[ox0 + 3], R = s[ox0 + 8];
                const float4 M = *reinterpret_cast<const float4 *>(s + ox0 + 4);
                acc[0] = __fmaf_rn(M.y, w2, __fmaf_rn(M.x, w1, __fmaf_rn(L, w0, acc[0])));
                acc[1] = __fmaf_rn(M.z, w2, __fmaf_rn(M.y, w1, __fmaf_rn(M.x, w0, acc[1])));
                acc[2] = __fmaf_rn(M.w, w2, __fmaf_rn(M.z, w1, __fmaf_rn(M.y, w0, acc[2])));
                acc[3] = __fmaf_rn(R, w2, __fmaf_rn(M.w, w1, __fmaf_rn(M.z, w0, acc[3])));
            } else {
                const float L = s[2 * ox0 + 3];
                const float4 M0 = *reinterpret_cast<const float4 *>(s + 2 * ox0 + 4);
                const float4 M1 = *reinterpret_cast<const float4 *>(s + 2 * ox0 + 8);
                acc[0] = __fmaf_rn(M0.y, w2, __fmaf_rn(M0.x, w1, __fmaf_rn(L, w0, acc[0])));
                acc[1] = __fmaf_rn(M0.w, w2, __fmaf_rn(M0.z, w1, __fmaf_rn(M0.y, w0, acc[1])));
                acc[2] = __fmaf_rn(M1.y, w2, __fmaf_rn(M1.x, w1, __fmaf_rn(M0.w, w0, acc[2])));
                acc[3] = __fmaf_rn(M1.w, w2, __fmaf_rn(M1.z, w1, __fmaf_rn(M1.y, w0, acc[3])));
            }
        }
        float *yp = a.y + (((int64_t)n * a.C + c0 + c) * a.OH + r0 + orow) * a.OW + ox0;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            if (a.bias) acc[t] = __fadd_rn(acc[t], bv);
            if (a.relu) acc[t] = fmaxf(acc[t], 0.0f);
        }
        if ((a.OW & 3) == 0) {
            *reinterpret_cast<float4 *>(yp) = make_float4(acc[0], acc[1], acc[2], acc[3]);
            lo = fminf(lo, fminf(fminf(acc[0], acc[1]), fminf(acc[2], acc[3])));
            hi = fmaxf(hi, fmaxf(fmaxf(acc[0], acc[1]), fmaxf(acc[2], acc[3])));
        } else {
#pragma unroll
            for (int t = 0; t < 4; ++t)
                if (ox0 + t < a.OW) { yp[t] = acc[t]; lo = fminf(lo, acc[t]); hi = fmaxf(hi, acc[t]); }
        }
    }
    if (a.stats) stats_flush(a.stats, n, lo, hi, red);
}

// ---------------------------------------------------------------------------------------------------------
// Depthwise 3x3, pad 1, dilation 1, stride 1 or 2, rows of 4*L floats (L <= 32): the register form.  No shared
// tile, no block barrier before the epilogue.  `lanes` = W/4 adjacent lanes own one input row as 128-bit words; a
// warp carries G = 32/lanes such row groups, each walking `R` output rows of its own (channel, row block) of the
// same image with a three-row window of quantised values in registers.  The left / right neighbours of a lane's
// four columns come from the adjacent lanes by shuffle (exact 0 at the plane's edges, like the rows above and
// below the image: the reference quantises before F.conv2d pads).  Input rows are requested two (stride 1) or
// four (stride 2) rows ahead of their use, so the loads of later rows are in flight while the current row is
// computed.  Tap order per output is that of the shared-memory form (bit-identical results).
struct DwRowArgs {
    const float *x; float *y; const float *w; const float *bias; const RangeSlot *slot; unsigned *stats;
    int C, H, W, OH, OW, relu, bits;
    int lanes, G, R, nblk, items, wps;   // lanes per row, row groups per warp, output rows per item, row blocks
                                         // per plane, items per image (C * nblk), warps per image
};

struct DwRow { float4 m; float l, r; };

// one input row's three taps into 4 adjacent outputs (stride 1) / 2 outputs (stride 2); rows are passed by value so
// that they stay in registers
__device__ __forceinline__ void taps4(float (&acc)[4], const DwRow rw, float w0, float w1, float w2) {
    acc[0] = __fmaf_rn(rw.m.y, w2, __fmaf_rn(rw.m.x, w1, __fmaf_rn(rw.l, w0, acc[0])));
    acc[1] = __fmaf_rn(rw.m.z, w2, __fmaf_rn(rw.m.y, w1, __fmaf_rn(rw.m.x, w0, acc[1])));
    acc[2] = __fmaf_rn(rw.m.w, w2, __fmaf_rn(rw.m.z, w1, __fmaf_rn(rw.m.y, w0, acc[2])));
    acc[3] = __fmaf_rn(rw.r, w2, __fmaf_rn(rw.m.w, w1, __fmaf_rn(rw.m.z, w0, acc[3])));
}
__device__ __forceinline__ void taps2(float (&acc)[2], const DwRow rw, float w0, float w1, float w2) {
    acc[0] = __fmaf_rn(rw.m.y, w2, __fmaf_rn(rw.m.x, w1, __fmaf_rn(rw.l, w0, acc[0])));
    acc[1] = __fmaf_rn(rw.m.w, w2, __fmaf_rn(rw.m.z, w1, __fmaf_rn(rw.m.y, w0, acc[1])));
}

template <int S>
__global__ void __launch_bounds__(kThreads, 4)
dwconv3x3_rows_kernel(DwRowArgs a) {
    __shared__ float red[2 * kWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = blockIdx.y;
    const int wi = blockIdx.x * kWarps + warp;
    const int g = lane / a.lanes, j = lane - g * a.lanes;
    const int item = wi * a.G + g;
    const bool active = g < a.G && wi < a.wps && item < a.items;
    int c = 0, blk = 0;
    if (active) { c = item / a.nblk; blk = item - c * a.nblk; }
    const int r0 = blk * a.R;
    const int nr = active ? min(a.R, a.OH - r0) : 0;
    const bool on = a.slot != nullptr;
    QuantConsts qc;
    if (on) qc = consts_from_slot(*a.slot, a.bits, 0, 0.f, 0.f);
    float wreg[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) wreg[t] = __ldg(a.w + (int64_t)c * 9 + t);
    const float bv = a.bias ? __ldg(a.bias + c) : 0.0f;
    float *yp = a.y + (((int64_t)n * a.C + c) * a.OH + r0) * a.OW + (S == 1 ? 4 : 2) * j;
    const bool first = j == 0, last = j == a.lanes - 1;
    const int ib = S * r0 - 1;                                   // first input row of the window
    // input rows this item really reads: [ir_lo, ir_lo + span]; everything else is zero padding (an inactive
    // lane gets an empty range).  One unsigned compare per row.
    const int ir_lo = max(ib, 0);
    const unsigned span = active ? (unsigned)(min(S * (r0 + nr - 1) + 1, a.H - 1) - ir_lo) : 0u;
    const int ir_base = active ? ir_lo : (1 << 30);
    auto valid = [&](int ir) { return (unsigned)(ir - ir_base) <= span; };
    const float4 *xrow = reinterpret_cast<const float4 *>(a.x + ((int64_t)n * a.C + c) * a.H * a.W + 4 * j);
    const int w4 = a.W >> 2;                                      // row pitch in 128-bit words
    auto ld = [&](int ir) {
        return valid(ir) ? __ldg(xrow + (int64_t)ir * w4) : make_float4(0.f, 0.f, 0.f, 0.f);
    };
    const int qmode = !on ? 0 : (qc.fast ? 1 : 2);               // warp-uniform
    auto qrow = [&](float4 v, int ir) {
        DwRow r;
        if (qmode == 1) {
            if (valid(ir)) v = quant_dequant4_fast(v, qc);
        } else if (qmode == 2) {
            if (valid(ir)) {
                v.x = quant_dequant(v.x, qc); v.y = quant_dequant(v.y, qc);
                v.z = quant_dequant(v.z, qc); v.w = quant_dequant(v.w, qc);
            }
        }
        r.m = v;
        const float l = __shfl_up_sync(0xffffffffu, v.w, 1), rr = __shfl_down_sync(0xffffffffu, v.x, 1);
        r.l = first ? 0.0f : l;
        r.r = last ? 0.0f : rr;
        return r;
    };
    float lo = INFINITY, hi = -INFINITY;
    if (S == 1) {
        DwRow top = qrow(ld(ib), ib), mid = qrow(ld(ib + 1), ib + 1);
        float4 p1 = ld(ib + 2), p2 = ld(ib + 3);
        for (int t = 0; t < a.R; ++t) {
            const float4 p3 = ld(ib + t + 4);
            const DwRow bot = qrow(p1, ib + t + 2);
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            taps4(acc, top, wreg[0], wreg[1], wreg[2]);
            taps4(acc, mid, wreg[3], wreg[4], wreg[5]);
            taps4(acc, bot, wreg[6], wreg[7], wreg[8]);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (a.bias) acc[u] = __fadd_rn(acc[u], bv);
                if (a.relu) acc[u] = fmaxf(acc[u], 0.0f);
            }
            if (t < nr) {
                *reinterpret_cast<float4 *>(yp + (int64_t)t * a.OW) = make_float4(acc[0], acc[1], acc[2], acc[3]);
                lo = fminf(lo, fminf(fminf(acc[0], acc[1]), fminf(acc[2], acc[3])));
                hi = fmaxf(hi, fmaxf(fmaxf(acc[0], acc[1]), fmaxf(acc[2], acc[3])));
            }
            top = mid; mid = bot; p1 = p2; p2 = p3;
        }
    } else {
        DwRow top = qrow(ld(ib), ib);
        float4 p1 = ld(ib + 1), p2 = ld(ib + 2), p3 = ld(ib + 3), p4 = ld(ib + 4);
        for (int t = 0; t < a.R; ++t) {
            const DwRow mid = qrow(p1, ib + 2 * t + 1), bot = qrow(p2, ib + 2 * t + 2);
            p1 = p3; p2 = p4;
            p3 = ld(ib + 2 * t + 5); p4 = ld(ib + 2 * t + 6);
            float acc[2] = {0.f, 0.f};
            taps2(acc, top, wreg[0], wreg[1], wreg[2]);
            taps2(acc, mid, wreg[3], wreg[4], wreg[5]);
            taps2(acc, bot, wreg[6], wreg[7], wreg[8]);
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                if (a.bias) acc[u] = __fadd_rn(acc[u], bv);
                if (a.relu) acc[u] = fmaxf(acc[u], 0.0f);
            }
            if (t < nr) {
                *reinterpret_cast<float2 *>(yp + (int64_t)t * a.OW) = make_float2(acc[0], acc[1]);
                lo = fminf(lo, fminf(acc[0], acc[1]));
                hi = fmaxf(hi, fmaxf(acc[0], acc[1]));
            }
            top = bot;
        }
    }
    if (a.stats) stats_flush(a.stats, n, lo, hi, red);
}

// The same register form for narrow planes (W <= 32, any W: 14 x 14 and 7 x 7 in MobileNetV2): one lane per input
// column, G = 32/W whole planes side by side in a warp, K planes one after the other per lane group.  With
// stride 2 output column j gathers input columns 2j-1, 2j, 2j+1 from the lanes that hold them.
struct DwColArgs {
    const float *x; float *y; const float *w; const float *bias; const RangeSlot *slot; unsigned *stats;
    int C, H, W, OH, OW, relu, bits;
    int G, K, wps;                       // planes side by side, planes in sequence, warps per image
};

constexpr int kColsPre = 16;           // planes of at most this many rows are loaded whole before use
template <int S>
__global__ void __launch_bounds__(kThreads, 3)
dwconv3x3_cols_kernel(DwColArgs a) {
    __shared__ float red[2 * kWarps];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = blockIdx.y;
    const int wi = blockIdx.x * kWarps + warp;
    const int g = lane / a.W, j = lane - g * a.W;
    const bool on = a.slot != nullptr;
    QuantConsts qc;
    if (on) qc = consts_from_slot(*a.slot, a.bits, 0, 0.f, 0.f);
    const bool first = j == 0, last = j == a.W - 1;
    // stride 2: lanes of the three taps of output column j (clamped; masked below)
    const int base = g * a.W;
    const int sl = min(base + max(2 * j - 1, 0), 31), sm = min(base + 2 * j, 31), sr = min(base + 2 * j + 1, 31);
    const bool okl = 2 * j - 1 >= 0, okr = 2 * j + 1 < a.W, outlane = j < a.OW;
    const int ir_last = S * (a.OH - 1) + 1;
    float lo = INFINITY, hi = -INFINITY;
    for (int k = 0; k < a.K; ++k) {
        const int c = (wi * a.K + k) * a.G + g;
        const bool active = g < a.G && wi < a.wps && c < a.C;
        const int cc = active ? c : 0;
        float wreg[9];
#pragma unroll
        for (int t = 0; t < 9; ++t) wreg[t] = __ldg(a.w + (int64_t)cc * 9 + t);
        const float bv = a.bias ? __ldg(a.bias + cc) : 0.0f;
        const float *xp = a.x + ((int64_t)n * a.C + cc) * a.H * a.W + j;
        float *yp = a.y + ((int64_t)n * a.C + cc) * a.OH * a.OW + j;
        auto valid = [&](int ir) { return active && ir >= 0 && ir < a.H && ir <= ir_last; };
        auto ld = [&](int ir) { return valid(ir) ? __ldg(xp + (int64_t)ir * a.W) : 0.0f; };
        auto q = [&](float v, int ir) { return (on && valid(ir)) ? quant_dequant(v, qc) : v; };
        if (S == 1 && a.H <= kColsPre) {
            // short planes (14 x 14, 7 x 7): ALL rows of the plane are requested before the first is used -- with two
            // rows of look-ahead a warp had ~60 instructions to cover a memory round trip
            float pre[kColsPre + 1];
#pragma unroll
            for (int r = 0; r <= kColsPre; ++r) pre[r] = r < a.H ? ld(r) : 0.0f;
            float t0 = 0.0f, t1 = q(pre[0], 0);                     // row -1 is padding
            float tl = 0.0f, tr = 0.0f;
            float ml = __shfl_up_sync(0xffffffffu, t1, 1), mr = __shfl_down_sync(0xffffffffu, t1, 1);
            if (first) ml = 0.f;
            if (last) mr = 0.f;
#pragma unroll
            for (int t = 0; t < kColsPre; ++t) {
                if (t >= a.OH) break;
                const float b1 = q(pre[t + 1], t + 1);
                float bl = __shfl_up_sync(0xffffffffu, b1, 1), br = __shfl_down_sync(0xffffffffu, b1, 1);
                if (first) bl = 0.f;
                if (last) br = 0.f;
                float acc = __fmaf_rn(tr, wreg[2], __fmaf_rn(t0, wreg[1], __fmaf_rn(tl, wreg[0], 0.0f)));
                acc = __fmaf_rn(mr, wreg[5], __fmaf_rn(t1, wreg[4], __fmaf_rn(ml, wreg[3], acc)));
                acc = __fmaf_rn(br, wreg[8], __fmaf_rn(b1, wreg[7], __fmaf_rn(bl, wreg[6], acc)));
                if (a.bias) acc = __fadd_rn(acc, bv);
                if (a.relu) acc = fmaxf(acc, 0.0f);
                if (active) {
                    yp[(int64_t)t * a.OW] = acc;
                    lo = fminf(lo, acc); hi = fmaxf(hi, acc);
                }
                t0 = t1; tl = ml; tr = mr; t1 = b1; ml = bl; mr = br;
            }
        } else if (S == 1) {
            float t0 = q(ld(-1), -1), t1 = q(ld(0), 0);
            float tl = __shfl_up_sync(0xffffffffu, t0, 1), tr = __shfl_down_sync(0xffffffffu, t0, 1);
            float ml = __shfl_up_sync(0xffffffffu, t1, 1), mr = __shfl_down_sync(0xffffffffu, t1, 1);
            if (first) { tl = 0.f; ml = 0.f; }
            if (last) { tr = 0.f; mr = 0.f; }
            float p1 = ld(1), p2 = ld(2);
            for (int t = 0; t < a.OH; ++t) {
                const float p3 = ld(t + 3);
                const float b1 = q(p1, t + 1);
                float bl = __shfl_up_sync(0xffffffffu, b1, 1), br = __shfl_down_sync(0xffffffffu, b1, 1);
                if (first) bl = 0.f;
                if (last) br = 0.f;
                float acc = __fmaf_rn(tr, wreg[2], __fmaf_rn(t0, wreg[1], __fmaf_rn(tl, wreg[0], 0.0f)));
                acc = __fmaf_rn(mr, wreg[5], __fmaf_rn(t1, wreg[4], __fmaf_rn(ml, wreg[3], acc)));
                acc = __fmaf_rn(br, wreg[8], __fmaf_rn(b1, wreg[7], __fmaf_rn(bl, wreg[6], acc)));
                if (a.bias) acc = __fadd_rn(acc, bv);
                if (a.relu) acc = fmaxf(acc, 0.0f);
                if (active) {
                    yp[(int64_t)t * a.OW] = acc;
                    lo = fminf(lo, acc); hi = fmaxf(hi, acc);
                }
                t0 = t1; tl = ml; tr = mr; t1 = b1; ml = bl; mr = br; p1 = p2; p2 = p3;
            }
        } else {
            auto taps = [&](float v, float &l, float &m, float &r) {
                l = __shfl_sync(0xffffffffu, v, sl); m = __shfl_sync(0xffffffffu, v, sm); r = __shfl_sync(0xffffffffu, v, sr);
                if (!okl) l = 0.f;
                if (!okr) r = 0.f;
            };
            float tl, tm, tr;
            if (a.H <= kColsPre) {                                  // as above: the whole plane in flight
                float pre[kColsPre];
#pragma unroll
                for (int r = 0; r < kColsPre; ++r) pre[r] = r < a.H ? ld(r) : 0.0f;
                tl = tm = tr = 0.0f;                                // row -1 is padding
#pragma unroll
                for (int t = 0; t < kColsPre / 2; ++t) {
                    if (t >= a.OH) break;
                    float ml, mm, mr, bl, bm, br;
                    taps(q(pre[2 * t], 2 * t), ml, mm, mr);
                    taps(q(pre[2 * t + 1], 2 * t + 1), bl, bm, br);
                    float acc = __fmaf_rn(tr, wreg[2], __fmaf_rn(tm, wreg[1], __fmaf_rn(tl, wreg[0], 0.0f)));
                    acc = __fmaf_rn(mr, wreg[5], __fmaf_rn(mm, wreg[4], __fmaf_rn(ml, wreg[3], acc)));
                    acc = __fmaf_rn(br, wreg[8], __fmaf_rn(bm, wreg[7], __fmaf_rn(bl, wreg[6], acc)));
                    if (a.bias) acc = __fadd_rn(acc, bv);
                    if (a.relu) acc = fmaxf(acc, 0.0f);
                    if (active && outlane) {
                        yp[(int64_t)t * a.OW] = acc;
                        lo = fminf(lo, acc); hi = fmaxf(hi, acc);
                    }
                    tl = bl; tm = bm; tr = br;
                }
                continue;
            }
            taps(q(ld(-1), -1), tl, tm, tr);
            float p1 = ld(0), p2 = ld(1), p3 = ld(2), p4 = ld(3);
            for (int t = 0; t < a.OH; ++t) {
                float ml, mm, mr, bl, bm, br;
                taps(q(p1, 2 * t), ml, mm, mr);
                taps(q(p2, 2 * t + 1), bl, bm, br);
                p1 = p3; p2 = p4;
                p3 = ld(2 * t + 4); p4 = ld(2 * t + 5);
                float acc = __fmaf_rn(tr, wreg[2], __fmaf_rn(tm, wreg[1], __fmaf_rn(tl, wreg[0], 0.0f)));
                acc = __fmaf_rn(mr, wreg[5], __fmaf_rn(mm, wreg[4], __fmaf_rn(ml, wreg[3], acc)));
                acc = __fmaf_rn(br, wreg[8], __fmaf_rn(bm, wreg[7], __fmaf_rn(bl, wreg[6], acc)));
                if (a.bias) acc = __fadd_rn(acc, bv);
                if (a.relu) acc = fmaxf(acc, 0.0f);
                if (active && outlane) {
                    yp[(int64_t)t * a.OW] = acc;
                    lo = fminf(lo, acc); hi = fmaxf(hi, acc);
                }
                tl = bl; tm = bm; tr = br;
            }
        }
    }
    if (a.stats) stats_flush(a.stats, n, lo, hi, red);
}

// ---------------------------------------------------------------------------------------------------------
// Dense kxk convolution, direct form, for small C*kh*kw (MobileNetV2's first layer: 3 -> 32, 3x3, stride 2).
// A block owns `rows` output rows of one image for all O output channels: the quantised, zero-padded input strip
// and the whole weight tensor sit in shared memory; a thread owns one output pixel and walks the output channels
// 8 at a time (stores for a fixed channel are contiguous across the warp).
constexpr int kOT = 8;

__global__ void __launch_bounds__(kThreads)
conv_direct_kernel(ConvArgs a) {
    extern __shared__ float smem[];
    __shared__ float red[2 * kWarps];
    const int tid = threadIdx.x;
    const int strip = blockIdx.x % a.strips, n = blockIdx.x / a.strips;
    const int r0 = strip * a.rows, nr = min(a.rows, a.OH - r0);
    const int in_r0 = r0 * a.stride - a.pad;
    const int in_rows = (nr - 1) * a.stride + (a.kh - 1) * a.dil + 1;
    const int pw = a.W + 2 * a.pad;
    const int plane = in_rows * pw;
    const int kk = a.kh * a.kw, ckk = a.C * kk;
    float *sw = smem;                                   // [O][C*kh*kw]
    float *sx = smem + a.O * ckk;                       // [C][in_rows][pw]
    const bool on = a.slot != nullptr;
    QuantConsts qc;
    if (on) qc = consts_from_slot(*a.slot, a.bits, 0, 0.f, 0.f);
    for (int idx = tid; idx < a.O * ckk; idx += kThreads) sw[idx] = __ldg(a.w + idx);
    for (int idx = tid; idx < a.C * plane; idx += kThreads) {
        int c = idx / plane, rem = idx - c * plane;
        int r = rem / pw, col = rem - r * pw - a.pad;
        int gr = in_r0 + r;
        float v = 0.0f;
        if (gr >= 0 && gr < a.H && col >= 0 && col < a.W)
            v = quant_or_pass(__ldg(a.x + (((int64_t)n * a.C + c) * a.H + gr) * a.W + col), qc, on);
        sx[idx] = v;
    }
    __syncthreads();
    float lo = INFINITY, hi = -INFINITY;
    const int outs = nr * a.OW;
    for (int p = tid; p < outs; p += kThreads) {
        int orow = p / a.OW, ocol = p - orow * a.OW;
        const float *s0 = sx + (orow * a.stride) * pw + ocol * a.stride;
        for (int o0 = 0; o0 < a.O; o0 += kOT) {
            float acc[kOT];
#pragma unroll
            for (int t = 0; t < kOT; ++t) acc[t] = 0.0f;
            for (int c = 0; c < a.C; ++c)
                for (int i = 0; i < a.kh; ++i)
                    for (int j = 0; j < a.kw; ++j) {
                        float v = s0[c * plane + i * a.dil * pw + j * a.dil];
                        const float *wp = sw + (int64_t)o0 * ckk + c * kk + i * a.kw + j;
#pragma unroll
                        for (int t = 0; t < kOT; ++t)
                            if (o0 + t < a.O) acc[t] = __fmaf_rn(v, wp[t * ckk], acc[t]);
                    }
#pragma unroll
            for (int t = 0; t < kOT; ++t) {
                if (o0 + t >= a.O) break;
                float r = acc[t];
                if (a.bias) r = __fadd_rn(r, __ldg(a.bias + o0 + t));
                if (a.relu) r = fmaxf(r, 0.0f);
                a.y[(((int64_t)n * a.O + o0 + t) * a.OH + r0 + orow) * a.OW + ocol] = r;
                lo = fminf(lo, r); hi = fmaxf(hi, r);
            }
        }
    }
    if (a.stats) stats_flush(a.stats, n, lo, hi, red);
}

// Register-blocked form of the direct convolution for tiny C*kh*kw (the 3 -> 32, 3x3, stride 2 stem): the thread's
// C*KH*KW quantised inputs live in registers, the weights are read as [tap][O] rows with 128-bit broadcast loads, 16
// output channels accumulate at a time.  Same tap order per output as conv_direct_kernel.
template <int C, int KH, int KW>
__global__ void __launch_bounds__(kThreads)
conv_direct_reg_kernel(ConvArgs a) {
    constexpr int TAPS = C * KH * KW;
    extern __shared__ __align__(16) float smem[];
    __shared__ float red[2 * kWarps];
    const int tid = threadIdx.x;
    const int strip = blockIdx.x % a.strips, n = blockIdx.x / a.strips;
    const int r0 = strip * a.rows, nr = min(a.rows, a.OH - r0);
    const int in_r0 = r0 * a.stride - a.pad;
    const int in_rows = (nr - 1) * a.stride + (KH - 1) * a.dil + 1;
    const int pw = a.W + 2 * a.pad;
    const int plane = in_rows * pw;
    const int Op = (a.O + 3) & ~3;                      // weight row length, padded to 128-bit words
    float *sw = smem;                                   // [TAPS][Op]
    float *sb = smem + TAPS * Op;                       // [Op] bias
    float *sx = sb + Op;                                // [C][in_rows][pw]
    const bool on = a.slot != nullptr;
    QuantConsts qc;
    if (on) qc = consts_from_slot(*a.slot, a.bits, 0, 0.f, 0.f);
    for (int idx = tid; idx < TAPS * Op; idx += kThreads) {
        const int t = idx / Op, o = idx - t * Op;
        sw[idx] = (o < a.O) ? __ldg(a.w + (int64_t)o * TAPS + t) : 0.0f;
    }
    for (int o = tid; o < Op; o += kThreads) sb[o] = (a.bias && o < a.O) ? __ldg(a.bias + o) : 0.0f;
    {   // warp per input row: 128-bit loads, quantise, scalar stores behind the left pad; no per-element division
        const int warp = tid >> 5, lane = tid & 31;
        const bool vec = (a.W & 3) == 0 && (((uintptr_t)a.x) & 15) == 0;
        for (int rid = warp; rid < C * in_rows; rid += kWarps) {
            const int c = rid / in_rows, r = rid - c * in_rows;
            const int gr = in_r0 + r;
            float *dst = sx + c * plane + r * pw;
            if (gr < 0 || gr >= a.H) {
                for (int col = lane; col < pw; col += 32) dst[col] = 0.0f;
                continue;
            }
            const float *src = a.x + (((int64_t)n * C + c) * a.H + gr) * a.W;
            for (int col = lane; col < a.pad; col += 32) { dst[col] = 0.0f; dst[a.pad + a.W + col] = 0.0f; }
            if (vec) {
                for (int v4 = lane; v4 < (a.W >> 2); v4 += 32) {
                    const float4 v = __ldg(reinterpret_cast<const float4 *>(src) + v4);
                    float *d = dst + a.pad + 4 * v4;
                    d[0] = quant_or_pass(v.x, qc, on); d[1] = quant_or_pass(v.y, qc, on);
                    d[2] = quant_or_pass(v.z, qc, on); d[3] = quant_or_pass(v.w, qc, on);
                }
            } else {
                for (int col = lane; col < a.W; col += 32) dst[a.pad + col] = quant_or_pass(__ldg(src + col), qc, on);
            }
        }
    }
    __syncthreads();
    float lo = INFINITY, hi = -INFINITY;
    const int outs = nr * a.OW;
    for (int p = tid; p < outs; p += kThreads) {
        const int orow = p / a.OW, ocol = p - orow * a.OW;
        const float *s0 = sx + (orow * a.stride) * pw + ocol * a.stride;
        float xin[TAPS];
#pragma unroll
        for (int c = 0; c < C; ++c)
#pragma unroll
            for (int i = 0; i < KH; ++i)
#pragma unroll
                for (int j = 0; j < KW; ++j) xin[(c * KH + i) * KW + j] = s0[c * plane + i * a.dil * pw + j * a.dil];
        float *yp = a.y + (((int64_t)n * a.O) * a.OH + r0 + orow) * a.OW + ocol;
        const int64_t ostride = (int64_t)a.OH * a.OW;
        for (int o0 = 0; o0 < a.O; o0 += 16) {
            float acc[16];
#pragma unroll
            for (int t = 0; t < 16; ++t) acc[t] = 0.0f;
#pragma unroll
            for (int t = 0; t < TAPS; ++t) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    if (o0 + 4 * q < Op) {
                        const float4 w = *reinterpret_cast<const float4 *>(sw + t * Op + o0 + 4 * q);
                        acc[4 * q] = __fmaf_rn(xin[t], w.x, acc[4 * q]);
                        acc[4 * q + 1] = __fmaf_rn(xin[t], w.y, acc[4 * q + 1]);
                        acc[4 * q + 2] = __fmaf_rn(xin[t], w.z, acc[4 * q + 2]);
                        acc[4 * q + 3] = __fmaf_rn(xin[t], w.w, acc[4 * q + 3]);
                    }
                }
            }
#pragma unroll
            for (int t = 0; t < 16; ++t) {
                if (o0 + t < a.O) {
                    float r = acc[t];
                    if (a.bias) r = __fadd_rn(r, sb[o0 + t]);
                    if (a.relu) r = fmaxf(r, 0.0f);
                    yp[(o0 + t) * ostride] = r;
                    lo = fminf(lo, r); hi = fmaxf(hi, r);
                }
            }
        }
    }
    if (a.stats) stats_flush(a.stats, n, lo, hi, red);
}

// ---------------------------------------------------------------------------------------------------------
// Pointwise (1x1) convolution / linear layer in fp32 on CUDA cores: the exact-semantics path for layers whose
// weights are not on an 8-bit grid (QuantNConv2d with arbitrary weights) and the on-device comparator of the
// tcgen05 path.  Y[n,o,p] = sum_i W[o,i] * Q(X[n,i,p]) + b[o].
// Generic element addressing: X element (n, i, p) at x + n*xs_n + i*xs_i + p*xs_p (conv: xs_i = HW, xs_p = 1;
// linear: one "image", p = sample, xs_p = K, xs_i = 1); Y likewise.
struct PwArgs {
    const float *x; float *y; const float *w; const float *bias; const RangeSlot *slot; unsigned *stats;
    int N, K, O, P;                       // images, input channels, output channels, pixels per image
    int64_t xs_n, xs_i, xs_p, ys_n, ys_o, ys_p;
    int relu, bits, stats_per_pixel;      // stats_per_pixel: linear layer, the "sample" is the pixel index
};

constexpr int kPwPix = kThreads, kPwK = 32, kPwO = 16;

__global__ void __launch_bounds__(kPwPix)
pw_fp32_kernel(PwArgs a) {
    __shared__ float sx[kPwK][kPwPix];
    __shared__ float sw[kPwO][kPwK + 1];
    __shared__ float red[2 * kWarps];
    const int tid = threadIdx.x;
    const int ptiles = (a.P + kPwPix - 1) / kPwPix;
    const int pt = blockIdx.x % ptiles, n = blockIdx.x / ptiles;
    const int p = pt * kPwPix + tid;
    const bool live = p < a.P;
    const bool on = a.slot != nullptr;
    QuantConsts qc;
    if (on) qc = consts_from_slot(*a.slot, a.bits, 0, 0.f, 0.f);
    float lo = INFINITY, hi = -INFINITY;
    for (int o0 = 0; o0 < a.O; o0 += kPwO) {
        float acc[kPwO];
#pragma unroll
        for (int t = 0; t < kPwO; ++t) acc[t] = 0.0f;
        for (int k0 = 0; k0 < a.K; k0 += kPwK) {
            __syncthreads();
            for (int kk = 0; kk < kPwK; ++kk) {
                float v = 0.0f;
                if (live && k0 + kk < a.K)
                    v = quant_or_pass(__ldg(a.x + n * a.xs_n + (int64_t)(k0 + kk) * a.xs_i + p * a.xs_p), qc, on);
                sx[kk][tid] = v;
            }
            for (int idx = tid; idx < kPwO * kPwK; idx += kPwPix) {
                int t = idx / kPwK, kk = idx - t * kPwK;
                sw[t][kk] = (o0 + t < a.O && k0 + kk < a.K) ? __ldg(a.w + (int64_t)(o0 + t) * a.K + k0 + kk) : 0.0f;
            }
            __syncthreads();
#pragma unroll 4
            for (int kk = 0; kk < kPwK; ++kk) {
                float v = sx[kk][tid];
#pragma unroll
                for (int t = 0; t < kPwO; ++t) acc[t] = __fmaf_rn(v, sw[t][kk], acc[t]);
            }
        }
        if (live) {
#pragma unroll
            for (int t = 0; t < kPwO; ++t) {
                if (o0 + t >= a.O) break;
                float r = acc[t];
                if (a.bias) r = __fadd_rn(r, __ldg(a.bias + o0 + t));
                if (a.relu) r = fmaxf(r, 0.0f);
                a.y[n * a.ys_n + (int64_t)(o0 + t) * a.ys_o + p * a.ys_p] = r;
                lo = fminf(lo, r); hi = fmaxf(hi, r);
            }
        }
    }
    if (a.stats) {
        if (a.stats_per_pixel) {               // linear: every "pixel" is a sample with its own extrema
            if (live) {
                atomicMax(a.stats + 2 * p, enc_ordered(-lo));
                atomicMax(a.stats + 2 * p + 1, enc_ordered(hi));
            }
        } else {
            stats_flush(a.stats, n, lo, hi, red);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Quantised residual add (utils/layer_transform.py:18-32): out = Q_a(a) + Q_b(b), each operand through its own
// QuantMeasure; per-sample extrema of `out` for the next layer's QuantMeasure.
__global__ void __launch_bounds__(kThreads)
qadd_kernel(const float *__restrict__ xa, const float *__restrict__ xb, float *__restrict__ out, int64_t chw,
            int chunks, const RangeSlot *sa, const RangeSlot *sb, int bits_a, int bits_b, unsigned *stats) {
    __shared__ float red[2 * kWarps];
    const int n = blockIdx.x / chunks, ch = blockIdx.x % chunks;
    QuantConsts ca, cb;
    if (sa) ca = consts_from_slot(*sa, bits_a, 0, 0.f, 0.f);
    if (sb) cb = consts_from_slot(*sb, bits_b, 0, 0.f, 0.f);
    const int64_t base = (int64_t)n * chw;
    const int64_t begin = (int64_t)ch * kChunk, end = min(begin + (int64_t)kChunk, chw);
    float lo = INFINITY, hi = -INFINITY;
    const bool vec = (((uintptr_t)(xa + base + begin) | (uintptr_t)(xb + base + begin) | (uintptr_t)(out + base + begin)) & 15) == 0;
    int64_t n4 = vec ? ((end - begin) >> 2) : 0;
    const float4 *a4 = reinterpret_cast<const float4 *>(xa + base + begin);
    const float4 *b4 = reinterpret_cast<const float4 *>(xb + base + begin);
    float4 *o4 = reinterpret_cast<float4 *>(out + base + begin);
    for (int64_t k = threadIdx.x; k < n4; k += kThreads) {
        float4 va = __ldg(a4 + k), vb = __ldg(b4 + k), r;
        r.x = __fadd_rn(quant_or_pass(va.x, ca, sa), quant_or_pass(vb.x, cb, sb));
        r.y = __fadd_rn(quant_or_pass(va.y, ca, sa), quant_or_pass(vb.y, cb, sb));
        r.z = __fadd_rn(quant_or_pass(va.z, ca, sa), quant_or_pass(vb.z, cb, sb));
        r.w = __fadd_rn(quant_or_pass(va.w, ca, sa), quant_or_pass(vb.w, cb, sb));
        o4[k] = r;
        lo = fminf(lo, fminf(fminf(r.x, r.y), fminf(r.z, r.w)));
        hi = fmaxf(hi, fmaxf(fmaxf(r.x, r.y), fmaxf(r.z, r.w)));
    }
    for (int64_t k = begin + (n4 << 2) + threadIdx.x; k < end; k += kThreads) {
        float r = __fadd_rn(quant_or_pass(xa[base + k], ca, sa), quant_or_pass(xb[base + k], cb, sb));
        out[base + k] = r;
        lo = fminf(lo, r); hi = fmaxf(hi, r);
    }
    if (stats) stats_flush(stats, n, lo, hi, red);
}

// Quantised spatial mean (utils/layer_transform.py:79-94 + MobileNetV2.py:112): out[n,c] = mean_hw Q(x[n,c,:]).
// A block owns 32 channels of one image (a warp walks 4 of them); fp32 tree sum, one division.
__global__ void __launch_bounds__(kThreads)
qmean_kernel(const float *__restrict__ x, float *__restrict__ out, int C, int HW, const RangeSlot *slot, int bits,
             unsigned *stats) {
    __shared__ float red[2 * kWarps];
    const int groups = (C + 31) / 32;
    const int n = blockIdx.x / groups, c0 = (blockIdx.x % groups) * 32;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    QuantConsts qc;
    if (slot) qc = consts_from_slot(*slot, bits, 0, 0.f, 0.f);
    float lo = INFINITY, hi = -INFINITY;
    // planes of <= 64 pixels (7 x 7 at the end of MobileNetV2): the loads of the warp's 4 channels are issued
    // together -- one channel at a time the warp waited out a memory round trip and a shuffle chain per channel
    float pre[4][2];
    const bool small = HW <= 64;
    if (small) {
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int c = c0 + warp * 4 + t;
            const float *p = x + ((int64_t)n * C + c) * HW;
            pre[t][0] = (c < C && lane < HW) ? __ldg(p + lane) : 0.0f;
            pre[t][1] = (c < C && lane + 32 < HW) ? __ldg(p + lane + 32) : 0.0f;
        }
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        int c = c0 + warp * 4 + t;
        if (c >= C) break;
        const float *p = x + ((int64_t)n * C + c) * HW;
        float s = 0.0f;
        if (small) {                                     // same order of additions as the loop below
            if (lane < HW) s = __fadd_rn(s, quant_or_pass(pre[t][0], qc, slot));
            if (lane + 32 < HW) s = __fadd_rn(s, quant_or_pass(pre[t][1], qc, slot));
        } else {
            for (int k = lane; k < HW; k += 32) s = __fadd_rn(s, quant_or_pass(__ldg(p + k), qc, slot));
        }
        s = warp_sum(s);
        float m = __fdiv_rn(s, (float)HW);
        if (lane == 0) out[(int64_t)n * C + c] = m;
        lo = fminf(lo, m); hi = fmaxf(hi, m);
    }
    if (stats) stats_flush(stats, n, lo, hi, red);
}

}  // namespace dfq

using namespace dfq;

// Chooses the kernel for one layer: depthwise 3x3 / small dense kxk / pointwise (fp32 CUDA-core form).
extern "C" int dfq_qconv2d_f32(const DfqConvDesc *d, void *stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (!d || d->N <= 0) return (int)cudaErrorInvalidValue;
    const int OH = (d->H + 2 * d->pad - d->dil * (d->kh - 1) - 1) / d->stride + 1;
    const int OW = (d->W + 2 * d->pad - d->dil * (d->kw - 1) - 1) / d->stride + 1;
    if (OH != d->OH || OW != d->OW) return (int)cudaErrorInvalidValue;
    if (d->kh == 1 && d->kw == 1 && d->groups == 1 && d->stride == 1 && d->pad == 0) {
        PwArgs a;
        a.x = d->x; a.y = d->y; a.w = d->w; a.bias = d->bias; a.slot = (const RangeSlot *)d->in_slot;
        a.stats = d->out_stats;
        a.N = d->N; a.K = d->C_in; a.O = d->C_out; a.P = d->H * d->W;
        a.xs_n = (int64_t)d->C_in * a.P; a.xs_i = a.P; a.xs_p = 1;
        a.ys_n = (int64_t)d->C_out * a.P; a.ys_o = a.P; a.ys_p = 1;
        a.relu = d->relu; a.bits = d->bits; a.stats_per_pixel = 0;
        int ptiles = (a.P + kPwPix - 1) / kPwPix;
        pw_fp32_kernel<<<(unsigned)(ptiles * d->N), kPwPix, 0, st>>>(a);
        g_launches += 1;
        return (int)cudaGetLastError();
    }
    ConvArgs a;
    a.x = d->x; a.y = d->y; a.w = d->w; a.bias = d->bias; a.slot = (const RangeSlot *)d->in_slot; a.stats = d->out_stats;
    a.N = d->N; a.C = d->C_in; a.H = d->H; a.W = d->W; a.O = d->C_out; a.OH = OH; a.OW = OW;
    a.stride = d->stride; a.pad = d->pad; a.dil = d->dil; a.relu = d->relu; a.bits = d->bits; a.kh = d->kh; a.kw = d->kw;
    const int pw = d->W + 2 * d->pad;
    if (d->groups == d->C_in && d->C_in == d->C_out && d->kh == 3 && d->kw == 3 && d->pad == 1 && d->dil == 1 &&
        (d->stride == 1 || d->stride == 2)) {
        if ((d->W % 4) == 0 && d->W >= 16 && d->W <= 128 && ((uintptr_t)d->x % 16) == 0 && ((uintptr_t)d->y % 16) == 0 &&
            (d->stride == 1 || (d->H % 2) == 0) && d->N <= 65535) {
            DwRowArgs r;
            r.x = d->x; r.y = d->y; r.w = d->w; r.bias = d->bias; r.slot = (const RangeSlot *)d->in_slot; r.stats = d->out_stats;
            r.C = d->C_in; r.H = d->H; r.W = d->W; r.OH = OH; r.OW = OW; r.relu = d->relu; r.bits = d->bits;
            r.lanes = d->W / 4; r.G = 32 / r.lanes;
            r.nblk = (OH + 15) / 16; r.R = (OH + r.nblk - 1) / r.nblk;
            r.items = d->C_in * r.nblk; r.wps = (r.items + r.G - 1) / r.G;
            dim3 grid((unsigned)((r.wps + kWarps - 1) / kWarps), (unsigned)d->N);
            if (d->stride == 1) dwconv3x3_rows_kernel<1><<<grid, kThreads, 0, st>>>(r);
            else dwconv3x3_rows_kernel<2><<<grid, kThreads, 0, st>>>(r);
            g_launches += 1;
            return (int)cudaGetLastError();
        }
        if (d->W <= 32 && d->N <= 65535) {
            DwColArgs r;
            r.x = d->x; r.y = d->y; r.w = d->w; r.bias = d->bias; r.slot = (const RangeSlot *)d->in_slot; r.stats = d->out_stats;
            r.C = d->C_in; r.H = d->H; r.W = d->W; r.OH = OH; r.OW = OW; r.relu = d->relu; r.bits = d->bits;
            r.G = 32 / d->W;
            static int cols_k = -1;                                  // DFQ_DW_COLS_K: planes in sequence per lane group (A/B)
            if (cols_k < 0) { const char *e = getenv("DFQ_DW_COLS_K"); cols_k = e ? atoi(e) : 0; }
            r.K = cols_k > 0 ? cols_k : (d->C_in >= 64 * r.G ? 4 : 1);
            r.wps = (d->C_in + r.G * r.K - 1) / (r.G * r.K);
            dim3 grid((unsigned)((r.wps + kWarps - 1) / kWarps), (unsigned)d->N);
            if (d->stride == 1) dwconv3x3_cols_kernel<1><<<grid, kThreads, 0, st>>>(r);
            else dwconv3x3_cols_kernel<2><<<grid, kThreads, 0, st>>>(r);
            g_launches += 1;
            return (int)cudaGetLastError();
        }
        DwArgs w;
        w.x = d->x; w.y = d->y; w.w = d->w; w.bias = d->bias; w.slot = (const RangeSlot *)d->in_slot; w.stats = d->out_stats;
        w.C = d->C_in; w.H = d->H; w.W = d->W; w.OH = OH; w.OW = OW; w.stride = d->stride; w.relu = d->relu; w.bits = d->bits;
        w.RS = (d->W + 5 + 3) / 4 * 4;
        w.G4 = (OW + 3) / 4;
        auto tile_bytes = [&](int rows, int cb) {
            return ((size_t)cb * ((rows - 1) * d->stride + 3) * w.RS + 8) * sizeof(float);
        };
        // a strip of one plane (about 1024 four-output items) or, for small planes, several whole planes
        int rows = OH, cb = 1;
        if (OH * w.G4 > 1536) rows = max(1, 1024 / w.G4);
        else cb = max(1, min(d->C_in, 1024 / (OH * w.G4)));
        while (tile_bytes(rows, cb) > 40 * 1024 && (cb > 1 || rows > 1)) { if (cb > 1) cb = (cb + 1) / 2; else rows = (rows + 1) / 2; }
        // whole planes of several channels must start on a 16-byte boundary for the vector path
        const bool vec = (d->W % 4) == 0 && ((uintptr_t)d->x % 16) == 0;
        w.rows = rows; w.cb = cb; w.strips = (OH + rows - 1) / rows;
        w.inv_W = 1.0f / (float)d->W; w.inv_W4 = 4.0f / (float)d->W; w.inv_G4 = 1.0f / (float)w.G4;
        const int blocks = d->N * ((d->C_in + cb - 1) / cb) * w.strips;
        if (vec) dwconv3x3_stream_kernel<true><<<(unsigned)blocks, kThreads, tile_bytes(rows, cb), st>>>(w);
        else dwconv3x3_stream_kernel<false><<<(unsigned)blocks, kThreads, tile_bytes(rows, cb), st>>>(w);
        g_launches += 1;
        return (int)cudaGetLastError();
    }
    if (d->groups == d->C_in && d->C_in == d->C_out && d->kh == 3 && d->kw == 3 && d->pad == 0 &&
        ((d->stride == 1 && (d->dil == 1 || d->dil == 2)) || (d->stride == 2 && d->dil == 1))) {
        // valid depthwise (DeepLab after F.pad): about 4096 outputs per block in row groups of 4, <= 48 KB of input;
        // planes of <= 2048 outputs: one plane per WARP, 8 channels per block
        const int S = d->stride, D = d->dil;
        int rows = OH, cb = 1;
        auto smem_of = [&](int r, int c) { return ((size_t)c * ((((r - 1) * S + 2 * D + 1) * d->W + 6) & ~3) + (size_t)c * 10) * sizeof(float); };
        bool warp_mode = OH * OW <= 2048 && d->C_in >= kWarps && smem_of(OH, kWarps) <= 48 * 1024;
        if (warp_mode) cb = kWarps;
        else if (OH * OW > 4096) rows = max(4, (4096 / OW) & ~3);
        else cb = max(1, min(d->C_in, 4096 / (OH * OW)));
        while (smem_of(rows, cb) > 48 * 1024 && (cb > 1 || rows > 4)) { if (cb > 1) cb = (cb + 1) / 2; else rows = max(4, (rows / 2 + 3) & ~3); }
        if (smem_of(rows, cb) <= 48 * 1024) {
            a.rows = rows; a.cb = cb; a.strips = (OH + rows - 1) / rows;
            const int64_t blocks = (int64_t)d->N * ((d->C_in + cb - 1) / cb) * a.strips;
            if (blocks > 0x7fffffffll) return (int)cudaErrorInvalidValue;
            const size_t smem = smem_of(rows, cb);
            const unsigned g = (unsigned)blocks;
            if (warp_mode) {
                if (S == 1 && D == 1) dwconv3x3_valid_kernel<1, 1, true><<<g, kThreads, smem, st>>>(a);
                else if (S == 2) dwconv3x3_valid_kernel<2, 1, true><<<g, kThreads, smem, st>>>(a);
                else dwconv3x3_valid_kernel<1, 2, true><<<g, kThreads, smem, st>>>(a);
            } else {
                if (S == 1 && D == 1) dwconv3x3_valid_kernel<1, 1, false><<<g, kThreads, smem, st>>>(a);
                else if (S == 2) dwconv3x3_valid_kernel<2, 1, false><<<g, kThreads, smem, st>>>(a);
                else dwconv3x3_valid_kernel<1, 2, false><<<g, kThreads, smem, st>>>(a);
            }
            g_launches += 1;
            return (int)cudaGetLastError();
        }
    }
    if (d->groups == d->C_in && d->C_in == d->C_out && d->kh == 3 && d->kw == 3) {
        // strip height: about 2048 outputs per block; small planes take several channels per block
        int rows = OH, cb = 1;
        if (OH * OW > 2048) rows = max(1, 2048 / OW);
        else cb = max(1, min(d->C_in, 2048 / (OH * OW)));
        auto smem_of = [&](int r, int c) { return (size_t)c * ((r - 1) * d->stride + 2 * d->dil + 1) * pw * sizeof(float); };
        while (smem_of(rows, cb) > 96 * 1024 && (cb > 1 || rows > 1)) { if (cb > 1) cb = (cb + 1) / 2; else rows = (rows + 1) / 2; }
        a.rows = rows; a.cb = cb; a.strips = (OH + rows - 1) / rows;
        size_t smem = smem_of(rows, cb);
        if (smem > 48 * 1024) cudaFuncSetAttribute(dwconv3x3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
        int blocks = d->N * ((d->C_in + cb - 1) / cb) * a.strips;
        dwconv3x3_kernel<<<(unsigned)blocks, kThreads, smem, st>>>(a);
        g_launches += 1;
        return (int)cudaGetLastError();
    }
    if (d->groups == 1) {
        size_t wbytes = (size_t)d->C_out * d->C_in * d->kh * d->kw * sizeof(float);
        int rows = max(1, 1024 / OW);
        auto smem_of = [&](int r) { return wbytes + (size_t)d->C_in * ((r - 1) * d->stride + (d->kh - 1) * d->dil + 1) * pw * sizeof(float); };
        while (smem_of(rows) > 96 * 1024 && rows > 1) rows = (rows + 1) / 2;
        if (smem_of(rows) > 96 * 1024) return (int)cudaErrorInvalidValue;   // large dense kxk: tensor-core path (not built yet)
        if (d->C_in == 3 && d->kh == 3 && d->kw == 3) {      // the stem: register-blocked form
            const int Op = (d->C_out + 3) & ~3;
            int rr = max(1, 1024 / OW);
            auto smem_reg = [&](int r) {
                return ((size_t)27 * Op + Op + (size_t)3 * ((r - 1) * d->stride + 2 * d->dil + 1) * pw) * sizeof(float);
            };
            while (smem_reg(rr) > 64 * 1024 && rr > 1) rr = (rr + 1) / 2;
            if (smem_reg(rr) <= 64 * 1024) {
                a.rows = rr; a.cb = 0; a.strips = (OH + rr - 1) / rr;
                static bool attr = false;
                if (!attr) {
                    cudaFuncSetAttribute(conv_direct_reg_kernel<3, 3, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
                    attr = true;
                }
                conv_direct_reg_kernel<3, 3, 3><<<(unsigned)(d->N * a.strips), kThreads, smem_reg(rr), st>>>(a);
                g_launches += 1;
                return (int)cudaGetLastError();
            }
        }
        a.rows = rows; a.cb = 0; a.strips = (OH + rows - 1) / rows;
        if (smem_of(rows) > 48 * 1024) cudaFuncSetAttribute(conv_direct_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
        conv_direct_kernel<<<(unsigned)(d->N * a.strips), kThreads, smem_of(rows), st>>>(a);
        g_launches += 1;
        return (int)cudaGetLastError();
    }
    return (int)cudaErrorInvalidValue;
}

// Linear layer through the same fp32 path: x [N, K], w [O, K], y [N, O]; out_stats per sample.
extern "C" int dfq_qlinear_f32(const float *x, float *y, const float *w, const float *bias, int64_t N, int K, int O,
                               const float *in_slot, int bits, int relu, unsigned *out_stats, void *stream) {
    if (N <= 0) return 0;
    PwArgs a;
    a.x = x; a.y = y; a.w = w; a.bias = bias; a.slot = (const RangeSlot *)in_slot; a.stats = out_stats;
    a.N = 1; a.K = K; a.O = O; a.P = (int)N;
    a.xs_n = 0; a.xs_i = 1; a.xs_p = K; a.ys_n = 0; a.ys_o = 1; a.ys_p = O;
    a.relu = relu; a.bits = bits; a.stats_per_pixel = 1;
    pw_fp32_kernel<<<(unsigned)((N + kPwPix - 1) / kPwPix), kPwPix, 0, (cudaStream_t)stream>>>(a);
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_qadd_f32(const float *a, const float *b, float *out, int64_t N, int64_t chw, const float *slot_a,
                            const float *slot_b, int bits_a, int bits_b, unsigned *out_stats, void *stream) {
    if (N <= 0 || chw <= 0) return 0;
    int chunks = (int)((chw + kChunk - 1) / kChunk);
    qadd_kernel<<<(unsigned)(N * chunks), kThreads, 0, (cudaStream_t)stream>>>(
        a, b, out, chw, chunks, (const RangeSlot *)slot_a, (const RangeSlot *)slot_b, bits_a, bits_b, out_stats);
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_qmean_f32(const float *x, float *out, int64_t N, int C, int HW, const float *slot, int bits,
                             unsigned *out_stats, void *stream) {
    if (N <= 0) return 0;
    qmean_kernel<<<(unsigned)(N * ((C + 31) / 32)), kThreads, 0, (cudaStream_t)stream>>>(
        x, out, C, HW, (const RangeSlot *)slot, bits, out_stats);
    g_launches += 1;
    return (int)cudaGetLastError();
}
