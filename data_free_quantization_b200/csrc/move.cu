// Kernel (e), data-movement ops of the quantised eval forward: bilinear interpolation and channel concatenation
// with their operands fake-quantised on load (utils/layer_transform.py:98-125 route the operands of
// F.interpolate / torch.cat through QuantMeasures first), and constant zero padding (utils/layer_transform.py
// ignore_op=['pad']: recorded in the graph, never quantised).  Each is ONE pass: read the source once, write the
// destination once, and reduce the per-sample min/max of what was written into the encoded [N,2] table the next
// QuantMeasure consumes (utils/quantize.py:110-111) -- no quantised temporary, no separate statistics pass.
//
// Planes of 129 x 129 or 513 x 513 floats are not 16-byte aligned, so cat and pad address the destination of one
// sample as a FLAT run of floats: a block owns kChunk consecutive floats of the run (in a coordinate shifted so
// that multiples of 4 are 16-byte aligned), a thread owns groups of 4 (one 128-bit store) and decodes (plane, row,
// column) once per group, then steps.  The interpolation works on bands of rows (see qinterp_kernel).
#include "common.cuh"

namespace dfq {

// Drives one sample's flat destination run.  f.four(k, v) fills v[0..3] for elements k..k+3, f.one(k) one element.
template <typename F>
__device__ __forceinline__ void flat_run(float *__restrict__ dst, int count, int chunk, F &f, float &lo, float &hi) {
    const int m = (int)((((uintptr_t)dst) >> 2) & 3);           // dst - m is 16-byte aligned
    const int j_end = m + count;
#pragma unroll 1
    for (int u = 0; u < kChunk / (4 * kThreads); ++u) {
        const int j0 = chunk * kChunk + 4 * ((int)threadIdx.x + u * kThreads);
        if (j0 >= j_end) break;
        if (j0 >= m && j0 + 4 <= j_end) {
            float v[4];
            f.four(j0 - m, v);
            *reinterpret_cast<float4 *>(dst - m + j0) = make_float4(v[0], v[1], v[2], v[3]);
            lo = fminf(lo, fminf(fminf(v[0], v[1]), fminf(v[2], v[3])));
            hi = fmaxf(hi, fmaxf(fmaxf(v[0], v[1]), fmaxf(v[2], v[3])));
        } else {
            for (int j = max(j0, m); j < min(j0 + 4, j_end); ++j) {
                const float r = f.one(j - m);
                dst[j - m] = r;
                lo = fminf(lo, r); hi = fmaxf(hi, r);
            }
        }
    }
}
static inline int flat_chunks(int64_t count) { return (int)((count + 3 + kChunk - 1) / kChunk); }

// ---------------------------------------------------------------------------------------------------------
// Bilinear interpolation of Q(x): out[n,c,oy,ox] = l0y (l0x a + l1x b) + l1y (l0x c + l1x d), the source index and
// the two weights formed as ATen does (UpSample.h area_pixel_compute_source_index / compute_scales_value: fp32
// scale, src = scale * dst for align_corners, scale * (dst + 0.5) - 0.5 clamped at 0 otherwise; i1 = i0 + (i0 <
// size - 1); l1 = src - i0, l0 = 1 - l1).  Products and sums individually rounded.
//
// A block owns a band of `band` output rows of one plane.  The source rows the band touches are one contiguous
// run: it is read coalesced, quantised ONCE and kept in shared memory next to a per-column table {x0, x1, l1x}, so
// an output costs four shared loads and the seven-operation blend (the first form of this kernel quantised its taps
// per output and sent the one group per row that wraps to a per-element path: 120 thread-instructions per output,
// half of every warp idle -- ncu, profiles/r1zz_move_ncu.md).  The band's destination is a flat run of floats.
__device__ __forceinline__ void interp_index(int o, float scale, int size, int align, int &i0, int &i1, float &l1) {
    const float s = align ? __fmul_rn(scale, (float)o)
                          : fmaxf(__fsub_rn(__fmul_rn(scale, __fadd_rn((float)o, 0.5f)), 0.5f), 0.0f);
    i0 = min((int)s, size - 1);
    i1 = i0 + (i0 < size - 1 ? 1 : 0);
    l1 = fminf(fmaxf(__fsub_rn(s, (float)i0), 0.0f), 1.0f);
}

__global__ void __launch_bounds__(kThreads)
qinterp_kernel(const float *__restrict__ x, float *__restrict__ y, int C, int H, int W, int OH, int OW, int align,
               float sy, float sx, int band, int bands, int src_rows_max, const RangeSlot *slot, int bits,
               unsigned *stats) {
    extern __shared__ __align__(16) float sm_dyn[];
    __shared__ float red[2 * kWarps];
    float *src = sm_dyn;                                          // [src_rows_max * W] quantised source rows
    // tables: per output column {x0 | x1 << 16, l1x}, per output row of the band {row offsets of y0 / y1, l1y}
    int2 *xtab = reinterpret_cast<int2 *>(sm_dyn + ((src_rows_max * W + 3) & ~3));      // 16-byte aligned
    int4 *ytab = reinterpret_cast<int4 *>(xtab + ((OW + 1) & ~1));
    int b = blockIdx.x;
    const int bi = b % bands; b /= bands;
    const int c = b % C, n = b / C;
    const int r0 = bi * band, r1 = min(r0 + band, OH);
    int ys0, ys1, t0, t1; float tl;
    interp_index(r0, sy, H, align, ys0, t1, tl);
    interp_index(r1 - 1, sy, H, align, t0, ys1, tl);
    if (ys1 < ys0) { const int t = ys0; ys0 = ys1; ys1 = t; }     // (monotone for every scale; kept for safety)
    const int nsrc = min(ys1 - ys0 + 1, src_rows_max) * W;
    const float *p = x + (((int64_t)n * C + c) * H + ys0) * W;
    const bool on = slot != nullptr;
    QuantConsts qc;
    if (on) qc = consts_from_slot(*slot, bits, 0, 0.f, 0.f);
    for (int i = threadIdx.x; i < nsrc; i += kThreads) src[i] = quant_or_pass(__ldg(p + i), qc, on);
    for (int ox = threadIdx.x; ox < OW; ox += kThreads) {
        int x0, x1; float l1;
        interp_index(ox, sx, W, align, x0, x1, l1);
        xtab[ox] = make_int2(x0 | (x1 << 16), __float_as_int(l1));
    }
    for (int r = threadIdx.x; r < r1 - r0; r += kThreads) {
        int y0, y1; float l1;
        interp_index(r0 + r, sy, H, align, y0, y1, l1);
        ytab[r] = make_int4((y0 - ys0) * W, (y1 - ys0) * W, __float_as_int(l1), 0);
    }
    __syncthreads();
    // A thread owns output COLUMNS (tid, tid + 256, ...) and walks the band's rows: its column's {x0, x1, l1x} live in
    // registers, the row table is a broadcast, and the horizontal blends `top` / `bot` are formed once per source row
    // pair -- when the image grows 4 x, four output rows share one, so an output costs the vertical blend, the store
    // and the extrema (the one-output-per-step form re-read both tables and all four taps for every output: 71
    // thread-instructions per output at 89 % issue-active under ncu).  Lanes sit on adjacent columns: coalesced 4-byte
    // stores whatever the row length.  Same operations per value in the same order: the same bits.
    float *dst = y + ((int64_t)n * C + c) * OH * OW + (int64_t)r0 * OW;
    const int nrows = r1 - r0;
    float lo = INFINITY, hi = -INFINITY;
    for (int ox = threadIdx.x; ox < OW; ox += kThreads) {
        const int2 e = xtab[ox];
        const int x0 = e.x & 0xffff, x1 = e.x >> 16;
        const float l1x = __int_as_float(e.y), l0x = __fsub_rn(1.0f, l1x);
        int cy0 = -1, cy1 = -1;
        float top = 0.0f, bot = 0.0f;
        float *d = dst + ox;
        for (int r = 0; r < nrows; ++r, d += OW) {
            const int4 yy = ytab[r];
            if (yy.x != cy0 || yy.y != cy1) {                      // (uniform across the block)
                cy0 = yy.x; cy1 = yy.y;
                const float *s0 = src + yy.x, *s1 = src + yy.y;
                top = __fadd_rn(__fmul_rn(l0x, s0[x0]), __fmul_rn(l1x, s0[x1]));
                bot = __fadd_rn(__fmul_rn(l0x, s1[x0]), __fmul_rn(l1x, s1[x1]));
            }
            const float l1y = __int_as_float(yy.z), l0y = __fsub_rn(1.0f, l1y);
            const float v = __fadd_rn(__fmul_rn(l0y, top), __fmul_rn(l1y, bot));
            *d = v;
            lo = fminf(lo, v); hi = fmaxf(hi, v);
        }
    }
    if (stats) stats_flush(stats, n, lo, hi, red);
}

// ---------------------------------------------------------------------------------------------------------
// One operand of a channel concatenation: dst[n, c_off : c_off + C_src] = Q(src[n]).  Also the plain quantised copy.
struct CopyFn {
    const float *x;            // this sample's source run
    QuantConsts qc;
    bool on, vec;              // vec: source and destination runs share their 16-byte phase
    __device__ __forceinline__ float one(int k) const { return quant_or_pass(__ldg(x + k), qc, on); }
    __device__ __forceinline__ void four(int k, float *v) const {
        if (vec) {
            const float4 t = __ldg(reinterpret_cast<const float4 *>(x + k));
            if (on && qc.fast) {
                const float4 r = quant_dequant4_fast(t, qc);
                v[0] = r.x; v[1] = r.y; v[2] = r.z; v[3] = r.w;
            } else {
                v[0] = quant_or_pass(t.x, qc, on); v[1] = quant_or_pass(t.y, qc, on);
                v[2] = quant_or_pass(t.z, qc, on); v[3] = quant_or_pass(t.w, qc, on);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) v[i] = one(k + i);
        }
    }
};

__global__ void __launch_bounds__(kThreads)
qcopy_kernel(const float *__restrict__ x, float *__restrict__ y, int count, int64_t src_stride, int64_t dst_stride,
             int chunks, const RangeSlot *slot, int bits, unsigned *stats) {
    __shared__ float red[2 * kWarps];
    const int n = blockIdx.x / chunks, ch = blockIdx.x - n * chunks;
    CopyFn f;
    f.x = x + (int64_t)n * src_stride;
    float *dst = y + (int64_t)n * dst_stride;
    f.on = slot != nullptr;
    if (slot) f.qc = consts_from_slot(*slot, bits, 0, 0.f, 0.f);
    f.vec = ((((uintptr_t)f.x) ^ ((uintptr_t)dst)) & 15) == 0;
    float lo = INFINITY, hi = -INFINITY;
    flat_run(dst, count, ch, f, lo, hi);
    if (stats) stats_flush(stats, n, lo, hi, red);
}

// ---------------------------------------------------------------------------------------------------------
// Constant zero padding of H and W (F.pad(x, (left, right, top, bottom))): no quantisation.
struct PadFn {
    const float *x;            // this sample's source, [C, H, W]
    int H, W, OH, OW, top, left;
    __device__ __forceinline__ float at(int plane, int oy, int ox) const {
        const int iy = oy - top, ix = ox - left;
        return ((unsigned)iy < (unsigned)H && (unsigned)ix < (unsigned)W)
                   ? __ldg(x + ((int64_t)plane * H + iy) * W + ix) : 0.0f;
    }
    __device__ __forceinline__ float one(int k) const {
        const int plane = k / (OH * OW), rem = k - plane * (OH * OW);
        const int oy = rem / OW;
        return at(plane, oy, rem - oy * OW);
    }
    __device__ __forceinline__ void four(int k, float *v) const {
        int plane = k / (OH * OW), rem = k - plane * (OH * OW);
        int oy = rem / OW, ox = rem - oy * OW;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[i] = at(plane, oy, ox);
            if (++ox == OW) { ox = 0; if (++oy == OH) { oy = 0; ++plane; } }
        }
    }
};

__global__ void __launch_bounds__(kThreads)
pad_kernel(const float *__restrict__ x, float *__restrict__ y, int C, int H, int W, int OH, int OW, int top, int left,
           int chunks, unsigned *stats) {
    __shared__ float red[2 * kWarps];
    const int n = blockIdx.x / chunks, ch = blockIdx.x - n * chunks;
    PadFn f;
    f.x = x + (int64_t)n * C * H * W;
    f.H = H; f.W = W; f.OH = OH; f.OW = OW; f.top = top; f.left = left;
    float lo = INFINITY, hi = -INFINITY;
    flat_run(y + (int64_t)n * C * OH * OW, C * OH * OW, ch, f, lo, hi);
    if (stats) stats_flush(stats, n, lo, hi, red);
}

// fp32 scale of ATen's compute_scales_value / area_pixel_compute_scale (no user scale factor)
static inline float interp_scale(int in, int out, int align) {
    if (align) return out > 1 ? (float)(in - 1) / (float)(out - 1) : 0.0f;
    return out > 0 ? (float)in / (float)out : 0.0f;
}

}  // namespace dfq

using namespace dfq;

extern "C" int dfq_qinterp_bilinear_f32(const float *x, float *y, int64_t N, int C, int H, int W, int OH, int OW,
                                        int align_corners, const float *in_slot, int bits, unsigned *out_stats,
                                        void *stream) {
    if (N <= 0 || C <= 0) return 0;
    if (H <= 0 || W <= 0 || OH <= 0 || OW <= 0 || W > 0x7fff || OW > 0xffff) return (int)cudaErrorInvalidValue;
    const int align = align_corners ? 1 : 0;
    const float sy = interp_scale(H, OH, align), sx = interp_scale(W, OW, align);
    // band of output rows per block: ~16 outputs per thread, shrunk until its source rows fit in shared memory
    int band = (kChunk + OW - 1) / OW;
    if (band > OH) band = OH;
    auto rows_of = [&](int bd) { int r = (int)ceilf((float)bd * fmaxf(sy, 0.0f)) + 3; return r < H ? r : H; };
    const size_t budget = 96 * 1024;
    auto smem_of = [&](int bd) { return (size_t)(((rows_of(bd) * W + 3) & ~3)) * 4 + (size_t)((OW + 1) & ~1) * 8 + (size_t)bd * 16; };
    while (band > 1 && smem_of(band) > budget) band = (band + 1) / 2;
    const size_t smem = smem_of(band);
    if (smem > 200 * 1024) return (int)cudaErrorInvalidValue;
    const int bands = (OH + band - 1) / band;
    const int64_t blocks = N * (int64_t)C * bands;
    if (blocks > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(qinterp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
    }
    qinterp_kernel<<<(unsigned)blocks, kThreads, smem, (cudaStream_t)stream>>>(
        x, y, C, H, W, OH, OW, align, sy, sx, band, bands, rows_of(band), (const RangeSlot *)in_slot, bits, out_stats);
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_qcopy_f32(const float *x, float *y, int64_t N, int64_t count, int64_t src_stride,
                             int64_t dst_stride, const float *in_slot, int bits, unsigned *out_stats, void *stream) {
    if (N <= 0 || count <= 0) return 0;
    if (count >= (1ll << 31) - kChunk) return (int)cudaErrorInvalidValue;
    const int chunks = flat_chunks(count);
    if (N * chunks > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    qcopy_kernel<<<(unsigned)(N * chunks), kThreads, 0, (cudaStream_t)stream>>>(
        x, y, (int)count, src_stride, dst_stride, chunks, (const RangeSlot *)in_slot, bits, out_stats);
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_pad2d_f32(const float *x, float *y, int64_t N, int C, int H, int W, int left, int right, int top,
                             int bottom, unsigned *out_stats, void *stream) {
    if (N <= 0 || C <= 0) return 0;
    const int OH = H + top + bottom, OW = W + left + right;
    if (H <= 0 || W <= 0 || left < 0 || right < 0 || top < 0 || bottom < 0) return (int)cudaErrorInvalidValue;
    const int64_t count = (int64_t)C * OH * OW;
    if (count >= (1ll << 31) - kChunk) return (int)cudaErrorInvalidValue;
    const int chunks = flat_chunks(count);
    if (N * chunks > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    pad_kernel<<<(unsigned)(N * chunks), kThreads, 0, (cudaStream_t)stream>>>(x, y, C, H, W, OH, OW, top, left, chunks,
                                                                              out_stats);
    g_launches += 1;
    return (int)cudaGetLastError();
}
