// Kernel (a): uniform fake-quantisation -- utils/quantize.py:24-78, utils/layer_transform.py:288-305.
// HBM-bound streaming passes: 128-bit coalesced loads, 4 independent loads per thread in flight,
// warp-shuffle + shared-memory block reductions, grid-level finish with atomicMax on an
// order-preserving integer encoding.  Algorithmic bytes: 4 B/elt (range pass) + 8 B/elt (apply).
#include "common.cuh"
#include "apply.cuh"

namespace dfq {

// ---------------------------------------------------------------------------------------------
// Work decomposition of the multi-tensor launches.  A block owns either one kChunk-element chunk of
// one range segment (per-tensor, or per-channel rows longer than kSmallRow) or kWarps short rows.
constexpr int kSmallRow = 1024;

struct BlockWork {
    int desc;
    int64_t row;        // first row (per-channel) or 0
    int64_t begin, end; // element range inside the tensor (chunk mode); begin<0 => warp-per-row mode
    int64_t cols;
};

__device__ __forceinline__ int find_desc(const DfqQuantDesc *d, int n, int64_t b) {
    int lo = 0, hi = n - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (d[mid].first_block <= b) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__device__ __forceinline__ BlockWork decode_work1(const DfqQuantDesc &d, int64_t j) {
    BlockWork w;
    w.desc = 0;
    if (!(d.flags & DFQ_Q_PER_CHANNEL)) {
        w.row = 0; w.cols = d.n;
        w.begin = j * kChunk;
        w.end = min(w.begin + (int64_t)kChunk, d.n);
    } else {
        w.cols = d.n / d.rows;
        if (w.cols > kSmallRow) {
            int64_t cpr = (w.cols + kChunk - 1) / kChunk;
            w.row = j / cpr;
            int64_t c0 = (j % cpr) * kChunk;
            w.begin = w.row * w.cols + c0;
            w.end = w.row * w.cols + min(c0 + (int64_t)kChunk, w.cols);
        } else {
            w.row = j * kWarps;
            w.begin = -1; w.end = -1;
        }
    }
    return w;
}

__device__ __forceinline__ BlockWork decode_work(const DfqQuantDesc *descs, int n_desc, int64_t b) {
    int t = find_desc(descs, n_desc, b);
    BlockWork w = decode_work1(descs[t], b - descs[t].first_block);
    w.desc = t;
    return w;
}

// Grid-stride range pass.  A block walks virtual chunks vb = blockIdx.x, +gridDim.x, ... and keeps a running
// (min, max) in registers while consecutive chunks belong to the same range slot; the block reduction and the
// two atomics happen once per slot change (once per block for a single big tensor) instead of once per 16 KB.
template <int kStride>
__device__ __forceinline__ void flush_slot(unsigned *__restrict__ ranges, int64_t slot, float lo, float hi, float *sm) {
    block_minmax(lo, hi, sm);
    if (threadIdx.x == 0) {
        atomicMax(ranges + kStride * slot, enc_ordered(-lo));
        atomicMax(ranges + kStride * slot + 1, enc_ordered(hi));
    }
}

template <int kStride>
__device__ __forceinline__ void warp_rows_minmax(const DfqQuantDesc &d, const BlockWork &w, unsigned *__restrict__ ranges) {
    int64_t r = w.row + (threadIdx.x >> 5);                 // one warp per short row, plain stores
    if (r >= d.rows) return;
    const float *p = d.x + r * w.cols;
    float lo = INFINITY, hi = -INFINITY;
    for (int64_t k = threadIdx.x & 31; k < w.cols; k += 32) {
        float t = p[k]; lo = fmin_nan(lo, t); hi = fmax_nan(hi, t);
    }
    lo = warp_min(lo); hi = warp_max(hi);
    if ((threadIdx.x & 31) == 0) {
        unsigned *slot = ranges + kStride * (d.range_slot + r);
        slot[0] = enc_ordered(-lo);
        slot[1] = enc_ordered(hi);
    }
}

// The descriptor a block is working in is looked up (binary search + dependent loads from the table) only when the
// walk leaves its block range: with a look-up per 16 KB chunk every chunk began with ~2 us of dependent loads before
// its first data load could issue, and the multi-tensor passes ran 15-30 % behind their single-tensor forms.
struct DescCursor {
    int t;
    int64_t first, end;     // virtual blocks [first, end) belong to descs[t]
    __device__ __forceinline__ DescCursor() : t(-1), first(0), end(0) {}
    __device__ __forceinline__ bool seek(const DfqQuantDesc *__restrict__ descs, int n_desc, int64_t total_blocks, int64_t vb) {
        if (vb >= first && vb < end) return false;
        t = find_desc(descs, n_desc, vb);
        first = descs[t].first_block;
        end = (t + 1 < n_desc) ? descs[t + 1].first_block : total_blocks;
        return true;
    }
};

__global__ void __launch_bounds__(kThreads, 6)
minmax_multi_kernel(const DfqQuantDesc *__restrict__ descs, int n_desc, int64_t total_blocks,
                    unsigned *__restrict__ ranges) {
    __shared__ float sm[2 * kWarps];
    float lo = INFINITY, hi = -INFINITY;
    int64_t cur = -1;
    DescCursor dc;
    for (int64_t vb = blockIdx.x; vb < total_blocks; vb += gridDim.x) {
        dc.seek(descs, n_desc, total_blocks, vb);
        const DfqQuantDesc &d = descs[dc.t];
        BlockWork w = decode_work1(d, vb - dc.first);
        if (w.begin >= 0) {
            int64_t slot = d.range_slot + w.row;
            if (slot != cur) {
                if (cur >= 0) flush_slot<4>(ranges, cur, lo, hi, sm);
                lo = INFINITY; hi = -INFINITY; cur = slot;
            }
            chunk_minmax(d.x, w.begin, w.end, lo, hi);
        } else {
            warp_rows_minmax<4>(d, w, ranges);
        }
    }
    if (cur >= 0) flush_slot<4>(ranges, cur, lo, hi, sm);
}

// one tensor, descriptor passed by value (dfq_minmax_f32 / dfq_minmax_rows_f32)
__global__ void __launch_bounds__(kThreads)
minmax_single_kernel(DfqQuantDesc d, int64_t total_blocks, unsigned *__restrict__ ranges) {
    __shared__ float sm[2 * kWarps];
    float lo = INFINITY, hi = -INFINITY;
    int64_t cur = -1;
    for (int64_t vb = blockIdx.x; vb < total_blocks; vb += gridDim.x) {
        BlockWork w = decode_work1(d, vb);
        if (w.begin >= 0) {
            int64_t slot = d.range_slot + w.row;
            if (slot != cur) {
                if (cur >= 0) flush_slot<2>(ranges, cur, lo, hi, sm);
                lo = INFINITY; hi = -INFINITY; cur = slot;
            }
            chunk_minmax(d.x, w.begin, w.end, lo, hi);
        } else {
            warp_rows_minmax<2>(d, w, ranges);
        }
    }
    if (cur >= 0) flush_slot<2>(ranges, cur, lo, hi, sm);
}

// encoded (max(-x), max(x)) -> plain (min, max) floats, in place
__global__ void decode_ranges_kernel(unsigned *ranges, int64_t slots) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= slots) return;
    float mn = -dec_ordered(ranges[2 * i]);
    float mx = dec_ordered(ranges[2 * i + 1]);
    reinterpret_cast<float *>(ranges)[2 * i] = mn;
    reinterpret_cast<float *>(ranges)[2 * i + 1] = mx;
}

// range table of the multi-tensor path: 4 floats per slot.  In: encoded (max(-x), max(x)) in [0], [1];
// out: RangeSlot {min, max, scale, 0} with the scale formed exactly as the host does (utils/quantize.py:51-68).
__global__ void decode_slots_kernel(const DfqQuantDesc *__restrict__ descs, int n_desc, unsigned *ranges, int64_t slots) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= slots) return;
    int lo = 0, hi = n_desc - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (descs[mid].range_slot <= i) lo = mid; else hi = mid - 1;
    }
    float mn = -dec_ordered(ranges[4 * i]);
    float mx = dec_ordered(ranges[4 * i + 1]);
    DfqQuantParams p = make_quant_params((double)mn, (double)mx, descs[lo].bits, descs[lo].flags);
    RangeSlot r = {mn, mx, p.scale, 0.0f};
    reinterpret_cast<RangeSlot *>(ranges)[i] = r;
}

// Apply pass: one block per virtual chunk (fresh blocks beat a persistent walk here: 452 vs 405 us on 1 GiB at the
// same occupancy).  One instantiation of the element loop for both code widths and both work modes keeps the kernel at
// 64 registers = 4 blocks per SM; the tensors are cudaMalloc'd, so the loop uses the global (not generic) space.
__global__ void __launch_bounds__(kThreads, 4)
fake_quant_multi_kernel(const DfqQuantDesc *__restrict__ descs, int n_desc, const RangeSlot *__restrict__ ranges) {
    // chunks are taken back to front: the range pass streamed the tensors front to back, so the last ~100 MB it read
    // are still in the 126 MB L2 when the first blocks of this pass ask for them
    const int64_t vb = (int64_t)gridDim.x - 1 - blockIdx.x;
    const int t = find_desc(descs, n_desc, vb);
    const DfqQuantDesc &d = descs[t];
    BlockWork w = decode_work1(d, vb - d.first_block);
    // chunk mode: the block walks [begin, end) of one slot; short rows: each warp walks its own row
    int64_t slot = d.range_slot + w.row, b = w.begin, e = w.end;
    int tid = threadIdx.x, nthr = kThreads;
    if (w.begin < 0) {
        int64_t r = w.row + (threadIdx.x >> 5);
        if (r >= d.rows) return;
        tid = threadIdx.x & 31; nthr = 32;
        slot = d.range_slot + r;
        b = r * w.cols; e = b + w.cols;
    }
    const int cb = d.bits <= 8 ? 1 : 2;
    const bool clamp_out = d.flags & DFQ_Q_CLAMP_OUT;
    const float *xg = reinterpret_cast<const float *>(__cvta_generic_to_global(d.x)) + b;
    float *yg = reinterpret_cast<float *>(__cvta_generic_to_global(d.y)) + b;
    if (w.begin >= 0 && e - b == kChunk && (((uintptr_t)xg | (uintptr_t)yg) & 15) == 0) {
        // full aligned chunk: the four data loads leave BEFORE the range slot is fetched.  With the slot first, every
        // block began with two dependent table loads (descriptor -> slot, ~1 us) and nothing in flight, a quarter of
        // its lifetime at 4 blocks per SM: 405 us against the 313 us of the single-tensor kernel on 1 GiB.
        const float4 *p4 = reinterpret_cast<const float4 *>(xg);
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = p4[threadIdx.x + u * kThreads];
        QuantConsts c = consts_from_slot(ranges[slot], d.bits, d.flags, d.clip_lo, d.clip_hi);
        uint8_t *codes = d.codes ? reinterpret_cast<uint8_t *>(__cvta_generic_to_global(d.codes)) + b * cb : nullptr;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int k = threadIdx.x + u * kThreads;
            apply_vec4(v[u], reinterpret_cast<float4 *>(yg) + k, codes ? codes + 4 * k * cb : nullptr, cb, c, clamp_out);
        }
        return;
    }
    QuantConsts c = consts_from_slot(ranges[slot], d.bits, d.flags, d.clip_lo, d.clip_hi);
    apply_range_w<0>(d.x, d.y, d.codes, cb, b, e, c, clamp_out, tid, nthr);
}

// Single-tensor elementwise pass with host-supplied constants (QuantMeasure / activations).
template <typename CodeT>
__global__ void __launch_bounds__(kThreads)
fake_quant_apply_kernel(const float *__restrict__ x, float *__restrict__ y, CodeT *__restrict__ codes,
                        int64_t n, DfqQuantParams p) {
    QuantConsts c = make_consts(p, 0.f, 0.f);
    int64_t begin = (int64_t)blockIdx.x * kChunk;
    int64_t end = min(begin + (int64_t)kChunk, n);
    apply_range<CodeT>(x, y, codes, begin, end, c, false, threadIdx.x, kThreads);
}

// codes -> fp32 with the two rounded operations that end the fake-quant chain (utils/quantize.py:76: q * scale, then
// + min), so a tensor shipped as codes + {min, max, scale} reproduces the sender's dequantised weights bit for bit.
// slots: one RangeSlot per row (per-channel) or one for the tensor; asymmetric grids (code = q - qmin = q).
template <typename CodeT>
__global__ void __launch_bounds__(kThreads)
dequant_codes_kernel(const CodeT *__restrict__ codes, float *__restrict__ y, int64_t n, int64_t cols,
                     const RangeSlot *__restrict__ slots, int per_row) {
    const int64_t begin = (int64_t)blockIdx.x * kChunk, end = min(begin + (int64_t)kChunk, n);
    for (int64_t k = begin + threadIdx.x; k < end; k += kThreads) {
        const RangeSlot r = slots[per_row ? k / cols : 0];
        y[k] = __fadd_rn(__fmul_rn((float)codes[k], r.scale), r.mn);
    }
}

// Multi-tensor form for the receiving side of the layer-sharded pass (dist.py): one launch dequantises every layer this
// rank does not own.  `codes` and `slot` may point into a PEER GPU's memory (NVLink P2P through a symmetric-memory
// buffer): the codes are then read straight from their owner -- 1 byte per element over the link, no staging copy, no
// all-gather -- and the fp32 values are written locally.  A block converts 16 K elements, 16 codes per thread-load.
constexpr int kDeqBlock = 16384;

__global__ void __launch_bounds__(kThreads)
dequant_multi_kernel(const DfqDequantDesc *__restrict__ descs, int n_desc) {
    int lo = 0, hi = n_desc - 1;
    while (lo < hi) {
        int mid = (lo + hi + 1) >> 1;
        if (descs[mid].first_block <= (int64_t)blockIdx.x) lo = mid; else hi = mid - 1;
    }
    const DfqDequantDesc d = descs[lo];
    const RangeSlot r = *reinterpret_cast<const RangeSlot *>(d.slot);
    const int64_t begin = ((int64_t)blockIdx.x - d.first_block) * kDeqBlock, end = min(begin + (int64_t)kDeqBlock, d.n);
    const uint8_t *c = reinterpret_cast<const uint8_t *>(d.codes);
    if (((((uintptr_t)c) | ((uintptr_t)d.y)) & 15) == 0) {
        // all four 16-byte code words of this thread are requested before the first conversion: over NVLink a load
        // takes microseconds, and one load in flight per thread left the link at a quarter of its bandwidth
        constexpr int U = kDeqBlock / (16 * kThreads);
        uint4 w[U];
        int64_t ks[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            ks[u] = begin + 16 * ((int64_t)threadIdx.x + (int64_t)u * kThreads);
            w[u] = (ks[u] + 16 <= end) ? __ldcs(reinterpret_cast<const uint4 *>(c + ks[u])) : make_uint4(0u, 0u, 0u, 0u);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t k = ks[u];
            if (k + 16 <= end) {
                const uint32_t ww[4] = {w[u].x, w[u].y, w[u].z, w[u].w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    float4 o;
                    o.x = __fadd_rn(__fmul_rn((float)(ww[j] & 0xffu), r.scale), r.mn);
                    o.y = __fadd_rn(__fmul_rn((float)((ww[j] >> 8) & 0xffu), r.scale), r.mn);
                    o.z = __fadd_rn(__fmul_rn((float)((ww[j] >> 16) & 0xffu), r.scale), r.mn);
                    o.w = __fadd_rn(__fmul_rn((float)(ww[j] >> 24), r.scale), r.mn);
                    __stcs(reinterpret_cast<float4 *>(d.y + k + 4 * j), o);
                }
            } else if (k < end) {
                for (int64_t e = k; e < end; ++e) d.y[e] = __fadd_rn(__fmul_rn((float)c[e], r.scale), r.mn);
            }
        }
    } else {
        for (int64_t k = begin + threadIdx.x; k < end; k += kThreads)
            d.y[k] = __fadd_rn(__fmul_rn((float)c[k], r.scale), r.mn);
    }
}

}  // namespace dfq

using namespace dfq;

extern "C" int dfq_dequant_plan_host(DfqDequantDesc *d, int n, int64_t *total_blocks) {
    int64_t b = 0;
    for (int i = 0; i < n; ++i) {
        if (d[i].n <= 0 || !d[i].codes || !d[i].y || !d[i].slot) return (int)cudaErrorInvalidValue;
        d[i].first_block = b;
        b += (d[i].n + kDeqBlock - 1) / kDeqBlock;
    }
    *total_blocks = b;
    return b > 0x7fffffffll ? (int)cudaErrorInvalidValue : 0;
}

extern "C" int dfq_dequant_multi_f32(const DfqDequantDesc *descs_dev, int n_desc, int64_t total_blocks, void *stream) {
    if (n_desc <= 0 || total_blocks <= 0) return 0;
    dequant_multi_kernel<<<(unsigned)total_blocks, kThreads, 0, (cudaStream_t)stream>>>(descs_dev, n_desc);
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_dequant_codes_f32(const void *codes, int code_bytes, float *y, int64_t n, int64_t rows,
                                     const float *slots_dev, void *stream) {
    if (n <= 0) return 0;
    if (rows < 1 || n % rows || (code_bytes != 1 && code_bytes != 2)) return (int)cudaErrorInvalidValue;
    const int64_t blocks = (n + kChunk - 1) / kChunk;
    if (blocks > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    cudaStream_t st = (cudaStream_t)stream;
    if (code_bytes == 1)
        dequant_codes_kernel<uint8_t><<<(unsigned)blocks, kThreads, 0, st>>>((const uint8_t *)codes, y, n, n / rows,
                                                                          (const RangeSlot *)slots_dev, rows > 1);
    else
        dequant_codes_kernel<uint16_t><<<(unsigned)blocks, kThreads, 0, st>>>((const uint16_t *)codes, y, n, n / rows,
                                                                           (const RangeSlot *)slots_dev, rows > 1);
    g_launches += 1;
    return (int)cudaGetLastError();
}

static inline int64_t desc_blocks(const DfqQuantDesc &d) {
    // virtual chunks of the range / apply passes
    if (!(d.flags & DFQ_Q_PER_CHANNEL)) return (d.n + kChunk - 1) / kChunk;
    int64_t cols = d.n / d.rows;
    if (cols > kSmallRow) return d.rows * ((cols + kChunk - 1) / kChunk);
    return (d.rows + kWarps - 1) / kWarps;
}

extern "C" void dfq_quant_params_host(double mn, double mx, int bits, int flags, DfqQuantParams *out) {
    *out = make_quant_params(mn, mx, bits, flags);
}

extern "C" int dfq_quant_plan_host(DfqQuantDesc *d, int n, int64_t *total_blocks, int64_t *total_slots) {
    int64_t b = 0, s = 0;
    for (int i = 0; i < n; ++i) {
        if (d[i].rows < 1) d[i].rows = 1;
        if (!(d[i].flags & DFQ_Q_PER_CHANNEL)) d[i].rows = 1;
        if (d[i].n <= 0 || d[i].n % d[i].rows) return (int)cudaErrorInvalidValue;
        d[i].first_block = b;
        d[i].range_slot = s;
        b += desc_blocks(d[i]);
        s += d[i].rows;
    }
    *total_blocks = b;
    *total_slots = s;
    return b > 0x7fffffffll ? (int)cudaErrorInvalidValue : 0;
}

extern "C" int dfq_fake_quant_multi_f32(const DfqQuantDesc *descs_dev, int n_desc, int64_t total_blocks,
                                        int64_t total_slots, float *ranges_dev, void *stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (n_desc <= 0 || total_blocks <= 0) return 0;
    cudaError_t e = cudaMemsetAsync(ranges_dev, 0, sizeof(float) * 4 * total_slots, st);
    if (e != cudaSuccess) return (int)e;
    minmax_multi_kernel<<<reduce_grid(minmax_multi_kernel, total_blocks), kThreads, 0, st>>>(descs_dev, n_desc, total_blocks,
                                                                        (unsigned *)ranges_dev);
    decode_slots_kernel<<<(unsigned)((total_slots + 255) / 256), 256, 0, st>>>(descs_dev, n_desc, (unsigned *)ranges_dev,
                                                                               total_slots);
    fake_quant_multi_kernel<<<(unsigned)total_blocks, kThreads, 0, st>>>(descs_dev, n_desc,
                                                                         (const RangeSlot *)ranges_dev);
    g_launches += 3;
    return (int)cudaGetLastError();
}

extern "C" int dfq_fake_quant_apply_f32(const float *x, float *y, void *codes, int64_t n, int bits,
                                        const DfqQuantParams *p, void *stream) {
    if (n <= 0) return 0;
    int64_t blocks = (n + kChunk - 1) / kChunk;
    if (blocks > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    cudaStream_t st = (cudaStream_t)stream;
    if (bits <= 8) fake_quant_apply_kernel<uint8_t><<<(unsigned)blocks, kThreads, 0, st>>>(x, y, (uint8_t *)codes, n, *p);
    else fake_quant_apply_kernel<uint16_t><<<(unsigned)blocks, kThreads, 0, st>>>(x, y, (uint16_t *)codes, n, *p);
    g_launches += 1;
    return (int)cudaGetLastError();
}

static int minmax_single(const float *x, int64_t n, int64_t rows, float *out, void *stream) {
    if (n <= 0 || rows <= 0 || n % rows) return (int)cudaErrorInvalidValue;
    DfqQuantDesc d = {};
    d.x = x; d.n = n; d.rows = rows; d.flags = rows > 1 ? DFQ_Q_PER_CHANNEL : 0;
    int64_t blocks = desc_blocks(d);
    if (blocks > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(out, 0, sizeof(float) * 2 * rows, st);
    if (e != cudaSuccess) return (int)e;
    minmax_single_kernel<<<reduce_grid(minmax_single_kernel, blocks), kThreads, 0, st>>>(d, blocks, (unsigned *)out);
    decode_ranges_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, st>>>((unsigned *)out, rows);
    g_launches += 2;
    return (int)cudaGetLastError();
}

extern "C" int dfq_minmax_f32(const float *x, int64_t n, float *out_minmax, void *stream) {
    return minmax_single(x, n, 1, out_minmax, stream);
}

extern "C" int dfq_minmax_rows_f32(const float *x, int64_t rows, int64_t cols, float *out_minmax, void *stream) {
    return minmax_single(x, rows * cols, rows, out_minmax, stream);
}
