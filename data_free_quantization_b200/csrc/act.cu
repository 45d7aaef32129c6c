// Activation-side pieces of the quantised eval forward (kernel e, part 1): QuantMeasure with the range kept on
// the device -- utils/quantize.py:94-126.  No host round trip per layer: the per-sample extrema, their mean over
// the batch, the running-range update (Python max()/min() semantics, :110-111) and the scale all stay in HBM, so
// a whole forward can be captured in a CUDA graph.
#include <cstdio>
#include "common.cuh"
#include "apply.cuh"

namespace dfq {

// One block: mean over samples of the per-sample (min, max) (fp64 accumulate), running-range update, scale.
//   stats   : [n_samples, 2] from minmax_single_kernel (may be NULL when !update_stat)
//   running : {running_min, running_max} device pointers (QuantMeasure buffers, updated in place)
//   slot    : RangeSlot {min, max, scale, 0} consumed by the apply kernel
__global__ void measure_update_kernel(const float *__restrict__ stats, int64_t n_samples, float *running_min,
                                      float *running_max, RangeSlot *slot, int bits, int flags, int update_stat,
                                      int encoded) {
    __shared__ double smin[kWarps], smax[kWarps];
    float rmin = *running_min, rmax = *running_max;
    if (update_stat) {                                       // block-uniform
        double lo = 0.0, hi = 0.0;
        for (int64_t i = threadIdx.x; i < n_samples; i += blockDim.x) {
            if (encoded) {       // {enc(max(-y)), enc(max(y))} accumulated by a producing conv's epilogue
                const unsigned *e = reinterpret_cast<const unsigned *>(stats);
                lo += (double)(-dec_ordered(e[2 * i]));
                hi += (double)dec_ordered(e[2 * i + 1]);
            } else {
                lo += (double)stats[2 * i];
                hi += (double)stats[2 * i + 1];
            }
        }
        lo = warp_sum(lo); hi = warp_sum(hi);
        if ((threadIdx.x & 31) == 0) { smin[threadIdx.x >> 5] = lo; smax[threadIdx.x >> 5] = hi; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double a = 0.0, b = 0.0;
            for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { a += smin[w]; b += smax[w]; }
            float mean_min = (float)(a / (double)n_samples), mean_max = (float)(b / (double)n_samples);
            // running_max = max(running_max, t): Python keeps the first argument unless t > it (NaN keeps it)
            if (mean_max > rmax) rmax = mean_max;
            if (mean_min < rmin) rmin = mean_min;
            *running_min = rmin;
            *running_max = rmax;
        }
    }
    if (threadIdx.x == 0) {
        // quantize(input, num_bits, float(min), float(max)): Python-float path, scale in double (:51-68)
        DfqQuantParams p = make_quant_params((double)rmin, (double)rmax, bits, flags);
        RangeSlot r = {rmin, rmax, p.scale, 0.0f};
        *slot = r;
    }
}

// Elementwise pass with the range read from the device slot (QuantMeasure.forward, :123).
template <typename CodeT>
__global__ void __launch_bounds__(kThreads)
fake_quant_slot_kernel(const float *__restrict__ x, float *__restrict__ y, CodeT *__restrict__ codes, int64_t n,
                       const RangeSlot *__restrict__ slot, int bits, int flags) {
    QuantConsts c = consts_from_slot(*slot, bits, flags, 0.f, 0.f);
    int64_t begin = (int64_t)blockIdx.x * kChunk;
    int64_t end = min(begin + (int64_t)kChunk, n);
    apply_range<CodeT>(x, y, codes, begin, end, c, false, threadIdx.x, kThreads);
}

}  // namespace dfq

using namespace dfq;

extern "C" int dfq_minmax_rows_f32(const float *x, int64_t rows, int64_t cols, float *out_minmax, void *stream);

extern "C" int dfq_measure_update(const float *x, int64_t n_samples, int64_t chw, float *running_min,
                                  float *running_max, float *slot_dev, int bits, int flags, int update_stat,
                                  float *stats_scratch, void *stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (update_stat) {
        if (n_samples <= 0 || chw <= 0 || !stats_scratch) return (int)cudaErrorInvalidValue;
        int e = dfq_minmax_rows_f32(x, n_samples, chw, stats_scratch, stream);      // per-sample extrema (:110-111)
        if (e) return e;
    }
    measure_update_kernel<<<1, 256, 0, st>>>(stats_scratch, n_samples, running_min, running_max,
                                             (RangeSlot *)slot_dev, bits, flags, update_stat, 0);
    g_launches += 1;
    return (int)cudaGetLastError();
}

// ---- allocation-free variant (no counterpart in the reference) ------
// Same work as dfq_measure_update, but the per-sample extrema live in a file-scope device array instead of a
// caller-provided scratch buffer, so a caller that may not allocate (CUDA graph capture, a tight serving loop) needs
// nothing but its inputs and outputs.  Every GPU that loads the library holds the 32 KB until exit.  All calls share
// the array: calls that overlap on the GPU (different streams) race on it; calls on one stream are ordered.
static __device__ __align__(256) float g_measure_scratch[2 * DFQ_MEASURE_NOALLOC_MAX_SAMPLES];

// The symbol's name is only a host-side handle; its device address is looked up once per device, with that device
// current (cudaGetSymbolAddress allocates nothing: the array is part of the loaded module).
static int measure_scratch_ptr(float **out) {
    static float *cached[64] = {nullptr};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
    if (!cached[dev]) {
        float *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_measure_scratch);
        if (e != cudaSuccess) return (int)e;
        cached[dev] = p;
    }
    *out = cached[dev];
    return 0;
}

extern "C" int dfq_measure_update_noalloc(const float *x, int64_t n_samples, int64_t chw, float *running_min,
                                          float *running_max, float *slot_dev, int bits, int flags, int update_stat,
                                          void *stream) {
    if (update_stat && n_samples > DFQ_MEASURE_NOALLOC_MAX_SAMPLES) {
        fprintf(stderr, "dfq_measure_update_noalloc: n_samples = %lld exceeds the built-in scratch of %d samples; "
                        "use dfq_measure_update with a caller-provided stats_scratch\n",
                (long long)n_samples, DFQ_MEASURE_NOALLOC_MAX_SAMPLES);
        return (int)cudaErrorInvalidValue;
    }
    float *scratch = nullptr;
    int e = measure_scratch_ptr(&scratch);
    if (e) return e;
    // dfq_minmax_rows_f32 clears the 2*n_samples words it accumulates into (cudaMemsetAsync on `stream`) before its
    // atomics run, so what the previous call left in the array is never read
    return dfq_measure_update(x, n_samples, chw, running_min, running_max, slot_dev, bits, flags, update_stat, scratch,
                              stream);
}

extern "C" int dfq_measure_from_stats(const void *stats, int encoded, int64_t n_samples, float *running_min,
                                      float *running_max, float *slot_dev, int bits, int flags, int update_stat,
                                      void *stream) {
    if (update_stat && (!stats || n_samples <= 0)) return (int)cudaErrorInvalidValue;
    measure_update_kernel<<<1, 256, 0, (cudaStream_t)stream>>>((const float *)stats, n_samples, running_min,
                                                               running_max, (RangeSlot *)slot_dev, bits, flags,
                                                               update_stat, encoded);
    g_launches += 1;
    return (int)cudaGetLastError();
}

extern "C" int dfq_fake_quant_apply_slot_f32(const float *x, float *y, void *codes, int64_t n, int bits, int flags,
                                             const float *slot_dev, void *stream) {
    if (n <= 0) return 0;
    int64_t blocks = (n + kChunk - 1) / kChunk;
    if (blocks > 0x7fffffffll) return (int)cudaErrorInvalidValue;
    cudaStream_t st = (cudaStream_t)stream;
    if (bits <= 8)
        fake_quant_slot_kernel<uint8_t><<<(unsigned)blocks, kThreads, 0, st>>>(x, y, (uint8_t *)codes, n,
                                                                               (const RangeSlot *)slot_dev, bits, flags);
    else
        fake_quant_slot_kernel<uint16_t><<<(unsigned)blocks, kThreads, 0, st>>>(x, y, (uint16_t *)codes, n,
                                                                                (const RangeSlot *)slot_dev, bits, flags);
    g_launches += 1;
    return (int)cudaGetLastError();
}
