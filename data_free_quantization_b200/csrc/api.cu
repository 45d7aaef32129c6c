// Library-level entry points of libdfq_b200.so.
#include "common.cuh"

namespace dfq {
long long g_launches = 0;
int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (!cached[dev]) {
        int n = 0;
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        cached[dev] = n > 0 ? n : 148;
    }
    return cached[dev];
}
int resident_blocks_per_sm(const void *kernel, int threads, size_t smem) {
    struct Entry { const void *k; int threads; size_t smem; int occ; };
    static Entry cache[64];
    static int n = 0;
    for (int i = 0; i < n; ++i)
        if (cache[i].k == kernel && cache[i].threads == threads && cache[i].smem == smem) return cache[i].occ;
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, threads, smem) != cudaSuccess || occ < 1) occ = 1;
    if (n < 64) cache[n++] = {kernel, threads, smem, occ};
    return occ;
}
}  // namespace dfq

extern "C" int dfq_abi_version(void) { return DFQ_ABI_VERSION; }
extern "C" const char *dfq_error_string(int err) { return cudaGetErrorString((cudaError_t)err); }
extern "C" int64_t dfq_launch_count(void) { return dfq::g_launches; }
