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
}  // namespace dfq

extern "C" int dfq_abi_version(void) { return DFQ_ABI_VERSION; }
extern "C" const char *dfq_error_string(int err) { return cudaGetErrorString((cudaError_t)err); }
extern "C" int64_t dfq_launch_count(void) { return dfq::g_launches; }
