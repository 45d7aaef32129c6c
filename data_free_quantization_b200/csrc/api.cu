// Library-level entry points of libdfq_b200.so.
#include "common.cuh"

namespace dfq { long long g_launches = 0; }

extern "C" int dfq_abi_version(void) { return DFQ_ABI_VERSION; }
extern "C" const char *dfq_error_string(int err) { return cudaGetErrorString((cudaError_t)err); }
extern "C" int64_t dfq_launch_count(void) { return dfq::g_launches; }
