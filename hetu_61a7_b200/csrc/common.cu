#include "common.cuh"
#include <cstdarg>
#include <cstring>

static thread_local char g_err[512] = "";

void hetu_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

HETU_API const char* hetu_last_error() { return g_err; }

HETU_API int hetu_version() { return 100; }

__global__ void increment_i64_kernel(long long* p) { *p += 1; }

// bump a device-resident step counter (dropout seed / Adam step) - graph capturable
HETU_API int hetu_increment_i64(void* p, cudaStream_t s) {
    increment_i64_kernel<<<1, 1, 0, s>>>((long long*)p);
    HETU_CHECK_LAUNCH("increment_i64");
    return 0;
}
