// Error plumbing and library-level queries of the C ABI (include/opb200.h).
#include "common.cuh"
#include <cstring>

namespace opb
{
    static thread_local char gError[1024] = "";

    void setError(const char* fmt, ...)
    {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(gError, sizeof(gError), fmt, ap);
        va_end(ap);
    }
    const char* getError() { return gError; }
}

extern "C" const char* opb200_last_error(void) { return opb::getError(); }
extern "C" int opb200_version(void) { return 1; }

// reference: op::getCudaGpuNumber, src/openpose/gpu/cuda.cpp:41-61
extern "C" int opb200_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess)
    {
        cudaGetLastError();
        return 0;
    }
    return n;
}
