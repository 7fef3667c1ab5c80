// maximum: per-channel argmax (x, y, value) of the hand/face heat-maps, first maximum in raster order.
// Replaces op::maximumCpu / op::maximumGpu (reference: src/openpose/net/maximumBase.cpp:8-57 via
// cv::minMaxLoc; src/openpose/net/maximumBase.cu:63-120 = one thrust::max_element + one <<<1,1>>> kernel
// + host sync PER CHANNEL). Here: one launch for all (frame, channel) planes. Each plane is cut into
// segments; every block reduces its segment with 128-bit loads to a 64-bit key
//     (order-preserving bits of the value) << 32 | ~pixelIndex
// so that an unsigned max picks the largest value and, among equal values, the smallest index; blocks merge
// with atomicMax and the last block of a plane (ticket) decodes the key into (x, y, value).
// HBM roofline: algorithmic bytes = 4*(C-1)*H*W per frame.
#include "../common.cuh"

namespace opb
{
    constexpr int kMaxThreads = 256;
    constexpr int kMaxSegs = 64;

    __device__ __forceinline__ unsigned orderedBits(float v)
    {
        if (v == 0.f)
            v = 0.f; // -0.0 compares equal to +0.0 in cv::minMaxLoc: canonicalise
        const unsigned u = __float_as_uint(v);
        return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    }
    __device__ __forceinline__ float unorderedBits(unsigned k)
    {
        return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
    }
    __device__ __forceinline__ unsigned long long makeKey(float v, int idx)
    {
        return ((unsigned long long)orderedBits(v) << 32) | (unsigned)(~(unsigned)idx);
    }

    struct MaxArgs
    {
        float* target;              // (planes, 3)
        unsigned long long* keys;   // (planes)
        int* tickets;               // (planes)
        const float* source;        // (N, C, H, W)
        int partsPerFrame, channels, hw, w, segs, segLen;
    };

    template <bool VEC>
    __global__ void __launch_bounds__(kMaxThreads) maximumKernel(const MaxArgs a)
    {
        const int plane = blockIdx.y; // n*(C-1) + part
        const int n = plane / a.partsPerFrame, part = plane - n * a.partsPerFrame;
        const float* __restrict__ s = a.source + ((size_t)n * a.channels + part) * a.hw;
        const int begin = blockIdx.x * a.segLen;
        const int end = min(begin + a.segLen, a.hw);
        unsigned long long best = 0ull; // below every real key (orderedBits > 0 for every non-NaN float)
        if (VEC)
        {
            for (int i = begin + threadIdx.x * 4; i < end; i += kMaxThreads * 4)
            {
                const float4 q = __ldg(reinterpret_cast<const float4*>(s + i));
                // keep the FIRST maximum of the four: strict '>' while walking forward
                float v = q.x;
                int idx = i;
                if (q.y > v) { v = q.y; idx = i + 1; }
                if (q.z > v) { v = q.z; idx = i + 2; }
                if (q.w > v) { v = q.w; idx = i + 3; }
                const unsigned long long k = makeKey(v, idx);
                best = k > best ? k : best;
            }
        }
        else
        {
            for (int i = begin + threadIdx.x; i < end; i += kMaxThreads)
            {
                const unsigned long long k = makeKey(__ldg(s + i), i);
                best = k > best ? k : best;
            }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1)
        {
            const unsigned long long o = __shfl_xor_sync(0xffffffffu, best, d);
            best = o > best ? o : best;
        }
        __shared__ unsigned long long warpBest[kMaxThreads / 32];
        __shared__ int isLast;
        if ((threadIdx.x & 31) == 0)
            warpBest[threadIdx.x >> 5] = best;
        __syncthreads();
        if (threadIdx.x == 0)
        {
            for (int wq = 1; wq < kMaxThreads / 32; wq++)
                best = warpBest[wq] > best ? warpBest[wq] : best;
            atomicMax(&a.keys[plane], best);
            __threadfence();
            const int t = atomicAdd(&a.tickets[plane], 1);
            isLast = (t == a.segs - 1);
            if (isLast)
            {
                __threadfence();
                const unsigned long long k = atomicMax(&a.keys[plane], 0ull); // coherent read
                const int idx = (int)(~(unsigned)(k & 0xffffffffu));
                float* out = a.target + (size_t)plane * 3;
                out[0] = (float)(idx % a.w);
                out[1] = (float)(idx / a.w);
                out[2] = unorderedBits((unsigned)(k >> 32));
                a.keys[plane] = 0ull;
                a.tickets[plane] = 0;
            }
        }
    }
}

extern "C" size_t opb200_maximum_scratch_bytes(const int target_shape[4], const int source_shape[4])
{
    (void)source_shape;
    const size_t planes = (size_t)target_shape[0] * target_shape[1] * target_shape[2];
    return planes * (sizeof(unsigned long long) + sizeof(int)) + 16;
}

extern "C" opb200_status opb200_maximum(float* target, void* scratch, const float* source,
                                        const int target_shape[4], const int source_shape[4], void* stream)
{
    using namespace opb;
    OPB_REQUIRE(target && scratch && source && target_shape && source_shape, "maximum: null argument");
    const int n = source_shape[0], c = source_shape[1], h = source_shape[2], w = source_shape[3];
    OPB_REQUIRE(n > 0 && c > 1 && h > 0 && w > 0, "maximum: empty source");
    // reference: maximumCaffe.cpp:66-70 -> top (N, 1, C-1, 3)
    OPB_REQUIRE(target_shape[0] == n && target_shape[1] == 1 && target_shape[2] == c - 1 && target_shape[3] == 3,
                "maximum: target must be (N, 1, C-1, 3)");
    OPB_REQUIRE((long long)h * w < (1ll << 31), "maximum: plane too large");
    const int planes = n * (c - 1);
    OPB_REQUIRE(planes <= 65535, "maximum: too many planes for one launch");

    MaxArgs a;
    a.target = target;
    a.keys = static_cast<unsigned long long*>(scratch);
    a.tickets = reinterpret_cast<int*>(a.keys + planes);
    a.source = source;
    a.partsPerFrame = c - 1;
    a.channels = c;
    a.hw = h * w;
    a.w = w;
    int segs = (kNumSMsB200 * 8 + planes - 1) / planes;
    const int maxByWork = (a.hw + kMaxThreads * 16 - 1) / (kMaxThreads * 16);
    segs = segs > kMaxSegs ? kMaxSegs : segs;
    segs = segs > maxByWork ? maxByWork : segs;
    segs = segs < 1 ? 1 : segs;
    a.segs = segs;
    a.segLen = ((a.hw + segs - 1) / segs + 3) / 4 * 4;

    cudaStream_t st = static_cast<cudaStream_t>(stream);
    OPB_CUDA(cudaMemsetAsync(scratch, 0, (size_t)planes * (sizeof(unsigned long long) + sizeof(int)), st));
    const bool vec = (a.hw % 4 == 0) && ((reinterpret_cast<uintptr_t>(source) & 15u) == 0);
    const dim3 grid(segs, planes);
    if (vec)
        maximumKernel<true><<<grid, kMaxThreads, 0, st>>>(a);
    else
        maximumKernel<false><<<grid, kMaxThreads, 0, st>>>(a);
    OPB_LAUNCH_CHECK();
    return OPB200_OK;
}
