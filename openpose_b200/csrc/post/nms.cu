// NMS: 3x3 peak detection + 7x7 score-weighted centroid, first maxPeaks peaks per channel in raster order.
//
// Replaces op::nmsCpu / op::nmsGpu (reference: src/openpose/net/nmsBase.cpp:7-170, nmsBase.cu:51-90,165-351).
// The reference GPU path is three passes (0/1 int map over the whole blob -> thrust scan of 25*H*W ints ->
// edge-detect + centroid): ~6x the compulsory traffic. Here ONE launch reads each heat-map value once:
//   * every warp streams a contiguous pixel segment of one channel with 128-bit loads and, only for the rare
//     values above the threshold, looks at the 8 neighbours (L1/L2 hits);
//   * peaks are compacted in raster order with ballot/shuffle prefix sums (warp) and a short concatenation
//     (block), keeping at most maxPeaks per segment -- later peaks can never be output;
//   * the last block of a channel to finish (atomic ticket) concatenates the band lists and runs the
//     centroid refinement, one thread per peak, with the reference's summation order and no FMA contraction.
// HBM roofline: algorithmic bytes = 4*parts*H*W (+ peaks), see DESIGN.md.
#include "../common.cuh"

namespace opb
{
    constexpr int kNmsThreads = 256;
    constexpr int kNmsWarps = kNmsThreads / 32;
    constexpr int kNmsMaxBands = 64;
    constexpr int kChunk = 128; // pixels per warp iteration (32 lanes x float4)

    // Peak predicate, CPU semantics (reference: nmsBase.cpp:17-67).
    __device__ __forceinline__ bool isPeakCpu(const float* __restrict__ s, int w, int h, float thr, int x, int y,
                                              float v)
    {
        if (1 < x && x < w - 2 && 1 < y && y < h - 2)
        {
            const float* r0 = s + (size_t)(y - 1) * w + x;
            const float* r1 = r0 + w;
            const float* r2 = r1 + w;
            return v > r0[-1] && v > r0[0] && v > r0[1] && v > r1[-1] && v > r1[1] && v > r2[-1] && v > r2[0]
                   && v > r2[1];
        }
        if (x == 1 || x == w - 2 || y == 1 || y == h - 2)
        {
            bool ok = true;
#pragma unroll
            for (int dy = -1; dy <= 1; dy++)
#pragma unroll
                for (int dx = -1; dx <= 1; dx++)
                {
                    if (dx == 0 && dy == 0)
                        continue;
                    const int xx = x + dx, yy = y + dy;
                    const float nb = (xx >= 0 && xx < w && yy >= 0 && yy < h) ? s[(size_t)yy * w + xx] : thr;
                    ok = ok && (v >= nb);
                }
            return ok;
        }
        return false;
    }

    // Peak predicate, CUDA semantics (reference: nmsBase.cu:64-89): strict '>' everywhere, outer border never.
    __device__ __forceinline__ bool isPeakCuda(const float* __restrict__ s, int w, int h, int x, int y, float v)
    {
        if (!(0 < x && x < w - 1 && 0 < y && y < h - 1))
            return false;
        const float* r0 = s + (size_t)(y - 1) * w + x;
        const float* r1 = r0 + w;
        const float* r2 = r1 + w;
        return v > r0[-1] && v > r0[0] && v > r0[1] && v > r1[-1] && v > r1[1] && v > r2[-1] && v > r2[0]
               && v > r2[1];
    }

    // 7x7 centroid (reference: nmsBase.cpp:71-107): row-major accumulation, int->float multiply, no FMA.
    __device__ __forceinline__ void refinePeak(float* __restrict__ out, const float* __restrict__ s, int px, int py,
                                               int w, int h, float offx, float offy)
    {
        float xAcc = 0.f, yAcc = 0.f, sAcc = 0.f;
        for (int dy = -3; dy <= 3; dy++)
        {
            const int y = py + dy;
            if (y < 0 || y >= h)
                continue;
            const float fy = (float)y;
#pragma unroll
            for (int dx = -3; dx <= 3; dx++)
            {
                const int x = px + dx;
                if (x < 0 || x >= w)
                    continue;
                const float sc = s[(size_t)y * w + x];
                if (sc > 0)
                {
                    xAcc = add(xAcc, mul((float)x, sc));
                    yAcc = add(yAcc, mul(fy, sc));
                    sAcc = add(sAcc, sc);
                }
            }
        }
        out[0] = add(divf(xAcc, sAcc), offx);
        out[1] = add(divf(yAcc, sAcc), offy);
        out[2] = s[(size_t)py * w + px];
    }

    struct NmsArgs
    {
        float* peaks;      // (CI, maxPeaks+1, 3)
        int* tickets;      // (CI)
        int* bandCounts;   // (CI, bands)
        int* bandLists;    // (CI, bands, maxPeaks)
        const float* heat; // (N, Ch, H, W)
        float thr, offx, offy;
        int parts, heatChannels, h, w, hw, maxPeaks, bands, segLen;
    };

    template <int MODE, bool VEC>
    __global__ void __launch_bounds__(kNmsThreads) nmsKernel(const NmsArgs a)
    {
        extern __shared__ int smem[];      // [kNmsWarps][maxPeaks] pixel indices, then scratch
        int* warpLists = smem;
        __shared__ int warpCounts[kNmsWarps];
        __shared__ int bandPrefix[kNmsMaxBands + 1];
        __shared__ int isLast;

        const int ci = blockIdx.y;                       // channel instance = n*parts + c
        const int n = ci / a.parts, c = ci - n * a.parts;
        const float* __restrict__ plane = a.heat + ((size_t)n * a.heatChannels + c) * a.hw;
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        const int band = blockIdx.x;
        const int segBegin = (band * kNmsWarps + warp) * a.segLen;
        const int segEnd = min(segBegin + a.segLen, a.hw);
        int* myList = warpLists + warp * a.maxPeaks;
        int count = 0;                                   // warp-uniform

        // kUnroll independent 128-bit loads per lane are issued before any of them is examined: the peak count
        // that steers the loop depends on the loaded values, so without this only one load per warp is in flight
        // and the scan is latency bound (measured 1.2 TB/s); sub-chunks are then processed in raster order.
        constexpr int kUnroll = 4;
        for (int base = segBegin; base < segEnd && count < a.maxPeaks; base += kChunk * kUnroll)
        {
            float v[kUnroll][4];
#pragma unroll
            for (int u = 0; u < kUnroll; u++)
            {
                const int pix = base + u * kChunk + lane * 4;
                if (VEC)
                {
                    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (pix < segEnd)
                        q = __ldg(reinterpret_cast<const float4*>(plane + pix));
                    v[u][0] = q.x; v[u][1] = q.y; v[u][2] = q.z; v[u][3] = q.w;
                }
                else
                {
#pragma unroll
                    for (int e = 0; e < 4; e++)
                        v[u][e] = (pix + e < segEnd) ? __ldg(plane + pix + e) : 0.f;
                }
            }
#pragma unroll
            for (int u = 0; u < kUnroll; u++)
            {
                const int pix = base + u * kChunk + lane * 4;
                unsigned mask = 0;
#pragma unroll
                for (int e = 0; e < 4; e++)
                {
                    // thr >= 0, so the zero fill of out-of-range lanes can never pass.
                    if (v[u][e] > a.thr && pix + e < segEnd)
                    {
                        const int y = (pix + e) / a.w, x = (pix + e) - y * a.w;
                        const bool pk = MODE == OPB200_SEMANTICS_CPU
                                            ? isPeakCpu(plane, a.w, a.h, a.thr, x, y, v[u][e])
                                            : isPeakCuda(plane, a.w, a.h, x, y, v[u][e]);
                        mask |= pk ? (1u << e) : 0u;
                    }
                }
                if (__ballot_sync(0xffffffffu, mask != 0) != 0)
                {
                    // raster-order compaction: lane-major, then element order inside the lane
                    const int mine = __popc(mask);
                    int incl = mine;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1)
                    {
                        const int t = __shfl_up_sync(0xffffffffu, incl, d);
                        if (lane >= d)
                            incl += t;
                    }
                    int slot = count + incl - mine;
#pragma unroll
                    for (int e = 0; e < 4; e++)
                        if (mask & (1u << e))
                        {
                            if (slot < a.maxPeaks)
                                myList[slot] = pix + e;
                            slot++;
                        }
                    count += __shfl_sync(0xffffffffu, incl, 31);
                }
            }
        }
        if (lane == 0)
            warpCounts[warp] = min(count, a.maxPeaks);
        __syncthreads();

        // concatenate the warp lists of this band (still raster order), clamp to maxPeaks
        int* gList = a.bandLists + ((size_t)ci * a.bands + band) * a.maxPeaks;
        int total = 0;
#pragma unroll
        for (int wq = 0; wq < kNmsWarps; wq++)
        {
            const int cnt = warpCounts[wq];
            for (int i = threadIdx.x; i < cnt && total + i < a.maxPeaks; i += kNmsThreads)
                gList[total + i] = warpLists[wq * a.maxPeaks + i];
            total += cnt;
        }
        total = min(total, a.maxPeaks);
        if (threadIdx.x == 0)
            a.bandCounts[ci * a.bands + band] = total;

        // last block of this channel finalises
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0)
        {
            const int t = atomicAdd(&a.tickets[ci], 1);
            isLast = (t == a.bands - 1);
        }
        __syncthreads();
        if (!isLast)
            return;
        __threadfence();
        if (threadIdx.x == 0)
        {
            int run = 0;
            for (int b = 0; b < a.bands; b++)
            {
                bandPrefix[b] = run;
                run += __ldcg(&a.bandCounts[ci * a.bands + b]);
            }
            bandPrefix[a.bands] = run;
            a.tickets[ci] = 0; // self-cleaning for the next launch
        }
        __syncthreads();
        const int nPeaks = min(bandPrefix[a.bands], a.maxPeaks);
        float* out = a.peaks + (size_t)ci * (a.maxPeaks + 1) * 3;
        for (int k = threadIdx.x; k < nPeaks; k += kNmsThreads)
        {
            int b = 0;
            while (bandPrefix[b + 1] <= k)
                b++;
            const int pix = __ldcg(&a.bandLists[((size_t)ci * a.bands + b) * a.maxPeaks + (k - bandPrefix[b])]);
            const int y = pix / a.w, x = pix - y * a.w;
            refinePeak(out + 3 * (k + 1), plane, x, y, a.w, a.h, a.offx, a.offy);
        }
        if (threadIdx.x == 0)
            out[0] = (float)nPeaks;
    }

    static int chooseBands(int channelInstances, int hw)
    {
        const int target = kNumSMsB200 * 8; // 8 resident 256-thread blocks per SM
        int bands = (target + channelInstances - 1) / channelInstances;
        const int maxByWork = (hw + kNmsWarps * kChunk - 1) / (kNmsWarps * kChunk);
        bands = bands < 1 ? 1 : bands;
        bands = bands > kNmsMaxBands ? kNmsMaxBands : bands;
        bands = bands > maxByWork ? maxByWork : bands;
        return bands < 1 ? 1 : bands;
    }
}

extern "C" size_t opb200_nms_scratch_bytes(const int peaks_shape[4], const int heat_shape[4])
{
    (void)heat_shape;
    const size_t ci = (size_t)peaks_shape[0] * peaks_shape[1];
    const size_t maxPeaks = (size_t)peaks_shape[2] - 1;
    return sizeof(int) * (ci + ci * opb::kNmsMaxBands * (1 + maxPeaks));
}

extern "C" opb200_status opb200_nms(float* peaks, void* scratch, const float* heat, float threshold,
                                    float offset_x, float offset_y, const int peaks_shape[4],
                                    const int heat_shape[4], int mode, void* stream)
{
    using namespace opb;
    OPB_REQUIRE(peaks && scratch && heat && peaks_shape && heat_shape, "nms: null argument");
    // reference: nmsBase.cpp:121-122
    OPB_REQUIRE(threshold >= 0.f && threshold <= 1.f, "threshold value invalid.");
    OPB_REQUIRE(mode == OPB200_SEMANTICS_CPU || mode == OPB200_SEMANTICS_CUDA, "nms: unknown semantics mode");
    const int n = heat_shape[0], ch = heat_shape[1], h = heat_shape[2], w = heat_shape[3];
    const int parts = peaks_shape[1], maxPeaks = peaks_shape[2] - 1;
    OPB_REQUIRE(n > 0 && ch > 0 && h > 0 && w > 0, "nms: sourceSize cannot be empty.");
    OPB_REQUIRE(peaks_shape[0] == n && parts > 0 && parts <= ch && maxPeaks > 0 && peaks_shape[3] == 3,
                "nms: targetSize inconsistent with sourceSize");
    OPB_REQUIRE((long long)h * w < (1ll << 31), "nms: plane too large");
    const int ci = n * parts;
    OPB_REQUIRE(ci <= 65535, "nms: too many channel instances for one launch");

    NmsArgs a;
    a.hw = h * w;
    a.bands = chooseBands(ci, a.hw);
    const int segs = a.bands * kNmsWarps;
    a.segLen = ((a.hw + segs - 1) / segs + kChunk - 1) / kChunk * kChunk;
    a.peaks = peaks;
    a.tickets = static_cast<int*>(scratch);
    a.bandCounts = a.tickets + ci;
    a.bandLists = a.bandCounts + (size_t)ci * kNmsMaxBands;
    a.heat = heat;
    a.thr = threshold;
    a.offx = offset_x;
    a.offy = offset_y;
    a.parts = parts;
    a.heatChannels = ch;
    a.h = h;
    a.w = w;
    a.maxPeaks = maxPeaks;

    cudaStream_t st = static_cast<cudaStream_t>(stream);
    OPB_CUDA(cudaMemsetAsync(a.tickets, 0, sizeof(int) * ci, st));
    const bool vec = (a.hw % 4 == 0) && ((reinterpret_cast<uintptr_t>(heat) & 15u) == 0);
    const dim3 grid(a.bands, ci);
    const size_t smem = sizeof(int) * kNmsWarps * maxPeaks;
    if (mode == OPB200_SEMANTICS_CPU)
    {
        if (vec)
            nmsKernel<OPB200_SEMANTICS_CPU, true><<<grid, kNmsThreads, smem, st>>>(a);
        else
            nmsKernel<OPB200_SEMANTICS_CPU, false><<<grid, kNmsThreads, smem, st>>>(a);
    }
    else
    {
        if (vec)
            nmsKernel<OPB200_SEMANTICS_CUDA, true><<<grid, kNmsThreads, smem, st>>>(a);
        else
            nmsKernel<OPB200_SEMANTICS_CUDA, false><<<grid, kNmsThreads, smem, st>>>(a);
    }
    OPB_LAUNCH_CHECK();
    return OPB200_OK;
}
