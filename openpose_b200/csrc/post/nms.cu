// NMS: 3x3 peak detection + 7x7 score-weighted centroid, first maxPeaks peaks per channel in raster order.
//
// Replaces op::nmsCpu / op::nmsGpu (reference: src/openpose/net/nmsBase.cpp:7-170, nmsBase.cu:51-90,165-351).
// The reference GPU path is three passes (0/1 int map over the whole blob -> thrust scan of 25*H*W ints ->
// edge-detect + centroid): ~6x the compulsory traffic. Here ONE launch reads each heat-map value once from HBM:
//   * persistent blocks (one per SM) each own an equal contiguous range of (channel, row tile) pairs; a producer
//     warp streams the tiles through a shared-memory ring with bulk async copies (UBLKCP + mbarrier), scanning
//     warps test them out of shared memory (warp-uniform fast path for pixel groups below the threshold, branch-free
//     4-neighbour pre-filter, exact reference predicate only for the survivors);
//   * peaks are compacted in raster order with ballot/shuffle prefix sums into per-warp lists; a merger warp
//     concatenates them per channel, off the scanning warps' critical path, keeping at most maxPeaks;
//   * the block that finishes a channel last (atomic ticket) runs the centroid refinement, one thread per peak,
//     with the reference's summation order and no FMA contraction.
// HBM roofline: algorithmic bytes = 4*parts*H*W (+ peaks), see DESIGN.md. Measured on B200 (16 frames, 25x368x656
// each): 83 us per launch = 4.7 TB/s, the rate of the same ring with the scan removed.
#include "../common.cuh"
#include "../net/sm100.cuh"
#include <algorithm>
#include <cstdlib>

namespace opb
{
    constexpr int kNmsMaxStages = 8;     // ring depth
    constexpr int kNmsWarpsPerStage = 3; // warps that share the scan of one tile (latency hiding needs ~24 warps/SM)
    constexpr int kNmsMaxWarps = kNmsMaxStages * kNmsWarpsPerStage;
    constexpr int kNmsFlushBands = 1;    // merged bands kept in shared memory between two output flushes
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

    // 7x7 centroid (reference: nmsBase.cpp:71-107): row-major accumulation, int->float multiply. CPU semantics: no
    // FMA (x86-64 build). CUDA semantics: xAcc += x*score / yAcc += y*score are FFMAs in the reference's compiled
    // writeResultKernel (nmsBase.cu:211-213 under nvcc's default -fmad=true), spelled out here.
    template <int MODE>
    __device__ __forceinline__ void refinePeak(float* __restrict__ out, const float* __restrict__ s, int px, int py,
                                               int w, int h, float offx, float offy)
    {
        // all 49 loads are issued before the (ordered, dependent) accumulation: one memory round trip, not 7
        float sc[7][7];
#pragma unroll
        for (int dy = -3; dy <= 3; dy++)
        {
            const int y = py + dy;
            const bool yIn = (y >= 0 && y < h);
            const float* __restrict__ row = s + (size_t)(yIn ? y : py) * w;
#pragma unroll
            for (int dx = -3; dx <= 3; dx++)
            {
                const int x = px + dx;
                sc[dy + 3][dx + 3] = (yIn && x >= 0 && x < w) ? row[x] : 0.f;   // 0 is skipped below, like outside
            }
        }
        float xAcc = 0.f, yAcc = 0.f, sAcc = 0.f;
#pragma unroll
        for (int dy = -3; dy <= 3; dy++)
        {
            const float fy = (float)(py + dy);
#pragma unroll
            for (int dx = -3; dx <= 3; dx++)
                if (sc[dy + 3][dx + 3] > 0)
                {
                    if (MODE == OPB200_SEMANTICS_CUDA)
                    {
                        xAcc = __fmaf_rn((float)(px + dx), sc[dy + 3][dx + 3], xAcc);
                        yAcc = __fmaf_rn(fy, sc[dy + 3][dx + 3], yAcc);
                    }
                    else
                    {
                        xAcc = add(xAcc, mul((float)(px + dx), sc[dy + 3][dx + 3]));
                        yAcc = add(yAcc, mul(fy, sc[dy + 3][dx + 3]));
                    }
                    sAcc = add(sAcc, sc[dy + 3][dx + 3]);
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
        int* bandCounts;   // (CI, kNmsMaxBands)
        int* bandLists;    // (CI, kNmsMaxBands, maxPeaks)
        const float* heat; // (N, Ch, H, W)
        float thr, offx, offy;
        int parts, heatChannels, h, w, hw, maxPeaks;
        int stages;        // ring depth
        int groups;        // warp groups; tile j of a block is scanned by group j % groups, out of stage j % stages
        int warpsPerTile;  // warps of a group (each scans an equal part of the tile); groups * warpsPerTile scanning warps
        int segLen;        // pixels per scanning warp within one tile (multiple of kChunk)
        int tileRows;      // rows per tile (the staged copy adds one halo row above and below)
        int tilesPerPlane; // ceil(h / tileRows)
        long long tiles;   // channel instances * tilesPerPlane
    };

    // Scans plane pixels [segBegin, segEnd) (whole rows) for peaks, appending their pixel indices to `list` in raster
    // order (at most maxPeaks kept); scan[p] must be readable for every pixel of the rows above/below too.
    // Returns the new warp-uniform count. kUnroll independent 128-bit loads per lane are issued before any of them
    // is examined (the loop is steered by loaded values, so without this one load per warp is in flight).
    template <int MODE, bool VEC>
    __device__ __forceinline__ int scanPixels(const NmsArgs& a, const float* __restrict__ scan, int segBegin, int segEnd,
                                              int* __restrict__ list, int count, int lane, float invW)
    {
        constexpr int kUnroll = 2;
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
                        q = *reinterpret_cast<const float4*>(scan + pix);
                    v[u][0] = q.x; v[u][1] = q.y; v[u][2] = q.z; v[u][3] = q.w;
                }
                else
                {
#pragma unroll
                    for (int e = 0; e < 4; e++)
                        v[u][e] = (pix + e < segEnd) ? scan[pix + e] : 0.f;
                }
            }
            {
                // cheapest, warp-uniform common path: nothing above the threshold in all kUnroll*128 pixels
                // (fmaxf drops NaNs, which can never be candidates either)
                float m = fmaxf(fmaxf(v[0][0], v[0][1]), fmaxf(v[0][2], v[0][3]));
#pragma unroll
                for (int u = 1; u < kUnroll; u++)
                    m = fmaxf(m, fmaxf(fmaxf(v[u][0], v[u][1]), fmaxf(v[u][2], v[u][3])));
                if (!__any_sync(0xffffffffu, m > a.thr))
                    continue;
            }
#pragma unroll
            for (int u = 0; u < kUnroll; u++)
            {
                const int pix = base + u * kChunk + lane * 4;
                // (thr >= 0, so the zero fill of out-of-range lanes can never pass)
                unsigned cand = (v[u][0] > a.thr ? 1u : 0u) | (v[u][1] > a.thr ? 2u : 0u) |
                                (v[u][2] > a.thr ? 4u : 0u) | (v[u][3] > a.thr ? 8u : 0u);
                if (__ballot_sync(0xffffffffu, cand != 0) == 0)
                    continue;
                unsigned mask = 0;
                float left = 0.f, right = 0.f;
                if (VEC)
                {
                    left = __shfl_up_sync(0xffffffffu, v[u][3], 1);      // pixel pix-1 (lanes > 0)
                    right = __shfl_down_sync(0xffffffffu, v[u][0], 1);   // pixel pix+4, or the zero fill (lanes < 31)
                }
                if (cand != 0)
                {
                    // row/column of the lane's first pixel: float reciprocal + one correction step (exact)
                    int y, x;
                    if (a.hw < (1 << 22))
                    {
                        y = (int)(((float)pix + 0.5f) * invW);
                        x = pix - y * a.w;
                        if (x < 0) { y--; x += a.w; }
                        else if (x >= a.w) { y++; x -= a.w; }
                    }
                    else
                    {
                        y = pix / a.w;
                        x = pix - y * a.w;
                    }
                    // Pre-filter, necessary under both semantics (>= also covers the strict tests): a peak is no
                    // smaller than its in-plane 4-neighbours. Inside a blob this leaves ~1 pixel for the exact
                    // 8-neighbour test (the zero fill past the segment end only lets more through).
                    unsigned keep = 0;
                    if (VEC)
                    {
                        // 4-neighbour pre-filter without branches (the lane's 4 pixels share a row: w % 4 == 0).
                        // Neighbours outside the plane are skipped (-inf); rows above/below come from the tile.
                        const float ninf = __int_as_float(0xff800000);
                        if (x == 0)
                            left = ninf;
                        else if (lane == 0)
                            left = scan[pix - 1];
                        if (x + 4 >= a.w)
                            right = ninf;
                        else if (lane == 31)
                            right = scan[pix + 4];
                        float4 up = make_float4(ninf, ninf, ninf, ninf), dn = up;
                        if (y > 0)
                            up = *reinterpret_cast<const float4*>(scan + pix - a.w);
                        if (y < a.h - 1)
                            dn = *reinterpret_cast<const float4*>(scan + pix + a.w);
                        const float v0 = v[u][0], v1 = v[u][1], v2 = v[u][2], v3 = v[u][3];
                        keep = ((v0 >= left && v0 >= v1 && v0 >= up.x && v0 >= dn.x) ? 1u : 0u) |
                               ((v1 >= v0 && v1 >= v2 && v1 >= up.y && v1 >= dn.y) ? 2u : 0u) |
                               ((v2 >= v1 && v2 >= v3 && v2 >= up.z && v2 >= dn.z) ? 4u : 0u) |
                               ((v3 >= v2 && v3 >= right && v3 >= up.w && v3 >= dn.w) ? 8u : 0u);
                        keep &= cand;
                    }
                    else
#pragma unroll
                    for (int e = 0; e < 4; e++)
                        if (cand & (1u << e))
                        {
                            int xe = x + e;
                            if (xe >= a.w)
                                xe -= a.w;
                            const float ve = v[u][e];
                            bool ok = true;
                            if (xe > 0)
                                ok = ve >= (e > 0 ? v[u][e > 0 ? e - 1 : 0] : scan[pix - 1]);
                            if (ok && xe < a.w - 1)
                                ok = ve >= (e < 3 ? v[u][e < 3 ? e + 1 : 3] : scan[pix + 4]);
                            keep |= ok ? (1u << e) : 0u;
                        }
                    cand = keep;
                    while (cand != 0)
                    {
                        const int e = __ffs(cand) - 1;
                        cand &= cand - 1;
                        int xe = x + e, ye = y;
                        if (xe >= a.w) { xe -= a.w; ye++; }
                        if (pix + e < segEnd)
                        {
                            const float ve = e == 0 ? v[u][0] : (e == 1 ? v[u][1] : (e == 2 ? v[u][2] : v[u][3]));
                            const bool pk = MODE == OPB200_SEMANTICS_CPU ? isPeakCpu(scan, a.w, a.h, a.thr, xe, ye, ve)
                                                                         : isPeakCuda(scan, a.w, a.h, xe, ye, ve);
                            mask |= pk ? (1u << e) : 0u;
                        }
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
                                list[slot] = pix + e;
                            slot++;
                        }
                    count += __shfl_sync(0xffffffffu, incl, 31);
                }
            }
        }
        return count;
    }

    // The job is the raster sequence of (channel instance, row tile) pairs, cut into gridDim.x equal contiguous
    // ranges, one persistent block (one per SM) per range -- balanced whatever the batch size. Warp roles:
    //   * producer (1 warp, one thread): streams the tiles through a ring of `stages` shared-memory buffers with bulk
    //     async copies (tile = tileRows rows + a halo row either side, one contiguous 16-byte aligned range of the
    //     plane), so ~200 KB per SM is in flight whatever the other warps are doing;
    //   * scanners (24 warps in `groups` groups of `warpsPerTile`): tile j of the block lives in ring stage j % stages and
    //     is scanned by group j % groups, an equal part per warp, out of shared memory -- the neighbour tests of
    //     above-threshold pixels never wait on global memory. 8 stages and 4 groups of 6: while a group scans a tile
    //     its next one is already loaded or loading. The only per-tile synchronisation is the stage's full/empty
    //     mbarrier pair.
    //     A warp appends its peaks to its own list and records one (start, count) group per tile;
    //   * merger (1 warp): when every scanner has left a band (the part of a channel instance inside this block) it
    //     concatenates the groups in (tile, warp) = raster order, then outputs the band: a channel held by one block
    //     is refined directly; otherwise the band list goes to global memory and the last band of the channel to
    //     finish (atomic ticket) concatenates the band lists and refines. Scanners run up to two bands ahead.
    template <int MODE, bool VEC>
    __global__ void __launch_bounds__((kNmsMaxWarps + 2) * 32) nmsKernel(const NmsArgs a)
    {
        extern __shared__ __align__(128) int smem[];   // warp lists | band list | ring of tiles
        const int S = a.stages, NG = a.groups, WPT = a.warpsPerTile, W = NG * WPT;
        int* warpLists = smem;                                   // [2][W][maxPeaks]
        int* bandLists = smem + 2 * W * a.maxPeaks;              // [kNmsFlushBands][maxPeaks]
        // per (tile of the band, scanning warp of the tile): (first entry in the warp's list) << 16 | entries
        int* groupInfo = bandLists + kNmsFlushBands * a.maxPeaks; // [2][tilesPerPlane * WPT]
        const int maxGroups = a.tilesPerPlane * WPT;
        float* ring = reinterpret_cast<float*>(smem + (((2 * W + kNmsFlushBands) * a.maxPeaks + 2 * maxGroups + 3) & ~3));
        __shared__ int slotCi[kNmsFlushBands], slotCnt[kNmsFlushBands], slotBands[kNmsFlushBands];
        __shared__ int slotOwned[kNmsFlushBands], slotN[kNmsFlushBands];
        __shared__ int slotPrefix[kNmsFlushBands][kNmsMaxBands + 1];
        __shared__ __align__(8) uint64_t fullBar[kNmsMaxStages], emptyBar[kNmsMaxStages], doneBar[2], freeBar[2];

        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        const long long G = gridDim.x;
        const long long g0 = (long long)blockIdx.x * a.tiles / G, g1 = ((long long)blockIdx.x + 1) * a.tiles / G;
        const int tileFloats = (a.tileRows + 2) * a.w;
        // block that owns global tile g (inverse of the range formula above)
        auto blockOf = [&](long long g) { return (int)(((g + 1) * G - 1) / a.tiles); };
        auto planeOf = [&](int ci) {
            const int n = ci / a.parts, c = ci - n * a.parts;
            return a.heat + ((size_t)n * a.heatChannels + c) * a.hw;
        };
        if (threadIdx.x == 0)
        {
            for (int s = 0; s < S; s++)
            {
                sm100::mbarInit(&fullBar[s], 1);
                sm100::mbarInit(&emptyBar[s], WPT);
            }
            for (int p = 0; p < 2; p++)
            {
                sm100::mbarInit(&doneBar[p], W);
                sm100::mbarInit(&freeBar[p], 1);
            }
            sm100::fenceBarrierInit();
        }
        __syncthreads();

        if (warp == W)
        {
            // ---- producer: one thread keeps the ring full ----
            if (VEC && lane == 0)
            {
                int ci = (int)(g0 / a.tilesPerPlane), tp = (int)(g0 - (long long)ci * a.tilesPerPlane);
                const float* plane = planeOf(ci);
                int stage = 0, round = 0;
                for (long long g = g0; g < g1; ++g)
                {
                    if (round > 0)
                        sm100::mbarWait(&emptyBar[stage], (round - 1) & 1);
                    const int r0 = tp * a.tileRows, r1 = min(a.h, r0 + a.tileRows);
                    const int tb = max(0, r0 - 1) * a.w, te = min(a.h, r1 + 1) * a.w;
                    const uint32_t bytes = (uint32_t)(te - tb) * 4u;
                    sm100::mbarExpectTx(&fullBar[stage], bytes);
                    sm100::bulkLoad1d(ring + (size_t)stage * tileFloats, plane + tb, bytes, &fullBar[stage]);
                    if (++stage == S) { stage = 0; round++; }
                    if (++tp == a.tilesPerPlane) { tp = 0; ci++; plane = planeOf(ci); }
                }
            }
            return;
        }

        const int ciFirst = (int)(g0 / a.tilesPerPlane), ciLast = (int)((g1 - 1) / a.tilesPerPlane);

        if (warp == W + 1)
        {
            // ---- merger warp: turns finished bands into output, off the scanning warps' critical path ----
            int nStored = 0;                             // merged bands waiting in shared memory for the flush
            // Flush: output of the stored bands in three phases (instead of one latency chain per band):
            // 1. bands of channels shared with other blocks go to global memory; 2. one lane per band takes the
            // channel's ticket -- the last band to finish owns the channel and prefixes its band counts; 3. all
            // peaks of all owned channels are refined in one flattened parallel pass.
            auto flush = [&]() {
                for (int s = 0; s < nStored; s++)
                {
                    const int ci = slotCi[s];
                    const long long planeFirst = (long long)ci * a.tilesPerPlane;
                    const int firstBlock = blockOf(planeFirst), lastBlock = blockOf(planeFirst + a.tilesPerPlane - 1);
                    const int bands = lastBlock - firstBlock + 1, band = (int)blockIdx.x - firstBlock;
                    if (bands > 1)
                    {
                        int* gList = a.bandLists + ((size_t)ci * kNmsMaxBands + band) * a.maxPeaks;
                        for (int i = lane; i < slotCnt[s]; i += 32)
                            gList[i] = bandLists[s * a.maxPeaks + i];
                        if (lane == 0)
                            a.bandCounts[ci * kNmsMaxBands + band] = slotCnt[s];
                    }
                    if (lane == 0)
                        slotBands[s] = bands;
                }
                __threadfence();
                __syncwarp();
                if (lane < nStored)
                {
                    const int s = lane, ci = slotCi[s], bands = slotBands[s];
                    int owned = 1, n = slotCnt[s];
                    if (bands > 1)
                    {
                        owned = (atomicAdd(&a.tickets[ci], 1) == bands - 1) ? 2 : 0;
                        n = 0;
                        if (owned)
                        {
                            __threadfence();
                            int run = 0;
                            for (int b = 0; b < bands; b++)
                            {
                                slotPrefix[s][b] = run;
                                run += __ldcg(&a.bandCounts[ci * kNmsMaxBands + b]);
                            }
                            slotPrefix[s][bands] = run;
                            n = min(run, a.maxPeaks);
                            a.tickets[ci] = 0;           // self-cleaning for the next launch
                        }
                    }
                    slotOwned[s] = owned;
                    slotN[s] = n;
                    if (owned)
                        a.peaks[(size_t)ci * (a.maxPeaks + 1) * 3] = (float)n;
                }
                __syncwarp();
                int total = 0;
                for (int s = 0; s < nStored; s++)
                    total += slotN[s];
                for (int item = lane; item < total; item += 32)
                {
                    int s = 0, k = item;
                    while (k >= slotN[s])
                        k -= slotN[s++];
                    const int ci = slotCi[s];
                    int pix;
                    if (slotOwned[s] == 1)
                        pix = bandLists[s * a.maxPeaks + k];
                    else
                    {
                        int b = 0;
                        while (slotPrefix[s][b + 1] <= k)
                            b++;
                        pix = __ldcg(&a.bandLists[((size_t)ci * kNmsMaxBands + b) * a.maxPeaks + (k - slotPrefix[s][b])]);
                    }
                    const int y = pix / a.w, x = pix - y * a.w;
                    refinePeak<MODE>(a.peaks + ((size_t)ci * (a.maxPeaks + 1) + k + 1) * 3, planeOf(ci), x, y, a.w, a.h,
                               a.offx, a.offy);
                }
                __syncwarp();                            // slots are reused
                nStored = 0;
            };

            int bandIdx = 0;
            for (int ci = ciFirst; ci <= ciLast; ++ci, ++bandIdx)
            {
                const int par = bandIdx & 1;
                const int* lists = warpLists + par * W * a.maxPeaks;
                const int* groups = groupInfo + par * maxGroups;
                const long long planeFirst = (long long)ci * a.tilesPerPlane;
                const long long ga = max(g0, planeFirst), gb = min(g1, planeFirst + a.tilesPerPlane);
                const int nGroups = (int)(gb - ga) * WPT;
                sm100::mbarWait(&doneBar[par], (bandIdx >> 1) & 1);      // every scanning warp finished this band
                // Concatenate the groups in (tile, warp) = raster order: exclusive prefix of the group sizes, then
                // each non-empty group that starts below maxPeaks is copied to its place (<= maxPeaks copies).
                int bandCount = 0;
                for (int gBase = 0; gBase < nGroups && bandCount < a.maxPeaks; gBase += 32)
                {
                    const int gi = gBase + lane;
                    const int info = gi < nGroups ? groups[gi] : 0;
                    const int cnt = info & 0xffff;
                    int incl = cnt;
#pragma unroll
                    for (int d = 1; d < 32; d <<= 1)
                    {
                        const int t = __shfl_up_sync(0xffffffffu, incl, d);
                        if (lane >= d)
                            incl += t;
                    }
                    const int dst = bandCount + incl - cnt;
                    if (cnt > 0 && dst < a.maxPeaks)
                    {
                        const int tl = gi / WPT, sub = gi - tl * WPT;
                        const int wq = (int)((ga - g0 + tl) % NG) * WPT + sub;
                        const int* src = lists + wq * a.maxPeaks + (info >> 16);
                        const int n = min(cnt, a.maxPeaks - dst);
                        for (int i = 0; i < n; i++)
                            bandLists[nStored * a.maxPeaks + dst + i] = src[i];
                    }
                    bandCount += __shfl_sync(0xffffffffu, incl, 31);
                }
                if (lane == 0)
                {
                    slotCi[nStored] = ci;
                    slotCnt[nStored] = min(bandCount, a.maxPeaks);
                }
                __syncwarp();
                if (lane == 0)
                    sm100::mbarArrive(&freeBar[par]);    // this set of warp lists may be overwritten (band + 2)
                if (++nStored == kNmsFlushBands || ci == ciLast)
                    flush();
            }
            return;
        }

        // ---- scanning warps: nothing but scanning; a warp runs ahead of the others by up to two bands ----
        const int grp = warp / WPT, sub = warp - grp * WPT;
        const float invW = 1.0f / (float)a.w;
        int bandIdx = 0;
        for (int ci = ciFirst; ci <= ciLast; ++ci, ++bandIdx)
        {
            const float* __restrict__ plane = planeOf(ci);
            const long long planeFirst = (long long)ci * a.tilesPerPlane;
            const long long ga = max(g0, planeFirst), gb = min(g1, planeFirst + a.tilesPerPlane);
            const int par = bandIdx & 1;
            int* myList = warpLists + (par * W + warp) * a.maxPeaks;
            int* groups = groupInfo + par * maxGroups;
            if (bandIdx >= 2)
                sm100::mbarWait(&freeBar[par], ((bandIdx >> 1) - 1) & 1);   // band - 2 has been merged
            // this warp's tiles of the band: j = g - g0 with j % NG == grp; tile j sits in ring stage j % S (its (j / S)-th use)
            long long g = ga + ((grp - (int)((ga - g0) % NG)) + NG) % NG;
            int count = 0;
            for (; g < gb; g += NG)
            {
                const int j = (int)(g - g0), stage = j % S;
                const int tp = (int)(g - planeFirst);
                const int r0 = tp * a.tileRows, r1 = min(a.h, r0 + a.tileRows);
                const int segBegin = r0 * a.w + sub * a.segLen, segEnd = min(segBegin + a.segLen, r1 * a.w);
                const int before = min(count, a.maxPeaks);
                if (VEC)
                {
                    sm100::mbarWait(&fullBar[stage], (j / S) & 1);
                    const float* scan = ring + (size_t)stage * tileFloats - max(0, r0 - 1) * a.w;
                    count = scanPixels<MODE, VEC>(a, scan, segBegin, segEnd, myList, count, lane, invW);
                    __syncwarp();
                    if (lane == 0)
                        sm100::mbarArrive(&emptyBar[stage]);
                }
                else
                    count = scanPixels<MODE, VEC>(a, plane, segBegin, segEnd, myList, count, lane, invW);
                if (lane == 0)
                    groups[(int)(g - ga) * WPT + sub] = (before << 16) | (min(count, a.maxPeaks) - before);
            }
            __syncwarp();
            if (lane == 0)
                sm100::mbarArrive(&doneBar[par]);
        }
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

    NmsArgs a;
    a.hw = h * w;
    // staged mode needs 16-byte aligned rows for the bulk copies; otherwise the warps scan global memory directly
    const bool vec = (w % 4 == 0) && ((reinterpret_cast<uintptr_t>(heat) & 15u) == 0);
    const size_t rowBytes = sizeof(float) * (size_t)w;
    const size_t smemMax = 227 * 1024 - 4096;            // static shared memory counts against the 227 KB too
    OPB_REQUIRE(maxPeaks < 32768, "nms: maxPeaks too large");
    // Shared memory: warp lists (two sets), band list, the group table of a whole plane, then the ring.
    // Ring geometry: 8 stages of 8-row tiles when they fit; wider rows get shorter tiles, then fewer stages.
    // The unstaged mode (rows not 16-byte aligned: warps scan global memory) keeps the same tiling without a ring.
    // Scanning warps: 3 per ring stage. With the full ring of 8 stages and tiles large enough the 24 warps form 4 groups of
    // 6 instead of 8 groups of 3: a group then always has its next tile loaded (or loading) while it scans, whereas a
    // group tied to one stage waits out the whole load of its next tile -- on maps where most tiles take the candidate
    // path (crowds, the noise of a random-weight net) the warps spent 37 % of their time in that wait.
    auto warpsPerTileFor = [&](int st, int rows) {
        // (the group count must divide the stage count: a stage is then always consumed by the same group, which waits for
        // tile j + stages only after it has consumed tile j -- with a rotating assignment a group could poll the stage's
        // barrier a whole phase early and the one parity bit would alias; 3 groups of 8 over 8 stages deadlocked)
        constexpr int wide = 2 * kNmsWarpsPerStage;
        static_assert(kNmsMaxStages % (kNmsMaxWarps / wide) == 0, "group count must divide the stage count");
        return (st == kNmsMaxStages && (long long)rows * w >= (long long)wide * 2 * kChunk) ? wide : kNmsWarpsPerStage;
    };
    auto listBytesFor = [&](int st, int rows) {
        const size_t groups = (size_t)((h + rows - 1) / rows) * warpsPerTileFor(st, rows);
        return sizeof(int) * ((((size_t)st * kNmsWarpsPerStage * 2 + kNmsFlushBands) * maxPeaks + 2 * groups + 3) & ~(size_t)3);
    };
    int stages = 0, tileRows = 0;
    for (int st = kNmsMaxStages; st >= 2 && tileRows == 0; st >>= 1)
        for (int rows = std::min(8, h); rows >= 1; rows--)
        {
            if (rows < std::min(4, h) && st > 2)
                break;                                   // prefer fewer stages to very short tiles
            if (listBytesFor(st, rows) + (vec ? (size_t)st * (rows + 2) * rowBytes : 0) <= smemMax)
            {
                stages = st;
                tileRows = rows;
                break;
            }
        }
    OPB_REQUIRE(tileRows > 0, "nms: heat-map rows too wide (or maxPeaks too large) for shared-memory staging");
    const size_t listBytes = listBytesFor(stages, tileRows);
    a.stages = stages;
    a.tileRows = tileRows;
    a.tilesPerPlane = (h + tileRows - 1) / tileRows;
    a.warpsPerTile = warpsPerTileFor(stages, tileRows);
    a.groups = stages * kNmsWarpsPerStage / a.warpsPerTile;
    a.segLen = ((tileRows * w + a.warpsPerTile - 1) / a.warpsPerTile + kChunk - 1) / kChunk * kChunk;
    a.tiles = (long long)ci * a.tilesPerPlane;
    // one block per SM; a channel may be cut into at most kNmsMaxBands bands
    long long gridLL = std::min<long long>(a.tiles, kNumSMsB200);
    gridLL = std::min<long long>(gridLL, (long long)ci * (kNmsMaxBands - 2));
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
    const size_t smem = listBytes + (vec ? (size_t)stages * rowBytes * (size_t)(tileRows + 2) : 0);
    const dim3 grid((unsigned)gridLL);
    const int threads = (stages * kNmsWarpsPerStage + 2) * 32;
    {
        static int configuredDev[64] = {0};   // > 48 KB of dynamic shared memory needs the opt-in, once per device
        int dev = 0;
        OPB_CUDA(cudaGetDevice(&dev));
        if (!configuredDev[dev & 63])
        {
            const int lim = (int)smemMax;
            OPB_CUDA(cudaFuncSetAttribute(nmsKernel<OPB200_SEMANTICS_CPU, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim));
            OPB_CUDA(cudaFuncSetAttribute(nmsKernel<OPB200_SEMANTICS_CPU, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim));
            OPB_CUDA(cudaFuncSetAttribute(nmsKernel<OPB200_SEMANTICS_CUDA, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim));
            OPB_CUDA(cudaFuncSetAttribute(nmsKernel<OPB200_SEMANTICS_CUDA, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim));
            configuredDev[dev & 63] = 1;
        }
    }
    if (mode == OPB200_SEMANTICS_CPU)
    {
        if (vec)
            nmsKernel<OPB200_SEMANTICS_CPU, true><<<grid, threads, smem, st>>>(a);
        else
            nmsKernel<OPB200_SEMANTICS_CPU, false><<<grid, threads, smem, st>>>(a);
    }
    else
    {
        if (vec)
            nmsKernel<OPB200_SEMANTICS_CUDA, true><<<grid, threads, smem, st>>>(a);
        else
            nmsKernel<OPB200_SEMANTICS_CUDA, false><<<grid, threads, smem, st>>>(a);
    }
    OPB_LAUNCH_CHECK();
    return OPB200_OK;
}

// ----------------------------------------------------------------------------------------------------
// opb200_nms_static: the same kernel with its scratch in a static device array, for callers that must not allocate.
// No counterpart in the reference.
// In host code `gNmsStaticScratch` is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
// ----------------------------------------------------------------------------------------------------
namespace opb
{
    static __device__ __align__(256) unsigned char gNmsStaticScratch[OPB200_NMS_STATIC_SCRATCH_BYTES];
}

#include <mutex>

extern "C" opb200_status opb200_nms_static(float* peaks, const float* heat, float threshold, float offset_x,
                                           float offset_y, const int peaks_shape[4], const int heat_shape[4], int mode,
                                           void* stream)
{
    using namespace opb;
    OPB_REQUIRE(peaks && heat && peaks_shape && heat_shape, "nms: null argument");
    OPB_REQUIRE(peaks_shape[0] > 0 && peaks_shape[1] > 0 && peaks_shape[2] > 1, "nms: targetSize inconsistent with sourceSize");
    const size_t need = opb200_nms_scratch_bytes(peaks_shape, heat_shape);
    if (need > (size_t)OPB200_NMS_STATIC_SCRATCH_BYTES)
    {
        setError("nms_static: these shapes need %zu bytes of scratch, the static array holds %u "
                 "(OPB200_NMS_STATIC_SCRATCH_BYTES); use opb200_nms with a caller-owned scratch buffer",
                 need, (unsigned)OPB200_NMS_STATIC_SCRATCH_BYTES);
        return OPB200_ERR_INVALID;
    }
    static std::mutex mtx;
    static void* devicePtr[64] = {nullptr};
    int dev = 0;
    OPB_CUDA(cudaGetDevice(&dev));
    OPB_REQUIRE(dev >= 0 && dev < 64, "nms_static: device ordinal out of range");
    void* ptr = nullptr;
    {
        std::lock_guard<std::mutex> lock(mtx);
        if (devicePtr[dev] == nullptr)
        {
            void* p = nullptr;
            OPB_CUDA(cudaGetSymbolAddress(&p, gNmsStaticScratch));   // address on the CURRENT device
            devicePtr[dev] = p;
        }
        ptr = devicePtr[dev];
    }
    // opb200_nms clears the tickets (the only part that must start at zero) on `stream` and allocates nothing
    return opb200_nms(peaks, ptr, heat, threshold, offset_x, offset_y, peaks_shape, heat_shape, mode, stream);
}
