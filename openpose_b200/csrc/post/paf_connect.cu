// bodyPartConnector on the device: PAF line-integral scores, per-limb sort + greedy one-to-one matching,
// person assembly, threshold filtering and output arrays -- bit-exact in both of the reference's semantics:
// the CPU_ONLY build (default) and the CUDA build (OPB200_SEMANTICS_CUDA: pafPtrIntoVector +
// pafVectorIntoPeopleVector, bodyPartConnectorBase.cpp:475-718, see assembleCudaKernel below).
//
// Replaces op::connectBodyPartsCpu / op::connectBodyPartsGpu (reference:
// src/openpose/net/bodyPartConnectorBase.cpp:12-75 getScoreAB, :157-472 createPeopleVector,
// :721-884 removePeopleBelowThresholdsAndFillFaces, :887-934 peopleVectorToPeopleArray;
// src/openpose/net/bodyPartConnectorBase.cu:15-66,108-145 for the CUDA-semantics score cube).
// The reference GPU path scores on the device, copies a 1.68 MB cube to the host every frame
// (bodyPartConnectorBase.cu:206) and sorts/assembles there. Here two launches keep everything on the device:
//   pafMatchKernel  one block per (limb, frame): scores -> candidates (> 1e-6) -> bitonic sort of 64-bit keys
//                   (score bits | i | j, i.e. std::greater<tuple<double,int,int>>) -> warp-parallel greedy
//                   matching -> <= min(nA,nB) connections;
//   assembleKernel  one block per frame: the 26 limbs in table order; inside a limb every connection is
//                   independent (one-to-one matching => distinct persons), so searches/updates run in
//                   parallel and only the append order is serialised with a prefix sum.
// Float arithmetic mirrors the reference's evaluation order with FMA contraction disabled.
#include "../common.cuh"
#include <algorithm>

namespace opb
{
    constexpr int kMaxParts = 25;
    constexpr int kMaxPairs = 26;
    constexpr int kMatchThreads = 1024;
    constexpr int kAsmThreads = 256;

    struct ModelTables
    {
        int parts, pairs;
        unsigned char pairTable[2 * kMaxPairs];
        unsigned char mapIdx[2 * kMaxPairs];
    };
    // data: reference src/openpose/pose/poseParameters.cpp:253-262 (POSE_MAP_INDEX), :416-424 (POSE_BODY_PART_PAIRS)
    static const ModelTables kBody25 = {
        25, 26,
        {1,8, 1,2, 1,5, 2,3, 3,4, 5,6, 6,7, 8,9, 9,10, 10,11, 8,12, 12,13, 13,14, 1,0, 0,15, 15,17, 0,16, 16,18,
         2,17, 5,18, 14,19, 19,20, 14,21, 11,22, 22,23, 11,24},
        {0,1, 14,15, 22,23, 16,17, 18,19, 24,25, 26,27, 6,7, 2,3, 4,5, 8,9, 10,11, 12,13, 30,31, 32,33, 36,37,
         34,35, 38,39, 20,21, 28,29, 40,41, 42,43, 44,45, 46,47, 48,49, 50,51}};
    static const ModelTables kCoco18 = {
        18, 19,
        {1,2, 1,5, 2,3, 3,4, 5,6, 6,7, 1,8, 8,9, 9,10, 1,11, 11,12, 12,13, 1,0, 0,14, 14,16, 0,15, 15,17, 2,16,
         5,17},
        {12,13, 20,21, 14,15, 16,17, 22,23, 24,25, 0,1, 2,3, 4,5, 6,7, 8,9, 10,11, 28,29, 30,31, 34,35, 32,33,
         36,37, 18,19, 26,27}};

    static const ModelTables* tablesFor(int model)
    {
        if (model == OPB200_POSE_BODY_25)
            return &kBody25;
        if (model == OPB200_POSE_COCO_18)
            return &kCoco18;
        return nullptr;
    }

    struct ConnectArgs
    {
        ModelTables t;
        const float* heat;   // (N, Ch, H, W)
        const float* peaks;  // (N, parts, maxPeaks+1, 3)
        float* cube;         // optional (N, pairs, maxPeaks, maxPeaks)
        int* conn;           // (N, pairs, 1 + 3*maxPeaks): count, i[], j[], scoreBits[]
        unsigned long long* candKeys;   // (N, pairs, maxPeaks^2): candidate keys of pafCandKernel, unordered
        int* candCount;                 // (N, pairs)
        float* outKeypoints; // (N, capacity, parts, 3)
        float* outScores;    // (N, capacity)
        int* outCount;       // (N)
        int capacity;
        int heatChannels, h, w, maxPeaks, sortCap;
        float interThreshold, interMinAbove, nmsThreshold, minSubsetScore, scale;
        double closeThreshold; // sqrt(double(w*h))/150, reference bodyPartConnectorBase.cpp:63
        float closeThresholdF; // sqrtf(w*h)/150, reference bodyPartConnectorBase.cu:59
        float closeScore;      // T(defaultNmsThreshold + 1e-6)
        int minSubsetCnt, maximizePositives, semantics, maxPeople;
    };

    // getScoreAB (CPU semantics) / process() (CUDA semantics). a, b -> (x, y, score).
    // CUDA semantics = process() as the reference's CUDA build evaluates it: nvcc's default -fmad=true fuses
    // norm^2 = fma(vx,vx,vy*vy), the sample position fma(lm,step,s) and the score fma(mapX,ux,mapY*uy) (SASS of
    // pafScoreKernel<float> in oracle/_ref); spelled out with __fmaf_rn, pinned on the reference's kernel by
    // tests/test_refcuda_gpu.py.
    template <int MODE>
    __device__ __forceinline__ float pafScore(const float* __restrict__ a, const float* __restrict__ b,
                                              const float* __restrict__ mapX, const float* __restrict__ mapY,
                                              const ConnectArgs& c)
    {
        const float ax = a[0], ay = a[1];
        const float vx = sub(b[0], ax), vy = sub(b[1], ay);
        const float vmax = fmaxf(fabsf(vx), fabsf(vy));
        const int n = max(5, min(25, posIntRound(sqrtf(mul(5.f, vmax)))));
        const float norm = MODE == OPB200_SEMANTICS_CUDA ? sqrtf(__fmaf_rn(vx, vx, mul(vy, vy)))
                                                         : sqrtf(add(mul(vx, vx), mul(vy, vy)));
        if ((double)norm > 1e-6)
        {
            const float ux = divf(vx, norm), uy = divf(vy, norm);
            const float fn = (float)n;
            const float stepx = divf(vx, fn), stepy = divf(vy, fn);
            // Smallest count that passes `count / n > interMinAbove` in the reference's own float arithmetic (the
            // quotient is monotonic in count). Once more than n - minCount samples have failed the pair cannot pass
            // any more and its result no longer depends on the remaining samples: stop reading the maps. Exact.
            int minCount = (int)mul(c.interMinAbove, fn);
            minCount = max(0, min(n + 1, minCount));
            while (minCount <= n && !(divf((float)minCount, fn) > c.interMinAbove))
                ++minCount;
            while (minCount > 0 && divf((float)(minCount - 1), fn) > c.interMinAbove)
                --minCount;
            const int failsAllowed = n - minCount;
            float sum = 0.f;
            unsigned count = 0;
            int fails = 0;
            for (int lm = 0; lm < n; lm++)
            {
                int mx, my;
                if (MODE == OPB200_SEMANTICS_CPU)
                {
                    mx = max(0, min(c.w - 1, posIntRound(add(ax, mul((float)lm, stepx)))));
                    my = max(0, min(c.h - 1, posIntRound(add(ay, mul((float)lm, stepy)))));
                }
                else
                {
                    mx = min(c.w - 1, posIntRound(__fmaf_rn((float)lm, stepx, ax)));
                    my = min(c.h - 1, posIntRound(__fmaf_rn((float)lm, stepy, ay)));
                }
                const int idx = my * c.w + mx;
                const float sc = MODE == OPB200_SEMANTICS_CPU
                                     ? add(mul(ux, __ldg(mapX + idx)), mul(uy, __ldg(mapY + idx)))
                                     : __fmaf_rn(__ldg(mapX + idx), ux, mul(__ldg(mapY + idx), uy));
                if (sc > c.interThreshold)
                {
                    sum = add(sum, sc);
                    count++;
                }
                else if (++fails > failsAllowed)
                    break;
            }
            if (divf((float)count, fn) > c.interMinAbove)
                return divf(sum, (float)count);
            if (MODE == OPB200_SEMANTICS_CPU)
            {
                if ((double)norm < c.closeThreshold)
                    return c.closeScore;
            }
            else if (norm < c.closeThresholdF)
                return c.closeScore;
        }
        return MODE == OPB200_SEMANTICS_CPU ? 0.f : -1.f;
    }

    // ------------------------------------------------------------------------------------------------
    // Score cube only (pafScoreKernel layout, -1 outside nA x nB).
    // ------------------------------------------------------------------------------------------------
    template <int MODE>
    __global__ void __launch_bounds__(256) pafCubeKernel(const ConnectArgs c)
    {
        const int pair = blockIdx.y, n = blockIdx.z;
        const int mp = c.maxPeaks, po = 3 * (mp + 1);
        const float* pk = c.peaks + (size_t)n * c.t.parts * po;
        const float* A = pk + c.t.pairTable[2 * pair] * po;
        const float* B = pk + c.t.pairTable[2 * pair + 1] * po;
        const size_t hw = (size_t)c.h * c.w;
        const float* mapX = c.heat + ((size_t)n * c.heatChannels + c.t.parts + 1 + c.t.mapIdx[2 * pair]) * hw;
        const float* mapY = c.heat + ((size_t)n * c.heatChannels + c.t.parts + 1 + c.t.mapIdx[2 * pair + 1]) * hw;
        const float nA = A[0], nB = B[0];
        float* out = c.cube + ((size_t)n * c.t.pairs + pair) * mp * mp;
        for (int t = blockIdx.x * 256 + threadIdx.x; t < mp * mp; t += gridDim.x * 256)
        {
            const int i = t / mp, j = t - i * mp;
            float v = -1.f;
            if ((float)i < nA && (float)j < nB)
                v = pafScore<MODE>(A + 3 * (i + 1), B + 3 * (j + 1), mapX, mapY, c);
            out[t] = v;
        }
    }

    // ------------------------------------------------------------------------------------------------
    // Kernel A0: candidate scores of every (limb, frame) at full occupancy. The line integral is a chain of dependent
    // gathers, so it wants as many resident warps as the SM holds; the sort of kernel A needs 128 KB of shared memory
    // per CTA and would cap that at one CTA per SM. Candidates (score > 1e-6) go to a per-(limb, frame) list in
    // arbitrary order: their keys are distinct, so the sort of kernel A makes the result deterministic.
    // ------------------------------------------------------------------------------------------------
    constexpr int kCandThreads = 256;
    __global__ void __launch_bounds__(kCandThreads) pafCandKernel(const ConnectArgs c)
    {
        const int pair = blockIdx.y, n = blockIdx.z;
        const int mp = c.maxPeaks, po = 3 * (mp + 1);
        const float* pk = c.peaks + (size_t)n * c.t.parts * po;
        const float* A = pk + c.t.pairTable[2 * pair] * po;
        const float* B = pk + c.t.pairTable[2 * pair + 1] * po;
        const int nA = posIntRound(A[0]), nB = posIntRound(B[0]);
        const size_t hw = (size_t)c.h * c.w;
        const float* mapX = c.heat + ((size_t)n * c.heatChannels + c.t.parts + 1 + c.t.mapIdx[2 * pair]) * hw;
        const float* mapY = c.heat + ((size_t)n * c.heatChannels + c.t.parts + 1 + c.t.mapIdx[2 * pair + 1]) * hw;
        // a few blocks per (limb, frame) stride over its nA * nB pairs: a grid sized for maxPeaks^2 pairs is 64 blocks per
        // (limb, frame), of which an 8-person frame uses one (73 216 blocks, 68 us, for 44 frames)
        for (int t = blockIdx.x * kCandThreads + threadIdx.x; t < nA * nB; t += gridDim.x * kCandThreads)
        {
            const int i = t / nB + 1, j = t - (i - 1) * nB + 1;
            const float s = pafScore<OPB200_SEMANTICS_CPU>(A + 3 * i, B + 3 * j, mapX, mapY, c);
            if ((double)s > 1e-6) // reference: bodyPartConnectorBase.cpp:317
            {
                const size_t slot = (size_t)n * c.t.pairs + pair;
                const int pos = atomicAdd(c.candCount + slot, 1);
                c.candKeys[slot * mp * mp + pos] =
                    ((unsigned long long)__float_as_uint(s) << 16) | ((unsigned)i << 8) | (unsigned)j;
            }
        }
    }

    // ------------------------------------------------------------------------------------------------
    // Kernel A: (scores, unless kernel A0 ran) + sort + greedy matching for one (limb, frame).
    // ------------------------------------------------------------------------------------------------
    __global__ void __launch_bounds__(kMatchThreads) pafMatchKernel(const ConnectArgs c)
    {
        extern __shared__ unsigned long long keys[]; // [sortCap]
        __shared__ int nCand;
        __shared__ unsigned usedA[8], usedB[8];      // bit k = peak index k used (k <= 255)
        __shared__ int nConShared;

        const int pair = blockIdx.x, n = blockIdx.y;
        const int mp = c.maxPeaks, po = 3 * (mp + 1);
        const float* pk = c.peaks + (size_t)n * c.t.parts * po;
        const float* A = pk + c.t.pairTable[2 * pair] * po;
        const float* B = pk + c.t.pairTable[2 * pair + 1] * po;
        int* conn = c.conn + ((size_t)n * c.t.pairs + pair) * (1 + 3 * mp);
        const int nA = posIntRound(A[0]), nB = posIntRound(B[0]);
        if (nA == 0 || nB == 0)
        {
            if (threadIdx.x == 0)
                conn[0] = 0;
            return;
        }
        if (threadIdx.x == 0)
        {
            nCand = 0;
            nConShared = 0;
        }
        if (threadIdx.x < 8)
        {
            usedA[threadIdx.x] = 0;
            usedB[threadIdx.x] = 0;
        }
        __syncthreads();
        const size_t hw = (size_t)c.h * c.w;
        const float* mapX = c.heat + ((size_t)n * c.heatChannels + c.t.parts + 1 + c.t.mapIdx[2 * pair]) * hw;
        const float* mapY = c.heat + ((size_t)n * c.heatChannels + c.t.parts + 1 + c.t.mapIdx[2 * pair + 1]) * hw;
        if (c.candKeys != nullptr)
        {
            const size_t slot = (size_t)n * c.t.pairs + pair;
            const int have = c.candCount[slot];
            const unsigned long long* src = c.candKeys + slot * mp * mp;
            for (int t = threadIdx.x; t < have; t += kMatchThreads)
                keys[t] = src[t];
            if (threadIdx.x == 0)
                nCand = have;
        }
        else
            for (int t = threadIdx.x; t < nA * nB; t += kMatchThreads)
            {
                const int i = t / nB + 1, j = t - (i - 1) * nB + 1;
                const float s = pafScore<OPB200_SEMANTICS_CPU>(A + 3 * i, B + 3 * j, mapX, mapY, c);
                if ((double)s > 1e-6) // reference: bodyPartConnectorBase.cpp:317
                {
                    const int pos = atomicAdd(&nCand, 1);
                    keys[pos] = ((unsigned long long)__float_as_uint(s) << 16) | ((unsigned)i << 8) | (unsigned)j;
                }
            }
        __syncthreads();
        const int cand = nCand;
        int sortN = 1;
        while (sortN < cand)
            sortN <<= 1;
        for (int t = cand + threadIdx.x; t < sortN; t += kMatchThreads)
            keys[t] = 0ull;
        __syncthreads();
        // bitonic sort, descending. score > 0 => its bit pattern orders like the value; ties fall to larger i, j.
        for (int k = 2; k <= sortN; k <<= 1)
            for (int j = k >> 1; j > 0; j >>= 1)
            {
                for (int t = threadIdx.x; t < sortN; t += kMatchThreads)
                {
                    const int p = t ^ j;
                    if (p > t)
                    {
                        const unsigned long long x = keys[t], y = keys[p];
                        const bool desc = (t & k) == 0;
                        if (desc ? (x < y) : (x > y))
                        {
                            keys[t] = y;
                            keys[p] = x;
                        }
                    }
                }
                __syncthreads();
            }
        // greedy one-to-one matching (reference :350-372), warp 0, 32 candidates per step
        if (threadIdx.x < 32)
        {
            const int lane = threadIdx.x;
            const int minAB = min(nA, nB);
            int nCon = 0;
            for (int base = 0; base < cand && nCon < minAB; base += 32)
            {
                const bool have = base + lane < cand;
                const unsigned long long key = have ? keys[base + lane] : 0ull;
                const int i = (int)((key >> 8) & 0xffu), j = (int)(key & 0xffu);
                unsigned alive = __ballot_sync(0xffffffffu, have);
                while (alive && nCon < minAB)
                {
                    const bool freePair = ((alive >> lane) & 1u) && !((usedA[i >> 5] >> (i & 31)) & 1u)
                                          && !((usedB[j >> 5] >> (j & 31)) & 1u);
                    const unsigned freeMask = __ballot_sync(0xffffffffu, freePair);
                    if (!freeMask)
                        break;
                    const int first = __ffs(freeMask) - 1;
                    if (lane == first)
                    {
                        conn[1 + nCon] = i;
                        conn[1 + mp + nCon] = j;
                        conn[1 + 2 * mp + nCon] = (int)(key >> 16);
                        usedA[i >> 5] |= 1u << (i & 31);
                        usedB[j >> 5] |= 1u << (j & 31);
                    }
                    nCon++;
                    alive = freeMask & ~((2u << first) - 1u); // only later lanes that were free can still be free
                    __syncwarp();
                }
            }
            if (lane == 0)
                conn[0] = nCon;
        }
    }
}

namespace opb
{
    // ------------------------------------------------------------------------------------------------
    // Kernel B: person assembly for one frame (createPeopleVector :157-472, filter :721-884, output :887-934).
    // People live in shared memory as structure-of-arrays: rows[part][person] = peak index k (1..maxPeaks,
    // 0 = none; the reference's flat index is part*3*(maxPeaks+1) + 3k + 2), cnt[person], score[person].
    // Slots beyond nPeople are always all-zero.
    // ------------------------------------------------------------------------------------------------
    __device__ __forceinline__ int blockExclusiveScan(int value, int* warpSums, int& total)
    {
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        int incl = value;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1)
        {
            const int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d)
                incl += t;
        }
        __syncthreads(); // warpSums may still be read from a previous call
        if (lane == 31)
            warpSums[warp] = incl;
        __syncthreads();
        int offset = 0, sum = 0;
#pragma unroll
        for (int wq = 0; wq < kAsmThreads / 32; wq++)
        {
            const int s = warpSums[wq];
            if (wq < warp)
                offset += s;
            sum += s;
        }
        total = sum;
        return offset + incl - value;
    }

    __global__ void __launch_bounds__(kAsmThreads) assembleKernel(const ConnectArgs c)
    {
        extern __shared__ unsigned char smraw[];
        const int cap = c.maxPeople;                       // multiple of 4
        float* score = reinterpret_cast<float*>(smraw);    // [cap]
        unsigned char* cnt = smraw + sizeof(float) * cap;  // [cap]
        unsigned char* rows = cnt + cap;                   // [parts][cap]
        __shared__ int warpSums[kAsmThreads / 32];
        __shared__ int found[256];                         // per connection / per peak: person index or -1
        __shared__ unsigned char byA[256], byB[256];       // ear limbs: partner lookup
        __shared__ int nPeopleS;

        const int n = blockIdx.x;
        const int parts = c.t.parts, pairs = c.t.pairs, mp = c.maxPeaks, po = 3 * (mp + 1);
        const float* pk = c.peaks + (size_t)n * parts * po;
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        constexpr int kWarps = kAsmThreads / 32;

        for (int i = threadIdx.x; i < (int)((sizeof(float) + 1 + parts) * cap / 4); i += kAsmThreads)
            reinterpret_cast<unsigned*>(smraw)[i] = 0u;
        if (threadIdx.x == 0)
            nPeopleS = 0;
        __syncthreads();

        const int earFirst = parts == 18 ? 17 : 18; // reference :400-405
        for (int pair = 0; pair < pairs; pair++)
        {
            const int partA = c.t.pairTable[2 * pair], partB = c.t.pairTable[2 * pair + 1];
            const float* A = pk + partA * po;
            const float* B = pk + partB * po;
            const int nA = posIntRound(A[0]), nB = posIntRound(B[0]);
            const int nPeople = nPeopleS;
            unsigned char* rowA = rows + partA * cap;
            unsigned char* rowB = rows + partB * cap;
            if (nA == 0 || nB == 0)
            {
                // reference :190-289 (non-MPI): singletons of the non-empty side not yet referenced
                const int part = nA == 0 ? partB : partA;
                const float* C = nA == 0 ? B : A;
                const int cntSide = nA == 0 ? nB : nA;
                unsigned char* row = rows + part * cap;
                for (int i = 1 + warp; i <= cntSide; i += kWarps)
                {
                    bool hit = false;
                    for (int q = lane; q < nPeople && !hit; q += 32)
                        hit = row[q] == i;
                    hit = __any_sync(0xffffffffu, hit);
                    if (lane == 0)
                        found[i] = hit ? 1 : 0;
                }
                __syncthreads();
                int total = 0;
                const int idx = threadIdx.x + 1; // peak index handled by this thread in the append
                const int isNew = (idx <= cntSide && !found[idx]) ? 1 : 0;
                const int slot = blockExclusiveScan(isNew, warpSums, total);
                if (isNew)
                {
                    const int q = nPeople + slot;
                    row[q] = (unsigned char)idx;
                    cnt[q] = 1;
                    score[q] = C[idx * 3 + 2];
                }
                __syncthreads();
                if (threadIdx.x == 0)
                    nPeopleS = nPeople + total;
                __syncthreads();
                continue;
            }
            const int* conn = c.conn + ((size_t)n * pairs + pair) * (1 + 3 * mp);
            const int nCon = conn[0];
            if (nCon == 0)
                continue; // block-uniform
            if (pair == 0)
            {
                // reference :378-393
                for (int k = threadIdx.x; k < nCon; k += kAsmThreads)
                {
                    const int i = conn[1 + k], j = conn[1 + mp + k];
                    const float s = __int_as_float(conn[1 + 2 * mp + k]);
                    const int q = nPeople + k;
                    rowA[q] = (unsigned char)i;
                    rowB[q] = (unsigned char)j;
                    cnt[q] = 2;
                    score[q] = add(add(A[3 * i + 2], B[3 * j + 2]), s);
                }
                __syncthreads();
                if (threadIdx.x == 0)
                    nPeopleS = nPeople + nCon;
                __syncthreads();
            }
            else if (pair == earFirst || pair == earFirst + 1)
            {
                // reference :399-428: fill the missing side in EVERY matching person; no count/score change
                for (int k = threadIdx.x; k < 256; k += kAsmThreads)
                {
                    byA[k] = 0;
                    byB[k] = 0;
                }
                __syncthreads();
                for (int k = threadIdx.x; k < nCon; k += kAsmThreads)
                {
                    const int i = conn[1 + k], j = conn[1 + mp + k];
                    byA[i] = (unsigned char)j;
                    byB[j] = (unsigned char)i;
                }
                __syncthreads();
                for (int q = threadIdx.x; q < nPeople; q += kAsmThreads)
                {
                    const int a = rowA[q], b = rowB[q];
                    if (a != 0 && b == 0)
                    {
                        if (byA[a])
                            rowB[q] = byA[a];
                    }
                    else if (b != 0 && a == 0)
                    {
                        if (byB[b])
                            rowA[q] = byB[b];
                    }
                }
                __syncthreads();
            }
            else
            {
                // reference :430-461: first person (list order) holding A gets B; otherwise a new person
                for (int k = warp; k < nCon; k += kWarps)
                {
                    const int i = conn[1 + k];
                    int hitAt = -1;
                    for (int base = 0; base < nPeople && hitAt < 0; base += 32)
                    {
                        const int q = base + lane;
                        const unsigned m = __ballot_sync(0xffffffffu, q < nPeople && rowA[q] == i);
                        if (m)
                            hitAt = base + __ffs(m) - 1;
                    }
                    if (lane == 0)
                    {
                        found[k] = hitAt;
                        if (hitAt >= 0)
                        {
                            const int j = conn[1 + mp + k];
                            const float s = __int_as_float(conn[1 + 2 * mp + k]);
                            rowB[hitAt] = (unsigned char)j;
                            cnt[hitAt] = (unsigned char)(cnt[hitAt] + 1);
                            score[hitAt] = add(score[hitAt], add(B[3 * j + 2], s));
                        }
                    }
                }
                __syncthreads();
                int total = 0;
                const int k = threadIdx.x;
                const int isNew = (k < nCon && found[k] < 0) ? 1 : 0;
                const int slot = blockExclusiveScan(isNew, warpSums, total);
                if (isNew)
                {
                    const int i = conn[1 + k], j = conn[1 + mp + k];
                    const float s = __int_as_float(conn[1 + 2 * mp + k]);
                    const int q = nPeople + slot;
                    rowA[q] = (unsigned char)i;
                    rowB[q] = (unsigned char)j;
                    cnt[q] = 2;
                    score[q] = add(add(A[3 * i + 2], B[3 * j + 2]), s);
                }
                __syncthreads();
                if (threadIdx.x == 0)
                    nPeopleS = nPeople + total;
                __syncthreads();
            }
        }

        // filter (reference :721-884) + ordered output (reference :887-934)
        const int nPeople = nPeopleS;
        const float inv = divf(1.f, (float)(parts + pairs));
        float* outKp = c.outKeypoints + (size_t)n * c.capacity * parts * 3;
        float* outSc = c.outScores + (size_t)n * c.capacity;
        int written = 0;
        for (int attempt = c.maximizePositives ? 1 : 0; attempt < 2; attempt++)
        {
            const bool maxPos = attempt == 1;
            written = 0;
            for (int base = 0; base < nPeople; base += kAsmThreads) // block-uniform trip count
            {
                const int q = base + threadIdx.x;
                int valid = 0;
                if (q < nPeople)
                {
                    int counter = cnt[q];
                    bool skip = false;
                    if (!maxPos && parts == 25)
                    {
                        int feet = 0;
                        for (int p = 19; p < 25; p++)
                            feet += rows[p * cap + q] > 0;
                        const int before = counter;
                        counter -= feet;
                        skip = counter != before && counter <= 4;
                    }
                    valid = !skip && counter >= c.minSubsetCnt
                            && divf(score[q], (float)counter) >= c.minSubsetScore;
                }
                int total = 0;
                const int slot = written + blockExclusiveScan(valid, warpSums, total);
                if (valid && slot < c.capacity)
                {
                    for (int p = 0; p < parts; p++)
                    {
                        const int k = rows[p * cap + q];
                        float* o = outKp + ((size_t)slot * parts + p) * 3;
                        if (k > 0)
                        {
                            const float* src = pk + p * po + 3 * k;
                            o[0] = mul(src[0], c.scale);
                            o[1] = mul(src[1], c.scale);
                            o[2] = src[2];
                        }
                        else
                        {
                            o[0] = 0.f;
                            o[1] = 0.f;
                            o[2] = 0.f;
                        }
                    }
                    outSc[slot] = mul(score[q], inv);
                }
                written += total;
            }
            if (written > 0)
                break; // reference :870-874: retry with maximizePositives only when nobody survived
        }
        if (threadIdx.x == 0)
            c.outCount[n] = written;
    }

    // ------------------------------------------------------------------------------------------------
    // Kernel C: CUDA-build semantics of the assembly (pafPtrIntoVector :475-542 + pafVectorIntoPeopleVector
    // :545-718), fed by the score cube of pafCubeKernel<CUDA>. One block per frame:
    //   1. every cube cell with score > 1e-6 becomes a connection keyed by the reference's tuple
    //      (total = paf + 0.1*scoreA + 0.1*scoreB, paf, limb, indexA, indexB) packed into 128 bits;
    //   2. ONE descending sort over all limbs (bitonic; in shared memory up to kCudaSortSmem keys, else in the
    //      global scratch -- slow but exact for pathological maps);
    //   3. the five-case assignment walked in order by the whole block in lock-step (state in shared memory,
    //      people rows hold the reference's flat peak index, merging two people re-labels `assigned` in parallel);
    //   4. the same filter and output as the CPU-semantics kernel, skipping merged-away people.
    // ------------------------------------------------------------------------------------------------
    constexpr int kCudaSortSmem = 4096;
    struct ConnKey
    {
        unsigned long long hi, lo;   // hi = ord(total) << 32 | ord(paf); lo = limb << 16 | indexA << 8 | indexB
    };
    __device__ __forceinline__ bool keyLess(const ConnKey& a, const ConnKey& b)
    {
        return a.hi < b.hi || (a.hi == b.hi && a.lo < b.lo);
    }
    __device__ __forceinline__ unsigned ordFloat(float f)   // monotone float -> unsigned
    {
        const unsigned u = __float_as_uint(f);
        return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    }
    __device__ __forceinline__ float unordFloat(unsigned u)
    {
        return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
    }
    // descending bitonic sort of n = 2^k keys by the whole block
    __device__ void bitonicSortDesc(ConnKey* keys, int n)
    {
        for (int size = 2; size <= n; size <<= 1)
            for (int stride = size >> 1; stride > 0; stride >>= 1)
            {
                for (int t = threadIdx.x; t < (n >> 1); t += kAsmThreads)
                {
                    const int lo = 2 * t - (t & (stride - 1));
                    const int hi = lo + stride;
                    const bool descending = (lo & size) == 0;
                    const ConnKey a = keys[lo], b = keys[hi];
                    if (keyLess(a, b) == descending)
                    {
                        keys[lo] = b;
                        keys[hi] = a;
                    }
                }
                __syncthreads();
            }
    }

    __global__ void __launch_bounds__(kAsmThreads) assembleCudaKernel(const ConnectArgs c, ConnKey* keysAll,
                                                                      size_t keysPerFrame)
    {
        extern __shared__ __align__(16) unsigned char smraw[];
        __shared__ int warpSums[kAsmThreads / 32];
        __shared__ int nKeysS, nPeopleS;
        const int n = blockIdx.x;
        const int parts = c.t.parts, pairs = c.t.pairs, mp = c.maxPeaks, po = 3 * (mp + 1);
        const float* pk = c.peaks + (size_t)n * parts * po;
        const float* cube = c.cube + (size_t)n * pairs * mp * mp;
        ConnKey* keys = keysAll + (size_t)n * keysPerFrame;

        // ---- 1. gather ----
        if (threadIdx.x == 0)
            nKeysS = 0;
        __syncthreads();
        for (int pair = 0; pair < pairs; pair++)
        {
            const int partA = c.t.pairTable[2 * pair], partB = c.t.pairTable[2 * pair + 1];
            const int nA = posIntRound(pk[partA * po]), nB = posIntRound(pk[partB * po]);
            for (int t = threadIdx.x; t < nA * nB; t += kAsmThreads)
            {
                const int a = t / nB, b = t - a * nB;
                const float s = cube[((size_t)pair * mp + a) * mp + b];
                if ((double)s > 1e-6)
                {
                    const float sA = pk[partA * po + (a + 1) * 3 + 2], sB = pk[partB * po + (b + 1) * 3 + 2];
                    const float total = add(add(s, mul(0.1f, sA)), mul(0.1f, sB));
                    ConnKey k;
                    k.hi = ((unsigned long long)ordFloat(total) << 32) | ordFloat(s);
                    k.lo = ((unsigned long long)pair << 16) | ((unsigned long long)(a + 1) << 8) | (unsigned)(b + 1);
                    keys[atomicAdd(&nKeysS, 1)] = k;
                }
            }
        }
        __syncthreads();
        const int nKeys = nKeysS;

        // ---- 2. one descending sort over all limbs (keys are unique: ties cannot reorder) ----
        int sortN = 1;
        while (sortN < nKeys)
            sortN <<= 1;
        if (nKeys > 1)
        {
            const ConnKey zero = {0ull, 0ull};            // sorts last
            if (sortN <= kCudaSortSmem)
            {
                ConnKey* sk = reinterpret_cast<ConnKey*>(smraw);
                for (int t = threadIdx.x; t < sortN; t += kAsmThreads)
                    sk[t] = t < nKeys ? keys[t] : zero;
                __syncthreads();
                bitonicSortDesc(sk, sortN);
                for (int t = threadIdx.x; t < nKeys; t += kAsmThreads)
                    keys[t] = sk[t];
            }
            else
            {
                for (int t = nKeys + threadIdx.x; t < sortN; t += kAsmThreads)
                    keys[t] = zero;
                __syncthreads();
                bitonicSortDesc(keys, sortN);
            }
        }
        __syncthreads();

        // ---- 3. assignment ----
        const int cap = c.maxPeople;                       // people slots (each new person consumes two free peaks)
        float* score = reinterpret_cast<float*>(smraw);                           // [cap]
        unsigned short* rows = reinterpret_cast<unsigned short*>(score + cap);    // [cap][parts] flat peak index
        short* assigned = reinterpret_cast<short*>(rows + (size_t)cap * parts);   // [parts][mp] person or -1
        unsigned char* cnt = reinterpret_cast<unsigned char*>(assigned + parts * mp);   // [cap]
        unsigned char* dead = cnt + cap;                                          // [cap]
        for (int i = threadIdx.x; i < cap * parts; i += kAsmThreads)
            rows[i] = 0;
        for (int i = threadIdx.x; i < parts * mp; i += kAsmThreads)
            assigned[i] = -1;
        for (int i = threadIdx.x; i < cap; i += kAsmThreads)
        {
            score[i] = 0.f;
            cnt[i] = 0;
            dead[i] = 0;
        }
        if (threadIdx.x == 0)
            nPeopleS = 0;
        __syncthreads();
        for (int k = 0; k < nKeys; k++)
        {
            // every thread reads the same state, takes the same branch; designated threads write; one barrier
            const ConnKey key = keys[k];
            const int pair = (int)(key.lo >> 16), ia = (int)((key.lo >> 8) & 0xff), ib = (int)(key.lo & 0xff);
            const float paf = unordFloat((unsigned)(key.hi & 0xffffffffu));
            const int partA = c.t.pairTable[2 * pair], partB = c.t.pairTable[2 * pair + 1];
            const int idxA = (partA * (mp + 1) + ia) * 3 + 2, idxB = (partB * (mp + 1) + ib) * 3 + 2;
            const int sa = partA * mp + ia - 1, sb = partB * mp + ib - 1;
            const int aAs = assigned[sa], bAs = assigned[sb];
            const int nPeople = nPeopleS;
            if (aAs < 0 && bAs < 0)
            {
                __syncthreads();                           // everybody has read nPeopleS / assigned
                if (nPeople < cap)
                {
                    if ((int)threadIdx.x < parts)
                        rows[nPeople * parts + threadIdx.x] =
                            (unsigned short)((int)threadIdx.x == partA ? idxA : ((int)threadIdx.x == partB ? idxB : 0));
                    if (threadIdx.x == 0)
                    {
                        cnt[nPeople] = 2;
                        score[nPeople] = add(add(pk[idxA], pk[idxB]), paf);
                        assigned[sa] = (short)nPeople;
                        assigned[sb] = (short)nPeople;
                        nPeopleS = nPeople + 1;
                    }
                }
            }
            else if ((aAs >= 0) != (bAs >= 0))
            {
                const int have = aAs >= 0 ? aAs : bAs;
                const int part2 = aAs >= 0 ? partB : partA, idx2 = aAs >= 0 ? idxB : idxA;
                const bool freeSlot = rows[have * parts + part2] == 0;
                __syncthreads();
                if (freeSlot && threadIdx.x == 0)
                {
                    rows[have * parts + part2] = (unsigned short)idx2;
                    cnt[have] = (unsigned char)(cnt[have] + 1);
                    score[have] = add(score[have], add(pk[idx2], paf));
                    assigned[aAs >= 0 ? sb : sa] = (short)have;
                }
            }
            else if (aAs == bAs)
            {
                __syncthreads();
                if (threadIdx.x == 0)
                    score[aAs] = add(score[aAs], paf);
            }
            else
            {
                const int a1 = min(aAs, bAs), a2 = max(aAs, bAs);
                const bool clash = (int)threadIdx.x < parts && rows[a1 * parts + threadIdx.x] > 0 &&
                                   rows[a2 * parts + threadIdx.x] > 0;
                if (!__syncthreads_or(clash ? 1 : 0))      // complementary: the two people share no body part
                {
                    if ((int)threadIdx.x < parts && rows[a1 * parts + threadIdx.x] == 0)
                        rows[a1 * parts + threadIdx.x] = rows[a2 * parts + threadIdx.x];
                    if (threadIdx.x == 0)
                    {
                        cnt[a1] = (unsigned char)(cnt[a1] + cnt[a2]);
                        score[a1] = add(score[a1], add(score[a2], paf));
                        dead[a2] = 1;                      // erased after the loop in the reference (:707-709)
                    }
                    for (int i = threadIdx.x; i < parts * mp; i += kAsmThreads)
                        if (assigned[i] == a2)
                            assigned[i] = (short)a1;
                }
            }
            __syncthreads();
        }

        // ---- 4. filter (reference :721-884) + ordered output (reference :887-934) ----
        const int nPeople = nPeopleS;
        const float inv = divf(1.f, (float)(parts + pairs));
        float* outKp = c.outKeypoints + (size_t)n * c.capacity * parts * 3;
        float* outSc = c.outScores + (size_t)n * c.capacity;
        int written = 0;
        for (int attempt = c.maximizePositives ? 1 : 0; attempt < 2; attempt++)
        {
            const bool maxPos = attempt == 1;
            written = 0;
            for (int base = 0; base < nPeople; base += kAsmThreads) // block-uniform trip count
            {
                const int q = base + threadIdx.x;
                int valid = 0;
                if (q < nPeople && !dead[q])
                {
                    int counter = cnt[q];
                    bool skip = false;
                    if (!maxPos && parts == 25)
                    {
                        int feet = 0;
                        for (int p = 19; p < 25; p++)
                            feet += rows[q * parts + p] > 0;
                        const int before = counter;
                        counter -= feet;
                        skip = counter != before && counter <= 4;
                    }
                    valid = !skip && counter >= c.minSubsetCnt
                            && divf(score[q], (float)counter) >= c.minSubsetScore;
                }
                int total = 0;
                const int slot = written + blockExclusiveScan(valid, warpSums, total);
                if (valid && slot < c.capacity)
                {
                    for (int p = 0; p < parts; p++)
                    {
                        const int idx = rows[q * parts + p];
                        float* o = outKp + ((size_t)slot * parts + p) * 3;
                        if (idx > 0)
                        {
                            o[0] = mul(pk[idx - 2], c.scale);
                            o[1] = mul(pk[idx - 1], c.scale);
                            o[2] = pk[idx];
                        }
                        else
                        {
                            o[0] = 0.f;
                            o[1] = 0.f;
                            o[2] = 0.f;
                        }
                    }
                    outSc[slot] = mul(score[q], inv);
                }
                written += total;
            }
            if (written > 0)
                break;
        }
        if (threadIdx.x == 0)
            c.outCount[n] = written;
    }

    static opb200_status fillArgs(ConnectArgs& c, const opb200_connect_params* p, const int heat_shape[4])
    {
        OPB_REQUIRE(p != nullptr && heat_shape != nullptr, "connector: null argument");
        const ModelTables* t = tablesFor(p->pose_model);
        // reference :165-167
        OPB_REQUIRE(t != nullptr, "Model not implemented for CPU body connector.");
        OPB_REQUIRE(p->max_peaks >= 1 && p->max_peaks <= 255, "connector: max_peaks must be in [1, 255]");
        c.t = *t;
        c.heatChannels = heat_shape[1];
        c.h = heat_shape[2];
        c.w = heat_shape[3];
        OPB_REQUIRE(heat_shape[0] > 0 && c.h > 0 && c.w > 0, "connector: empty heat-map blob");
        OPB_REQUIRE(c.heatChannels >= t->parts + 1 + 2 * t->pairs, "connector: heat-map blob has too few channels");
        c.maxPeaks = p->max_peaks;
        c.interThreshold = p->inter_threshold;
        c.interMinAbove = p->inter_min_above_threshold;
        c.nmsThreshold = p->default_nms_threshold;
        c.minSubsetScore = p->min_subset_score;
        c.minSubsetCnt = p->min_subset_cnt;
        c.scale = p->scale_factor;
        c.maximizePositives = p->maximize_positives;
        c.semantics = p->semantics;
        c.closeThreshold = std::sqrt((double)(c.w * c.h)) / 150;
        c.closeThresholdF = sqrtf((float)(c.w * c.h)) / 150;
        c.closeScore = (float)((double)p->default_nms_threshold + 1e-6);
        c.maxPeople = (t->pairs * p->max_peaks + 3) / 4 * 4;
        int cap = 1;
        while (cap < p->max_peaks * p->max_peaks)
            cap <<= 1;
        c.sortCap = cap;
        return OPB200_OK;
    }
}

extern "C" size_t opb200_connect_scratch_bytes(const opb200_connect_params* p, int n_frames)
{
    if (!p || n_frames <= 0)
        return 0;
    const opb::ModelTables* t = opb::tablesFor(p->pose_model);
    if (!t)
        return 0;
    if (p->semantics == OPB200_SEMANTICS_CUDA)
    {
        // score cube + the connection keys of every cube cell, rounded up to a power of two for the sort
        const size_t cells = (size_t)t->pairs * p->max_peaks * p->max_peaks;
        size_t keys = 1;
        while (keys < cells)
            keys <<= 1;
        return (size_t)n_frames * (sizeof(float) * cells + sizeof(opb::ConnKey) * keys) + 256;
    }
    // connection lists, then the candidate counters and key lists of pafCandKernel
    const size_t conn = (sizeof(int) * (size_t)n_frames * t->pairs * (1 + 3 * (size_t)p->max_peaks) + 255) & ~(size_t)255;
    const size_t counts = (sizeof(int) * (size_t)n_frames * t->pairs + 255) & ~(size_t)255;
    return conn + counts + sizeof(unsigned long long) * (size_t)n_frames * t->pairs * p->max_peaks * p->max_peaks;
}

extern "C" opb200_status opb200_paf_scores(float* cube, const float* heat, const float* peaks,
                                           const int heat_shape[4], const opb200_connect_params* p, void* stream)
{
    using namespace opb;
    OPB_REQUIRE(cube && heat && peaks, "paf_scores: null argument");
    ConnectArgs c;
    const opb200_status s = fillArgs(c, p, heat_shape);
    if (s != OPB200_OK)
        return s;
    c.heat = heat;
    c.peaks = peaks;
    c.cube = cube;
    c.candKeys = nullptr;
    c.candCount = nullptr;
    const dim3 grid((c.maxPeaks * c.maxPeaks + 255) / 256, c.t.pairs, heat_shape[0]);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (p->semantics == OPB200_SEMANTICS_CPU)
        pafCubeKernel<OPB200_SEMANTICS_CPU><<<grid, 256, 0, st>>>(c);
    else
        pafCubeKernel<OPB200_SEMANTICS_CUDA><<<grid, 256, 0, st>>>(c);
    OPB_LAUNCH_CHECK();
    return OPB200_OK;
}

extern "C" opb200_status opb200_connect(float* out_keypoints, float* out_scores, int* out_count, int capacity,
                                        void* scratch, const float* heat, const float* peaks,
                                        const int heat_shape[4], const opb200_connect_params* p, void* stream)
{
    using namespace opb;
    OPB_REQUIRE(out_keypoints && out_scores && out_count && scratch && heat && peaks, "connector: null argument");
    OPB_REQUIRE(capacity > 0, "connector: capacity must be positive");
    ConnectArgs c;
    const opb200_status s = fillArgs(c, p, heat_shape);
    if (s != OPB200_OK)
        return s;
    c.heat = heat;
    c.peaks = peaks;
    c.cube = nullptr;
    c.candKeys = nullptr;
    c.candCount = nullptr;
    c.conn = static_cast<int*>(scratch);
    c.outKeypoints = out_keypoints;
    c.outScores = out_scores;
    c.outCount = out_count;
    c.capacity = capacity;
    const int nFrames = heat_shape[0];
    OPB_REQUIRE(nFrames <= 65535, "connector: too many frames for one launch");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (p->semantics == OPB200_SEMANTICS_CUDA)
    {
        // CUDA-build semantics: CUDA-rule score cube -> global sort over all limbs -> five-case assignment
        const size_t cells = (size_t)c.t.pairs * c.maxPeaks * c.maxPeaks;
        size_t keysPerFrame = 1;
        while (keysPerFrame < cells)
            keysPerFrame <<= 1;
        c.cube = static_cast<float*>(scratch);
        const size_t cubeBytes = (sizeof(float) * cells * nFrames + 255) & ~(size_t)255;
        ConnKey* keys = reinterpret_cast<ConnKey*>(static_cast<unsigned char*>(scratch) + cubeBytes);
        const dim3 grid((c.maxPeaks * c.maxPeaks + 255) / 256, c.t.pairs, nFrames);
        pafCubeKernel<OPB200_SEMANTICS_CUDA><<<grid, 256, 0, st>>>(c);
        OPB_LAUNCH_CHECK();
        c.maxPeople = (c.t.parts * c.maxPeaks / 2 + 4) / 4 * 4;   // a new person takes two unassigned peaks
        const size_t state = sizeof(float) * c.maxPeople + sizeof(unsigned short) * (size_t)c.maxPeople * c.t.parts +
                             sizeof(short) * (size_t)c.t.parts * c.maxPeaks + 2 * (size_t)c.maxPeople;
        const size_t smemC = std::max(state, sizeof(ConnKey) * (size_t)kCudaSortSmem);
        OPB_REQUIRE(smemC <= 200 * 1024, "connector: max_peaks too large for the CUDA-semantics assembly");
        OPB_CUDA(cudaFuncSetAttribute(assembleCudaKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemC));
        assembleCudaKernel<<<nFrames, kAsmThreads, smemC, st>>>(c, keys, keysPerFrame);
        OPB_LAUNCH_CHECK();
        return OPB200_OK;
    }

    {
        const size_t connBytes = (sizeof(int) * (size_t)nFrames * c.t.pairs * (1 + 3 * (size_t)c.maxPeaks) + 255) & ~(size_t)255;
        const size_t countBytes = (sizeof(int) * (size_t)nFrames * c.t.pairs + 255) & ~(size_t)255;
        c.candCount = reinterpret_cast<int*>(static_cast<unsigned char*>(scratch) + connBytes);
        c.candKeys = reinterpret_cast<unsigned long long*>(static_cast<unsigned char*>(scratch) + connBytes + countBytes);
        OPB_CUDA(cudaMemsetAsync(c.candCount, 0, sizeof(int) * (size_t)nFrames * c.t.pairs, st));
        const int blocksX = std::min(8, (c.maxPeaks * c.maxPeaks + kCandThreads - 1) / kCandThreads);
        const dim3 grid(blocksX, c.t.pairs, nFrames);
        pafCandKernel<<<grid, kCandThreads, 0, st>>>(c);
        OPB_LAUNCH_CHECK();
    }
    const size_t smemA = sizeof(unsigned long long) * (size_t)c.sortCap;
    OPB_CUDA(cudaFuncSetAttribute(pafMatchKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemA));
    pafMatchKernel<<<dim3(c.t.pairs, nFrames), kMatchThreads, smemA, st>>>(c);
    OPB_LAUNCH_CHECK();

    const size_t smemB = (sizeof(float) + 1 + c.t.parts) * (size_t)c.maxPeople;
    OPB_CUDA(cudaFuncSetAttribute(assembleKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemB));
    assembleKernel<<<nFrames, kAsmThreads, smemB, st>>>(c);
    OPB_LAUNCH_CHECK();
    return OPB200_OK;
}
