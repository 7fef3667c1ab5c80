// Implicit-GEMM convolution for sm_100a: TMA -> swizzled smem -> tcgen05.mma (bf16 x bf16 -> fp32 in TMEM)
// -> fused epilogue (bias + ReLU/PReLU/identity, bf16 NHWC store at a channel offset = free concat, optional
// fp32 NCHW copy for net_output members).
//
// Replaces, for the whole graph, what the reference runs per layer as cuDNN conv + bias + activation kernel +
// concat copy (3rdparty/caffe/src/caffe/layers/cudnn_conv_layer.cu:21-37, relu_layer.cu:9, prelu_layer.cu:11-18,
// concat_layer.cu:9) and on the CPU as im2col + sgemm (base_conv_layer.cpp:258-281).
//
// GEMM view: M = output pixels, N = Cout, K = Cin x taps. No im2col buffer exists anywhere.
//   * M tile = 8 x 16 pixel patch of one frame (128 rows). Per 64-channel chunk and per filter COLUMN kx the
//     producer loads ONE box (64 ch, 8 px, 16 + k - 1 rows) shifted by (kx - pad, -pad): the k filter ROWS ky are
//     then just smem descriptor offsets of ky * 1024 B into that box (8 px x 128 B = one swizzle group), so an
//     activation byte is fetched k times per layer instead of k*k times. TMA zero-fills out-of-image elements
//     (= the conv's zero padding) and channels beyond a view's extent (= Cin padding to 64).
//   * Each CTA works on a PAIR of M tiles against the same weight tile: every B (weight) stage feeds 8 MMAs
//     (2 tiles x 4 k-steps), halving weight traffic per FLOP. 2 tiles x 2 accumulator stages x BN columns of TMEM.
// The L2->SM operand stream, not the tensor pipe, bounds these layers (Cout <= 128): both measures cut bytes/FLOP.
//
// Roles (384 threads, persistent over work items): warp 0 = TMA producer, warp 1 = MMA issuer (one thread),
// warp 2 = TMEM allocator, warps 4-11 = epilogue (TMEM lane <-> pixel of the tile; two warps per lane quadrant split
// the columns: short-K layers -- 1x1, Cin = 64 -- are paced by the epilogue, not by the tensor pipe).
#include "../common.cuh"
#include "conv_params.h"
#include "sm100.cuh"
#include <cuda_bf16.h>
#include <cstdio>
#include <cstdlib>

namespace opb
{
    using namespace sm100;

    constexpr int kConvThreads = 384;     // 4 role warps + 8 epilogue warps
    constexpr int kMaxAStages = 4, kMaxBStages = 18;
    constexpr int kSmemLimit = 227 * 1024;

    template <int BN>
    struct ConvCfg
    {
        static constexpr int kBBytes = BN * kConvBlockK * 2;
        static constexpr int kStagingBytes = (BN / 32) * (kConvBlockM * 64);   // one bf16 output tile, 32-ch pieces
        static constexpr int kAccStride = BN <= 32 ? 32 : (BN <= 64 ? 64 : 128);
        static constexpr int kTmemCols = 4 * kAccStride;                       // 2 tiles x 2 accumulator stages
        static_assert(BN <= 128 && BN % 16 == 0, "BN must be a multiple of 16, at most 128");
    };

    struct ConvAux   // barriers and small tables, placed after the pipeline buffers
    {
        uint64_t aFull[kMaxAStages], aEmpty[kMaxAStages];
        uint64_t bFull[kMaxBStages], bEmpty[kMaxBStages];
        uint64_t tmemFull[2], tmemEmpty[2];
        uint64_t stFull[2], stEmpty[2];       // output staging buffers: written by the epilogue warps, drained by the store warp
        uint32_t tmemBase;
        uint32_t pad;
        alignas(16) float bias[128];          // bias of the current Cout tile (read as float4)
        float slope[128];                     // negative-side slope of the current Cout tile
    };
    constexpr int kAuxBytes = 2048;
    static_assert(sizeof(ConvAux) <= kAuxBytes, "aux block too large");

    // Debug timeline (OPB200_CONV_TRACE) and ablation switches (OPB200_CONV_DEBUG): compiled only into trace builds
    // (make TRACE=1 -> -DOPB200_CONV_TRACE_BUILD); the release kernel carries neither the stamps nor the mode tests.
#ifdef OPB200_CONV_TRACE_BUILD
    constexpr bool kConvDebug = true;
#else
    constexpr bool kConvDebug = false;
#endif
    // clock64 stamps of CTA 0's roles per work item, printed by the kernel itself
    __device__ unsigned long long gConvTrace[kConvDebug ? 16 * 16 + 2 : 1];
    __device__ __forceinline__ void traceStamp(bool on, int slot, int work)
    {
        if (kConvDebug && on && work < 16)
            reinterpret_cast<volatile unsigned long long*>(gConvTrace)[slot * 16 + work] = clock64();
    }

    struct TileCoord
    {
        int x0, y0, n;
    };
    // M tile index -> pixel origin. Indices past the end give a tile entirely outside the image: its loads are
    // zero-filled and its stores clipped away by TMA (used for the odd tile of the last pair).
    __device__ __forceinline__ TileCoord tileCoord(const ConvParams& p, int mt)
    {
        TileCoord c;
        if (mt >= p.numMTiles)
        {
            c.x0 = 0;
            c.y0 = p.tilesY * kConvTileH;
            c.n = p.N - 1;
            return c;
        }
        const int perFrame = p.tilesX * p.tilesY;
        c.n = mt / perFrame;
        const int r = mt - c.n * perFrame;
        c.y0 = (r / p.tilesX) * kConvTileH;
        c.x0 = (r % p.tilesX) * kConvTileW;
        return c;
    }

    // PAIR: two CTAs of a cluster (same TPC) work as one cta_group::2 unit on FOUR M tiles (two each) of the same Cout
    // tile: every CTA loads only half of each weight stage (BN/2 rows), the leader CTA issues M = 256 MMAs that read
    // both halves, and each CTA's TMEM receives (and its epilogue drains) its own two tiles. Halves the weight
    // traffic from L2 per SM, which is what bounds the single-CTA kernel on the large layers.
    template <int BN, bool PAIR>
    __global__ void __launch_bounds__(kConvThreads, 1) convIgemmKernel(const __grid_constant__ ConvParams p)
    {
        using Cfg = ConvCfg<BN>;
        constexpr int kBStageBytes = PAIR ? Cfg::kBBytes / 2 : Cfg::kBBytes;   // this CTA's part of a weight stage
        const uint32_t ctaRank = PAIR ? clusterCtaRank() : 0u;
        extern __shared__ uint8_t smemRaw[];
        uint8_t* smem = smemRaw + ((1024u - (smemAddr(smemRaw) & 1023u)) & 1023u);   // 1024-byte aligned
        const int aStageBytes = 2 * p.aTileBytes;
        const uint32_t boxBytes = (uint32_t)p.boxW * (kConvTileH + p.k - 1) * 128u;   // what one TMA box delivers
        const uint32_t groupPitch = (uint32_t)p.boxW * 128u;                         // bytes between pixel rows
        uint8_t* sA = smem;                                         // aStages x (tile0 | tile1)
        uint8_t* sB = sA + p.aStages * aStageBytes;                 // bStages x (BN x 128 B)
        const int bStageBytes = p.bGroup * kBStageBytes;            // one weight stage = bGroup K blocks
        uint8_t* sOut = sB + p.bStages * bStageBytes;               // staging for the TMA store
        ConvAux* aux = reinterpret_cast<ConvAux*>(sOut + p.stagingBufs * Cfg::kStagingBytes);

        // Programmatic dependent launch: let the next layer's CTAs start their prologue on SMs this grid has
        // already vacated; they block in griddepcontrol.wait (below) until this grid has completed and flushed.
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);   // warp-uniform FOR THE COMPILER: role branches
                                                                                  // become uniform branches, loop state may live in uniform registers
        const int lane = threadIdx.x & 31;
        // a work item = (group of M tiles, Cout tile); the group is a pair per CTA, i.e. a quad per CTA pair
        const int tpi = p.tilesPerItem;                             // M tiles per CTA and work item: 2, or 1 (small batches)
        const bool two = tpi == 2;
        const int tilesPerGroup = (PAIR ? 2 : 1) * tpi;             // a group = the tiles of one CTA, or of one CTA pair
        const int numGroups = (p.numMTiles + tilesPerGroup - 1) / tilesPerGroup;
        const int numWork = numGroups * p.nTilesN;
        const int workStart = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
        const int workStride = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;

        if (warp == 0 && lane == 0)
        {
            for (int s = 0; s < p.nSeg; ++s)
                prefetchTensorMap(&p.aMap[s]);
            prefetchTensorMap(&p.bMap);
            prefetchTensorMap(&p.oMap);
        }
        if (warp == 1 && lane == 0)
        {
            for (int i = 0; i < p.aStages; ++i)
            {
                mbarInit(&aux->aFull[i], 1);
                mbarInit(&aux->aEmpty[i], 1);
            }
            for (int i = 0; i < p.bStages; ++i)
            {
                mbarInit(&aux->bFull[i], 1);
                mbarInit(&aux->bEmpty[i], 1);
            }
            for (int i = 0; i < 2; ++i)
            {
                mbarInit(&aux->tmemFull[i], 1);
                mbarInit(&aux->tmemEmpty[i], PAIR ? 512 : 256);   // every epilogue thread (of both CTAs) arrives
                mbarInit(&aux->stFull[i], 8);                     // one arrival per epilogue warp
                mbarInit(&aux->stEmpty[i], 1);
            }
            fenceBarrierInit();
        }
        if (warp == 2)
        {
            if (PAIR)
                tmemAllocPair<Cfg::kTmemCols>(&aux->tmemBase);
            else
                tmemAlloc<Cfg::kTmemCols>(&aux->tmemBase);
        }
        tcFenceBefore();
        if (PAIR)
            clusterSync();      // the peer's barriers must be initialised before anything signals them
        else
            __syncthreads();
        tcFenceAfter();
        const uint32_t tmemBase = aux->tmemBase;
        // everything above touched only kernel parameters, smem and TMEM; from here on we read the previous
        // layer's output and overwrite buffers it may still be reading
        asm volatile("griddepcontrol.wait;" ::: "memory");
        const bool traceOn = kConvDebug && (p.debugMode & 4) && blockIdx.x == 0;
        if (traceOn && threadIdx.x == 0)
            reinterpret_cast<volatile unsigned long long*>(gConvTrace)[256] = clock64();

        if (warp == 0)
        {
            // ============================ TMA producer: activations (A) ===============================
            // the whole (converged) warp runs the loop and waits; the elected lane issues. Activations and weights
            // have their own producer warps: the weight ring is short (it follows the MMAs closely), and a single
            // producer blocked on it could not request the NEXT item's activation boxes early enough -- a box takes
            // 3-4 k cycles to arrive when the TMA queue is busy (OPB200_CONV_TRACE timeline).
            const bool leader = electOne();
            int as = 0;
            uint32_t aPhase = 0;
            for (int work = workStart; work < numWork; work += workStride)
            {
                const int item = p.reverse ? numWork - 1 - work : work;
                const int pair = PAIR ? 2 * (item / p.nTilesN) + (int)ctaRank : item / p.nTilesN;
                const TileCoord t0 = tileCoord(p, tpi * pair), t1 = tileCoord(p, tpi * pair + 1);
                const int wi = (work - workStart) / workStride;
                traceStamp(traceOn && leader, 0, wi);
                for (int s = 0; s < p.nSeg; ++s)
                    for (int c = 0; c < p.segChunks[s]; ++c)
                        for (int kx = 0; kx < p.k; kx += p.kxSpan)
                        {
                            // one box serves kxSpan filter columns (and all k filter rows)
                            mbarWait(&aux->aEmpty[as], aPhase ^ 1);
                            uint8_t* dst = sA + as * aStageBytes;
                            if (leader)
                            {
                                if (PAIR)
                                {
                                    // both CTAs' boxes are counted on the leader CTA's barrier
                                    const uint32_t bar = mapaShared(smemAddr(&aux->aFull[as]), 0);
                                    if (ctaRank == 0)
                                        mbarExpectTx(&aux->aFull[as], 2u * (uint32_t)tpi * boxBytes);
                                    tmaLoad4dPair(dst, &p.aMap[s], bar, c * kConvBlockK, t0.x0 + kx - p.pad,
                                                  t0.y0 - p.pad, t0.n);
                                    if (two)
                                        tmaLoad4dPair(dst + p.aTileBytes, &p.aMap[s], bar, c * kConvBlockK,
                                                      t1.x0 + kx - p.pad, t1.y0 - p.pad, t1.n);
                                }
                                else
                                {
                                    mbarExpectTx(&aux->aFull[as], (uint32_t)tpi * boxBytes);
                                    tmaLoad4d(dst, &p.aMap[s], &aux->aFull[as], c * kConvBlockK,
                                              t0.x0 + kx - p.pad, t0.y0 - p.pad, t0.n);
                                    if (two)
                                        tmaLoad4d(dst + p.aTileBytes, &p.aMap[s], &aux->aFull[as], c * kConvBlockK,
                                                  t1.x0 + kx - p.pad, t1.y0 - p.pad, t1.n);
                                }
                            }
                            __syncwarp();
                            if (++as == p.aStages)
                            {
                                as = 0;
                                aPhase ^= 1;
                            }
                        }
                traceStamp(traceOn && leader, 1, wi);
            }
        }
        else if (warp == 3)
        {
            // ============================== TMA producer: weights (B) =================================
            const bool leader = electOne();
            int bs = 0;
            uint32_t bPhase = 0;
            for (int work = workStart; work < numWork; work += workStride)
            {
                if (p.bResident && work != workStart)
                    break;                 // the whole weight tile already sits in shared memory
                const int item = p.reverse ? numWork - 1 - work : work;
                const int nt = item % p.nTilesN;
                const int kBlocks = p.chunksTotal * p.k * p.k;
                // one box = bGroup filter rows (consecutive K blocks) of this Cout tile
                for (int kb = 0; kb < kBlocks; kb += p.bGroup)
                {
                    mbarWait(&aux->bEmpty[bs], bPhase ^ 1);
                    if (leader)
                    {
                        if (PAIR)
                        {
                            // this CTA's half of the Cout tile (bMap box = BN/2 rows)
                            if (ctaRank == 0)
                                mbarExpectTx(&aux->bFull[bs], p.bGroup * Cfg::kBBytes);
                            tmaLoad3dPair(sB + bs * bStageBytes, &p.bMap, mapaShared(smemAddr(&aux->bFull[bs]), 0), 0,
                                          nt * BN + (int)ctaRank * (BN / 2), kb);
                        }
                        else
                        {
                            mbarExpectTx(&aux->bFull[bs], p.bGroup * Cfg::kBBytes);
                            tmaLoad3d(sB + bs * bStageBytes, &p.bMap, &aux->bFull[bs], 0, nt * BN, kb);
                        }
                    }
                    __syncwarp();
                    if (++bs == p.bStages)
                    {
                        bs = 0;
                        bPhase ^= 1;
                    }
                }
            }
        }
        else if (warp == 1)
        {
            // ===================================== MMA issuer ========================================
            // (pair mode: the leader CTA's warp issues for both CTAs; the peer's warp has nothing to do)
            const bool leader = electOne();
            if (!PAIR || ctaRank == 0)
            {
                constexpr uint32_t idesc = instrDescBf16(PAIR ? 2 * kConvBlockM : kConvBlockM, BN);
                int as = 0, bs = 0, acc = 0;
                uint32_t aPhase = 0, bPhase = 0, accPhase = 0;
                const int axPerChunk = p.k / p.kxSpan;
                const int numAx = p.chunksTotal * axPerChunk;           // A stages consumed per work item
                const uint64_t bBase = smemDescSw128(smemAddr(sB));
                const uint64_t kyStep = (uint64_t)(p.boxW * 8);         // one box pixel row down, in 16-byte units
                for (int work = workStart; work < numWork; work += workStride)
                {
                    const int wi = (work - workStart) / workStride;
                    traceStamp(traceOn && leader, 2, wi);
                    mbarWait(&aux->tmemEmpty[acc], accPhase ^ 1);
                    tcFenceAfter();
                    traceStamp(traceOn && leader, 3, wi);
                    const uint32_t tmemD0 = tmemBase + (acc * 2) * Cfg::kAccStride;
                    const uint32_t tmemD1 = tmemD0 + Cfg::kAccStride;
                    int chunk = 0, axInChunk = 0;
                    for (int ax = 0; ax < numAx; ++ax)
                    {
                        // 16-channel k-steps that hold real channels in this chunk (the rest is TMA zero fill)
                        const int steps = p.chunkSteps[chunk];
                        if (++axInChunk == axPerChunk)
                        {
                            axInChunk = 0;
                            ++chunk;
                        }
                        mbarWait(&aux->aFull[as], aPhase);
                        if (ax == 0)
                            traceStamp(traceOn && leader, 4, wi);
                        // tap (kxi, ky) = the 16 pixel rows starting at box pixel (kxi, ky): 8-pixel groups groupPitch
                        // bytes apart, the first one (ky * boxW + kxi) * 128 bytes into the box. Descriptors are built
                        // once per stage and advanced by adding 16-byte units to the start-address field.
                        const uint64_t aBase0 = smemDescSw128(smemAddr(sA + as * aStageBytes), groupPitch);
                        const uint64_t aBase1 = aBase0 + (uint64_t)(p.aTileBytes >> 4);
                        // one weight box = bGroup filter rows of one filter column: wait for it, issue all its MMAs in one
                        // statement (sm100.cuh: mmaGroupBf16), hand the slot back
                        uint32_t accFirst = ax != 0;
                        for (int kxi = 0; kxi < p.kxSpan; ++kxi)
                        {
                            uint64_t aDesc0 = aBase0 + (uint64_t)(kxi * 8);
                            uint64_t aDesc1 = aBase1 + (uint64_t)(kxi * 8);
                            for (int ky = 0; ky < p.k; ky += p.bGroup)
                            {
                                mbarWait(&aux->bFull[bs], bPhase);
                                tcFenceAfter();
                                const uint64_t bDesc = bBase + (uint64_t)(bs * p.bGroup * (kBStageBytes >> 4));
                                if (leader && !(kConvDebug && (p.debugMode & 2)))
                                {
                                    if (p.bGroup == 3)
                                        mmaGroupBf16<3, PAIR>(tmemD0, tmemD1, aDesc0, aDesc1, bDesc, (uint32_t)kyStep,
                                                              (uint32_t)(kBStageBytes >> 4), idesc, accFirst, (uint32_t)steps,
                                                              two ? 1u : 0u);
                                    else
                                        mmaGroupBf16<1, PAIR>(tmemD0, tmemD1, aDesc0, aDesc1, bDesc, (uint32_t)kyStep,
                                                              (uint32_t)(kBStageBytes >> 4), idesc, accFirst, (uint32_t)steps,
                                                              two ? 1u : 0u);
                                }
                                if (leader && !p.bResident)   // weight slot free (in both CTAs) once these MMAs retire
                                {
                                    if (PAIR)
                                        mmaCommitPair(&aux->bEmpty[bs]);
                                    else
                                        mmaCommit(&aux->bEmpty[bs]);
                                }
                                __syncwarp();
                                accFirst = 1;
                                // resident weights: the slots are never recycled, their (only) phase stays complete
                                if (++bs == p.bStages)
                                {
                                    bs = 0;
                                    bPhase ^= p.bResident ? 0u : 1u;
                                }
                                aDesc0 += kyStep * (uint64_t)p.bGroup;
                                aDesc1 += kyStep * (uint64_t)p.bGroup;
                            }
                        }
                        if (leader)                               // activation slot free
                        {
                            if (PAIR)
                                mmaCommitPair(&aux->aEmpty[as]);
                            else
                                mmaCommit(&aux->aEmpty[as]);
                        }
                        __syncwarp();
                        if (++as == p.aStages)
                        {
                            as = 0;
                            aPhase ^= 1;
                        }
                    }
                    traceStamp(traceOn && leader, 5, wi);
                    if (leader)                                   // both accumulators complete -> epilogue(s)
                    {
                        if (PAIR)
                            mmaCommitPair(&aux->tmemFull[acc]);
                        else
                            mmaCommit(&aux->tmemFull[acc]);
                    }
                    __syncwarp();
                    if (++acc == 2)
                    {
                        acc = 0;
                        accPhase ^= 1;
                    }
                }
            }
        }
        else if (warp >= 4)
        {
            // ===================================== epilogue ==========================================
            // 8 warps: warp (id % 4) may only touch TMEM lanes [32*quad, 32*quad+32); the two warps of a quadrant
            // split the 32-column pieces of the tile between them (half 0: even pieces, half 1: odd pieces), and
            // each half stages and stores its own pieces (own named barrier, own bulk-store group).
            const int quad = warp & 3;
            const int half = (warp - 4) >> 2;
            const int t = quad * 32 + lane;           // 0..127 = TMEM lane = pixel of the tile
            const int et = threadIdx.x - 128;         // 0..255 over all epilogue threads
            const int tw = t % kConvTileW, th = t / kConvTileW;
            constexpr int kPieces = BN / 32;
            int acc = 0;
            uint32_t accPhase = 0;
            int lastNt = -1;
            const bool stores = (p.storeBf16 | p.poolFused) != 0;
            uint32_t stUses[2] = {0u, 0u};
            for (int work = workStart; work < numWork; work += workStride)
            {
                const int item = p.reverse ? numWork - 1 - work : work;
                const int nt = item % p.nTilesN;
                const int pair = PAIR ? 2 * (item / p.nTilesN) + (int)ctaRank : item / p.nTilesN;
                if (nt != lastNt)
                {
                    namedBarrier(1, 256);             // everybody is done with the previous table
                    for (int i = et; i < BN; i += 256)
                    {
                        aux->bias[i] = __ldg(p.bias + nt * BN + i);
                        aux->slope[i] = __ldg(p.slope + nt * BN + i);
                    }
                    lastNt = nt;
                    namedBarrier(1, 256);
                }
                const int wi = (work - workStart) / workStride;
                mbarWait(&aux->tmemFull[acc], accPhase);
                tcFenceAfter();
                traceStamp(traceOn && et == 0, 6, wi);
                traceStamp(traceOn && lane == 0 && wi == 1, 15, warp - 4);
                bool released = false;
#pragma unroll 1
                for (int j = 0; j < tpi; ++j)
                {
                    const int mt = tpi * pair + j;
                    if (mt >= p.numMTiles)
                        break;                         // odd tile count: the last pair has no second tile
                    const TileCoord tc = tileCoord(p, mt);
                    // staging buffer of this tile: one per tile of the item, or alternating between single-tile items. It is
                    // free once the store warp's TMA stores of its previous use have read it.
                    const int sbuf = p.stagingBufs == 2 ? (two ? j : (wi & 1)) : 0;
                    if (stores)
                        mbarWait(&aux->stEmpty[sbuf], ((stUses[sbuf] & 1u) ^ 1u));
                    traceStamp(traceOn && et == 0, 8 + 2 * j, wi);
                    const uint32_t taddr =
                        tmemBase + (acc * 2 + j) * Cfg::kAccStride + ((uint32_t)(quad * 32) << 16);
                    const int x = tc.x0 + tw, y = tc.y0 + th;
                    const bool inImage = (x < p.W) && (y < p.H);
                    const bool isSeparator = p.flatPitch != 0 && (x % p.flatPitch) >= p.flatW;
#pragma unroll 1
                    for (int q = half; q < kPieces && !(kConvDebug && (p.debugMode & 1)); q += 2)
                    {
                        uint32_t v[32];
                        tmemLoad32(taddr + q * 32, v);
                        const uint32_t biasAddr = smemAddr(&aux->bias[q * 32]);
                        float f[32];
                        float4 b4[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            b4[i] = ldsF4(biasAddr + i * 16);
                        tmemLoadWait();

                        if (p.actMode == 1)            // ReLU everywhere (the VGG trunk): add + max
                        {
#pragma unroll
                            for (int i = 0; i < 8; ++i)
                            {
                                // (the max with 0 happens in the bf16 conversion below; ReLU layers have no fp32 output)
                                f[4 * i + 0] = __uint_as_float(v[4 * i + 0]) + b4[i].x;
                                f[4 * i + 1] = __uint_as_float(v[4 * i + 1]) + b4[i].y;
                                f[4 * i + 2] = __uint_as_float(v[4 * i + 2]) + b4[i].z;
                                f[4 * i + 3] = __uint_as_float(v[4 * i + 3]) + b4[i].w;
                            }
                        }
                        else
                        {
                            const uint32_t slopeAddr = smemAddr(&aux->slope[q * 32]);
#pragma unroll
                            for (int i = 0; i < 8; ++i)
                            {
                                const float4 s4 = ldsF4(slopeAddr + i * 16);
                                const float a0 = __uint_as_float(v[4 * i + 0]) + b4[i].x;
                                const float a1 = __uint_as_float(v[4 * i + 1]) + b4[i].y;
                                const float a2 = __uint_as_float(v[4 * i + 2]) + b4[i].z;
                                const float a3 = __uint_as_float(v[4 * i + 3]) + b4[i].w;
                                f[4 * i + 0] = a0 > 0.f ? a0 : a0 * s4.x;
                                f[4 * i + 1] = a1 > 0.f ? a1 : a1 * s4.y;
                                f[4 * i + 2] = a2 > 0.f ? a2 : a2 * s4.z;
                                f[4 * i + 3] = a3 > 0.f ? a3 : a3 * s4.w;
                            }
                        }
                        if (p.out32 != nullptr && inImage)
                        {
                            const int cBase = nt * BN + q * 32;
                            int fn = tc.n, fx = x, fw = p.W;
                            if (p.flatPitch)           // flat level: x runs over all frames and their separators
                            {
                                fn = x / p.flatPitch;
                                fx = x - fn * p.flatPitch;
                                fw = p.flatW;
                            }
                            float* o = p.out32 + (((size_t)fn * p.out32C + p.out32Off + cBase) * p.H + y) * fw + fx;
                            const size_t plane = (size_t)p.H * fw;
                            if (fx < fw && fn < (p.flatPitch ? p.flatN : p.N))   // not a separator column
                            {
#pragma unroll
                                for (int i = 0; i < 32; ++i)
                                    if (cBase + i < p.cout)
                                        o[i * plane] = f[i];
                            }
                        }
                        if (p.storeBf16 | p.poolFused)
                        {
                            uint32_t pk[16];           // channels (2i, 2i+1) as packed bf16
                            if (p.actMode == 1)
                            {
#pragma unroll
                                for (int i = 0; i < 16; ++i)
                                    pk[i] = packBf16Relu(f[2 * i], f[2 * i + 1]);
                            }
                            else
                            {
#pragma unroll
                                for (int i = 0; i < 16; ++i)
                                    pk[i] = packBf16(f[2 * i], f[2 * i + 1]);
                            }
                            if (isSeparator)           // flat level: the zero column between two frames stays zero
                            {
#pragma unroll
                                for (int i = 0; i < 16; ++i)
                                    pk[i] = 0u;
                            }
                            const uint32_t pieceAddr = smemAddr(sOut + sbuf * Cfg::kStagingBytes + q * (kConvBlockM * 64));
                            if (p.poolFused)
                            {
                                // fused 2x2 / stride 2 MAX pooling (Caffe ceil mode: windows are clipped at the image
                                // edge): the tile origin is even, so a window never straddles tiles, and its four pixels
                                // are lanes l, l^1 (x) and l^8 (y) of this warp. Rounding to bf16 is monotonic, so the
                                // max is taken on the packed values (same bits as max-then-round). The four lanes split
                                // the 32 channels: each exchange halves what a lane keeps, and every lane ends up with 8
                                // channels (one 16-byte chunk) of the pooled pixel. Pixels outside the image are -inf.
                                if (!inImage)
                                {
#pragma unroll
                                    for (int i = 0; i < 16; ++i)
                                        pk[i] = 0xff80ff80u;
                                }
                                const bool xOdd = lane & 1, yOdd = lane & 8;
                                uint32_t h[8], g[4];
#pragma unroll
                                for (int i = 0; i < 8; ++i)
                                {
                                    const uint32_t got = __shfl_xor_sync(0xffffffffu, xOdd ? pk[i] : pk[8 + i], 1);
                                    h[i] = maxBf16x2(xOdd ? pk[8 + i] : pk[i], got);
                                }
#pragma unroll
                                for (int i = 0; i < 4; ++i)
                                {
                                    const uint32_t got = __shfl_xor_sync(0xffffffffu, yOdd ? h[i] : h[4 + i], 8);
                                    g[i] = maxBf16x2(yOdd ? h[4 + i] : h[i], got);
                                }
                                // row (y/2)*4 + x/2 of a 32-row piece; 16-byte chunk c lands at c ^ ((r >> 1) & 3)
                                const int r = (th >> 1) * (kConvTileW / 2) + (tw >> 1);
                                const int c = (xOdd ? 2 : 0) + (yOdd ? 1 : 0);
                                stsU4(pieceAddr + r * 64 + ((c ^ ((r >> 1) & 3)) << 4), make_uint4(g[0], g[1], g[2], g[3]));
                            }
                            else
                            {
                                // row t of a 128 x 64-byte piece; 16-byte chunk c lands at c ^ ((t >> 1) & 3) (SWIZZLE_64B)
                                const uint32_t rowAddr = pieceAddr + t * 64;
                                const int sw = (t >> 1) & 3;
#pragma unroll
                                for (int c = 0; c < 4; ++c)
                                    stsU4(rowAddr + ((c ^ sw) << 4),
                                          make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]));
                            }
                        }
                    }
                    traceStamp(traceOn && et == 0, 9 + 2 * j, wi);
                    traceStamp(traceOn && lane == 0 && wi == 1, j ? 12 : 14, warp - 4);
                    if (j == tpi - 1 || mt + 1 >= p.numMTiles)
                    {
                        // all TMEM reads of this thread are complete (wait::ld above): hand both accumulators back NOW,
                        // before the stores -- a TMA store queues behind every load the producer has in flight
                        // (thousands of cycles), and the MMA issuer needs only the accumulator, not the stores
                        tcFenceBefore();
                        if (PAIR)   // the MMA issuer lives in the leader CTA: both CTAs' epilogues report to its barrier
                            mbarArriveCluster(mapaShared(smemAddr(&aux->tmemEmpty[acc]), 0));
                        else
                            mbarArrive(&aux->tmemEmpty[acc]);
                        released = true;
                    }
                    if (stores)
                    {
                        // hand the staged tile to the store warp: a TMA store queues behind every load the producers have in
                        // flight (the issuing thread sits 1.5-4 k cycles in the instruction), which used to hold up all
                        // epilogue warps at the next barrier -- now only the store warp waits
                        fenceProxyAsyncShared();
                        __syncwarp();
                        if (lane == 0)
                            mbarArrive(&aux->stFull[sbuf]);
                        ++stUses[sbuf];
                        traceStamp(traceOn && lane == 0 && wi == 1 && j == 1, 13, warp - 4);
                    }
                }
                if (!released)   // no tile of this item belongs to this CTA (ragged end of a quad)
                {
                    tcFenceBefore();
                    if (PAIR)
                        mbarArriveCluster(mapaShared(smemAddr(&aux->tmemEmpty[acc]), 0));
                    else
                        mbarArrive(&aux->tmemEmpty[acc]);
                }
                traceStamp(traceOn && et == 0, 7, wi);
                if (++acc == 2)
                {
                    acc = 0;
                    accPhase ^= 1;
                }
            }
        }
        else if (warp == 2 && (p.storeBf16 | p.poolFused))
        {
            // ====================================== store warp ========================================
            // (the TMEM allocator warp, idle in between) walks the same tiles as the epilogue warps: waits for a staged
            // tile, issues its TMA stores, and frees the staging buffer once the stores have read it
            constexpr int kPieces = BN / 32;
            const bool leader = electOne();
            uint32_t stUses[2] = {0u, 0u};
            for (int work = workStart; work < numWork; work += workStride)
            {
                const int item = p.reverse ? numWork - 1 - work : work;
                const int nt = item % p.nTilesN;
                const int pair = PAIR ? 2 * (item / p.nTilesN) + (int)ctaRank : item / p.nTilesN;
                const int wi = (work - workStart) / workStride;
                for (int j = 0; j < tpi; ++j)
                {
                    const int mt = tpi * pair + j;
                    if (mt >= p.numMTiles)
                        break;
                    const TileCoord tc = tileCoord(p, mt);
                    const int sbuf = p.stagingBufs == 2 ? (two ? j : (wi & 1)) : 0;
                    mbarWait(&aux->stFull[sbuf], stUses[sbuf] & 1u);
                    if (leader)
                    {
                        // pooled: oMap describes the pooled blob with a (32 ch, 4, 8) box
                        const int sx = p.poolFused ? tc.x0 >> 1 : tc.x0, sy = p.poolFused ? tc.y0 >> 1 : tc.y0;
                        for (int q = 0; q < kPieces; ++q)
                            if (nt * BN + q * 32 < p.cout)
                            {
                                const void* src = sOut + sbuf * Cfg::kStagingBytes + q * (kConvBlockM * 64);
                                const int c0 = nt * BN + q * 32;
                                if (p.poolFused && p.poolOutFlat)   // per-frame map of a flat level: (c, x, n, y)
                                    tmaStore4d(&p.oMap, src, c0, sx, tc.n, sy);
                                else
                                    tmaStore4d(&p.oMap, src, c0, sx, sy, tc.n);
                            }
                        tmaStoreCommit();
                        tmaStoreWaitRead0();
                        mbarArrive(&aux->stEmpty[sbuf]);
                    }
                    __syncwarp();
                    ++stUses[sbuf];
                }
            }
            if (leader)
                tmaStoreWaitAll();
            __syncwarp();
        }
        tcFenceBefore();
        if (traceOn)
        {
            __syncthreads();
            if (threadIdx.x == 0)
            {
                const volatile unsigned long long* tr = reinterpret_cast<volatile unsigned long long*>(gConvTrace);
                const unsigned long long z = tr[256];
                const int n = (numWork - workStart + workStride - 1) / workStride;
                printf("conv trace: cout %d chunks %d k %d pair %d aStages %d bStages %d x group %d resident %d, %d work items, end at %llu clk\n",
                       p.cout, p.chunksTotal, p.k, (int)PAIR, p.aStages, p.bStages, p.bGroup, p.bResident, n, clock64() - z);
                for (int w = 0; w < n && w < 16; ++w)
                    printf("  item %2d: producer %6llu..%6llu | mma start %6llu tmem-free %6llu first-A %6llu issued %6llu | epilogue %6llu [t0 free %6llu math %6llu | t1 free %6llu math %6llu] %6llu\n",
                           w, tr[0 * 16 + w] - z, tr[1 * 16 + w] - z, tr[2 * 16 + w] - z, tr[3 * 16 + w] - z,
                           tr[4 * 16 + w] - z, tr[5 * 16 + w] - z, tr[6 * 16 + w] - z, tr[8 * 16 + w] - z, tr[9 * 16 + w] - z,
                           tr[10 * 16 + w] - z, tr[11 * 16 + w] - z, tr[7 * 16 + w] - z);
                for (int w = 0; w < 8; ++w)
                    printf("  item 1, epilogue warp %d: start %6llu, tile 0 math done %6llu, tile 1 math done %6llu, past barrier %6llu\n",
                           w + 4, tr[15 * 16 + w] - z, tr[14 * 16 + w] - z, tr[12 * 16 + w] - z, tr[13 * 16 + w] - z);
            }
        }
        if (PAIR)
        {
            clusterSync();      // neither CTA may exit (or free TMEM) while the other can still signal or read it
            if (warp == 2)
                tmemFreePair<Cfg::kTmemCols>(tmemBase);
        }
        else
        {
            __syncthreads();
            if (warp == 2)
                tmemFree<Cfg::kTmemCols>(tmemBase);
        }
    }

    // Pipeline depths for a given kernel size: 4 weight stages, as many activation stages as fit (2..4).
    static int convTileBytes(int k, int kxSpan)
    {
        return ((kConvTileW + kxSpan - 1) * (kConvTileH + k - 1) * 128 + 1023) & ~1023;
    }

    // bStagesAll = weight stages of one whole Cout tile (chunks * k * k); when there is a single Cout tile and they all
    // fit next to two activation stages, the weights are loaded once per CTA and stay resident.
    template <int BN, bool PAIR>
    static bool convSmemPlan(int k, int kxSpan, int bGroup, int bBlocksAll, bool singleNTile, int& aStages, int& bStages,
                             int& resident, int& smemBytes, int stagingBufs, bool deepB = false)
    {
        using Cfg = ConvCfg<BN>;
        const int bStage = bGroup * (PAIR ? Cfg::kBBytes / 2 : Cfg::kBBytes);   // per CTA: bGroup K blocks
        const int bStagesAll = bBlocksAll / bGroup;
        const int aStage = 2 * convTileBytes(k, kxSpan);
        // weight stages: as many as fit next to two activation stages, up to `want` K blocks (one K block is only
        // 8 MMAs of tensor time, so the stream needs depth to cover the TMA round trip), at least two stages
        int want = (kxSpan > 1 || BN <= 64 || PAIR) ? 6 : 4;
        if (const char* e = std::getenv("OPB200_BSTAGES"))
            want = std::atoi(e);
        want = (want + bGroup - 1) / bGroup;
        if (want < 2)
            want = 2;
        // small-batch launches (one single-tile work item per CTA): the MMA phase is a chain of weight-box round trips
        // (timeline: 146 clk per MMA with 2 stages in flight), so the ring takes the whole Cout tile if it fits
        if (deepB && bStagesAll > want)
            want = bStagesAll;
        const int fitB = (kSmemLimit - (stagingBufs * Cfg::kStagingBytes + kAuxBytes + 1024) - 2 * aStage) / bStage;
        bStages = want < fitB ? want : fitB;
        if (bStages > kMaxBStages)
            bStages = kMaxBStages;
        if (bStages < 2)
            return false;
        resident = 0;
        if (singleNTile && bStagesAll <= kMaxBStages &&
            stagingBufs * Cfg::kStagingBytes + kAuxBytes + 1024 + bStagesAll * bStage + 2 * aStage <= kSmemLimit)
        {
            bStages = bStagesAll;
            resident = 1;
        }
        const int fixed = stagingBufs * Cfg::kStagingBytes + kAuxBytes + 1024 /*alignment slack*/ + bStages * bStage;
        aStages = (kSmemLimit - fixed) / aStage;
        if (aStages > kMaxAStages)
            aStages = kMaxAStages;
        smemBytes = fixed + aStages * aStage;
        return aStages >= 2;
    }

    template <int BN, bool PAIR>
    static opb200_status launchConvT(ConvParams& p, int gridCap, cudaStream_t st)
    {
        int smemBytes = 0;
        const bool deepB = p.tilesPerItem == 1;
        // two output staging buffers when they fit next to a working pipeline (OPB200_CONV_STAGING=1: always one)
        static const int stagingEnv = std::getenv("OPB200_CONV_STAGING") ? std::atoi(std::getenv("OPB200_CONV_STAGING")) : 2;
        p.stagingBufs = 1;
        {
            int a2 = 0, b2 = 0, r2 = 0, s2 = 0, a1 = 0, b1 = 0, r1 = 0, s1 = 0;
            const bool ok1 = convSmemPlan<BN, PAIR>(p.k, p.kxSpan, p.bGroup, p.chunksTotal * p.k * p.k, p.nTilesN == 1, a1, b1, r1, s1, 1, deepB);
            const bool ok2 = stagingEnv == 2 &&
                             convSmemPlan<BN, PAIR>(p.k, p.kxSpan, p.bGroup, p.chunksTotal * p.k * p.k, p.nTilesN == 1, a2, b2, r2, s2, 2, deepB);
            // not at the price of resident weights, and only with the same weight depth
            if (ok1 && ok2 && r2 == r1 && b2 == b1)
                p.stagingBufs = 2;
        }
        if (!convSmemPlan<BN, PAIR>(p.k, p.kxSpan, p.bGroup, p.chunksTotal * p.k * p.k, p.nTilesN == 1, p.aStages, p.bStages, p.bResident,
                              smemBytes, p.stagingBufs, deepB))
            OPB_FAIL(OPB200_ERR_UNSUPPORTED, "conv: kernel size too large for the shared-memory pipeline");
        p.aTileBytes = convTileBytes(p.k, p.kxSpan);
        static const int debugMode = std::getenv("OPB200_CONV_DEBUG") ? std::atoi(std::getenv("OPB200_CONV_DEBUG")) : 0;
        p.debugMode = debugMode;
        // OPB200_CONV_TRACE=chunks:cout:k[:nth] prints the role timeline of CTA 0 for the nth such launch (default 3rd)
        if (const char* e = std::getenv("OPB200_CONV_TRACE"))
        {
            int tc = 0, to = 0, tk = 0, nth = 3;
            static int seen = 0;
            if (std::sscanf(e, "%d:%d:%d:%d", &tc, &to, &tk, &nth) >= 3 && tc == p.chunksTotal && to == p.cout && tk == p.k &&
                ++seen == nth)
                p.debugMode |= 4;
        }
        static int configuredDev[64] = {0};   // attribute is per function per device; kSmemLimit covers every k
        int dev = 0;
        OPB_CUDA(cudaGetDevice(&dev));
        if (!configuredDev[dev & 63])
        {
            OPB_CUDA(cudaFuncSetAttribute(convIgemmKernel<BN, PAIR>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      kSmemLimit));
            configuredDev[dev & 63] = 1;
        }
        // one CTA per pair of M tiles; in pair mode one 2-CTA cluster per quad
        if (p.tilesPerItem != 1)
            p.tilesPerItem = 2;
        const int tilesPerGroup = (PAIR ? 2 : 1) * p.tilesPerItem;
        const int numWork = (p.numMTiles + tilesPerGroup - 1) / tilesPerGroup * p.nTilesN;
        int grid = PAIR ? 2 * (numWork < gridCap / 2 ? numWork : gridCap / 2) : (numWork < gridCap ? numWork : gridCap);
        if (grid < (PAIR ? 2 : 1))
            grid = PAIR ? 2 : 1;
        static const bool usePdl = std::getenv("OPB200_NO_PDL") == nullptr;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid);
        cfg.blockDim = dim3(kConvThreads);
        cfg.dynamicSmemBytes = smemBytes;
        cfg.stream = st;
        cudaLaunchAttribute attr[2];
        int na = 0;
        if (PAIR)
        {
            attr[na].id = cudaLaunchAttributeClusterDimension;
            attr[na].val.clusterDim.x = 2;
            attr[na].val.clusterDim.y = 1;
            attr[na].val.clusterDim.z = 1;
            ++na;
        }
        if (usePdl)
        {
            attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[na].val.programmaticStreamSerializationAllowed = 1;
            ++na;
        }
        cfg.attrs = attr;
        cfg.numAttrs = na;
        OPB_CUDA(cudaLaunchKernelEx(&cfg, convIgemmKernel<BN, PAIR>, p));
        OPB_LAUNCH_CHECK();
        return OPB200_OK;
    }

    opb200_status launchConv(ConvParams& p, int bn, int gridCap, cudaStream_t st)
    {
        if (p.k < 1 || p.k > kConvMaxKernel || p.k % 2 == 0)
            OPB_FAIL(OPB200_ERR_UNSUPPORTED, "conv: kernel size must be odd and at most 7");
        switch (bn)
        {
        case 32: return launchConvT<32, false>(p, gridCap, st);
        case 64: return p.pairMode ? launchConvT<64, true>(p, gridCap, st) : launchConvT<64, false>(p, gridCap, st);
        case 96: return p.pairMode ? launchConvT<96, true>(p, gridCap, st) : launchConvT<96, false>(p, gridCap, st);
        case 128: return p.pairMode ? launchConvT<128, true>(p, gridCap, st) : launchConvT<128, false>(p, gridCap, st);
        default: OPB_FAIL(OPB200_ERR_INVALID, "conv: unsupported BLOCK_N");
        }
    }
}
