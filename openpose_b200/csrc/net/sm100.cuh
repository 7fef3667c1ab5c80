// sm_100a primitives used by the implicit-GEMM convolution: mbarrier, TMA (cp.async.bulk.tensor), TMEM
// allocation, tcgen05.mma / commit / ld, UMMA descriptors. Inline PTX only; compiled for sm_100a.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>

namespace opb
{
namespace sm100
{
    __device__ __forceinline__ uint32_t smemAddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

    __device__ __forceinline__ bool electOne()
    {
        uint32_t pred = 0;
        asm volatile(
            "{\n\t.reg .pred P;\n\t"
            "elect.sync _|P, 0xffffffff;\n\t"
            "selp.u32 %0, 1, 0, P;\n\t}"
            : "=r"(pred));
        return pred != 0;
    }

    // ---------------------------------------------------------------- mbarrier
    __device__ __forceinline__ void mbarInit(uint64_t* bar, uint32_t count)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smemAddr(bar)), "r"(count));
    }
    __device__ __forceinline__ void fenceBarrierInit()
    {
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __device__ __forceinline__ void mbarExpectTx(uint64_t* bar, uint32_t bytes)
    {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)), "r"(bytes)
                     : "memory");
    }
    __device__ __forceinline__ void mbarArrive(uint64_t* bar)
    {
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smemAddr(bar)) : "memory");
    }
    __device__ __forceinline__ bool mbarTryWait(uint64_t* bar, uint32_t parity)
    {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred P;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, P;\n\t}"
            : "=r"(done)
            : "r"(smemAddr(bar)), "r"(parity)
            : "memory");
        return done != 0;
    }
    __device__ __forceinline__ void mbarWait(uint64_t* bar, uint32_t parity)
    {
        while (!mbarTryWait(bar, parity)) {}
    }

    // ---------------------------------------------------------------- TMA
    __device__ __forceinline__ void prefetchTensorMap(const CUtensorMap* m)
    {
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
    }
    __device__ __forceinline__ void tmaLoad2d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1)
    {
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
            ::"r"(smemAddr(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smemAddr(bar)), "r"(c0), "r"(c1)
            : "memory");
    }
    __device__ __forceinline__ void tmaLoad3d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2)
    {
        asm volatile(
            "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
            ::"r"(smemAddr(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smemAddr(bar)), "r"(c0), "r"(c1), "r"(c2)
            : "memory");
    }
    __device__ __forceinline__ void tmaLoad4d(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2,
                                              int c3)
    {
        asm volatile(
            "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
            ::"r"(smemAddr(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smemAddr(bar)), "r"(c0), "r"(c1), "r"(c2),
            "r"(c3)
            : "memory");
    }
    // 1-D bulk copy global -> shared (no tensor map): size and both addresses multiples of 16 bytes.
    __device__ __forceinline__ void bulkLoad1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
    {
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smemAddr(dst)), "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(smemAddr(bar))
                     : "memory");
    }
    __device__ __forceinline__ void tmaStore4d(const CUtensorMap* m, const void* src, int c0, int c1, int c2, int c3)
    {
        asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
                     ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smemAddr(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
                     : "memory");
    }
    __device__ __forceinline__ void tmaStoreCommit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
    __device__ __forceinline__ void tmaStoreWaitRead0()
    {
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
    __device__ __forceinline__ void tmaStoreWaitRead1()   // all but the most recent store group have read their source
    {
        asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
    }
    __device__ __forceinline__ void tmaStoreWaitAll() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
    // generic-proxy smem writes -> visible to the async proxy (TMA store reads them)
    __device__ __forceinline__ void fenceProxyAsyncShared()
    {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }

    // ---------------------------------------------------------------- TMEM + tcgen05
    template <uint32_t kCols>
    __device__ __forceinline__ void tmemAlloc(uint32_t* dstSmem)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smemAddr(dstSmem)),
                     "n"(kCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    template <uint32_t kCols>
    __device__ __forceinline__ void tmemFree(uint32_t taddr)
    {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
    }
    __device__ __forceinline__ void tcFenceBefore() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
    __device__ __forceinline__ void tcFenceAfter() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

    // D[tmem] (+)= A[smem] * B[smem]; kind::f16 covers bf16 inputs with fp32 accumulation.
    __device__ __forceinline__ void mmaBf16(uint32_t tmemD, uint64_t descA, uint64_t descB, uint32_t idesc,
                                            uint32_t accumulate)
    {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
            ::"r"(tmemD), "l"(descA), "l"(descB), "r"(idesc), "r"(accumulate)
            : "memory");
    }
    // Arrive on an mbarrier once every previously issued tcgen05.mma of this thread has completed.
    __device__ __forceinline__ void mmaCommit(uint64_t* bar)
    {
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smemAddr(bar))
                     : "memory");
    }
    // 32 lanes x 32 consecutive fp32 columns -> 32 registers per thread (thread = lane = accumulator row).
    __device__ __forceinline__ void tmemLoad32(uint32_t taddr, uint32_t (&v)[32])
    {
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
            "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
            "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
            : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
              "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
              "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
              "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
            : "r"(taddr));
    }
    __device__ __forceinline__ void tmemLoadWait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

    // UMMA shared-memory descriptor, K-major operand, 128-byte swizzle, rows of 64 bf16 (=128 B), 8-row
    // groups 1024 B apart (SBO). Field layout: cute/arch/mma_sm100_desc.hpp SmemDescriptor (version 1).
    __device__ __forceinline__ uint64_t smemDescSw128(uint32_t saddr, uint32_t strideBytes = 1024)
    {
        uint64_t d = 0;
        d |= (uint64_t)((saddr & 0x3FFFF) >> 4);          // start address, bits [0,14)
        d |= (uint64_t)0 << 16;                           // leading byte offset: unused for swizzled K-major
        d |= (uint64_t)(strideBytes >> 4) << 32;          // stride byte offset (8-row group pitch), bits [32,46)
        d |= (uint64_t)1 << 46;                           // descriptor version (sm_100)
        d |= (uint64_t)2 << 61;                           // layout type SWIZZLE_128B
        return d;
    }
    // Instruction descriptor: D=f32, A=B=bf16, both K-major, M x N tile (InstrDescriptor, same header).
    __host__ __device__ constexpr uint32_t instrDescBf16(uint32_t m, uint32_t n)
    {
        return (1u << 4) | (1u << 7) | (1u << 10) | ((n >> 3) << 17) | ((m >> 4) << 24);
    }

    // ---------------------------------------------------------------- CTA pair (cluster of 2, cta_group::2)
    __device__ __forceinline__ uint32_t clusterCtaRank()
    {
        uint32_t r;
        asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
        return r;
    }
    __device__ __forceinline__ void clusterSync()
    {
        asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    // shared::cta address of this CTA -> shared::cluster address of the same offset in CTA `rank` of the cluster
    __device__ __forceinline__ uint32_t mapaShared(uint32_t addr, uint32_t rank)
    {
        uint32_t r;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
        return r;
    }
    __device__ __forceinline__ void mbarArriveCluster(uint32_t clusterAddr)
    {
        asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(clusterAddr) : "memory");
    }
    // TMA loads of a CTA pair: data lands in this CTA's shared memory, the bytes are counted on the mbarrier at
    // shared::cluster address `barCluster` (the leader CTA's barrier).
    __device__ __forceinline__ void tmaLoad2dPair(void* dst, const CUtensorMap* m, uint32_t barCluster, int c0, int c1)
    {
        asm volatile(
            "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
            ::"r"(smemAddr(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(barCluster), "r"(c0), "r"(c1)
            : "memory");
    }
    __device__ __forceinline__ void tmaLoad3dPair(void* dst, const CUtensorMap* m, uint32_t barCluster, int c0, int c1,
                                                  int c2)
    {
        asm volatile(
            "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
            ::"r"(smemAddr(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(barCluster), "r"(c0), "r"(c1), "r"(c2)
            : "memory");
    }
    __device__ __forceinline__ void tmaLoad4dPair(void* dst, const CUtensorMap* m, uint32_t barCluster, int c0, int c1,
                                                  int c2, int c3)
    {
        asm volatile(
            "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
            ::"r"(smemAddr(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(barCluster), "r"(c0), "r"(c1), "r"(c2),
            "r"(c3)
            : "memory");
    }
    // collective over the pair: the same warp of both CTAs executes it; both get the same column range
    template <uint32_t kCols>
    __device__ __forceinline__ void tmemAllocPair(uint32_t* dstSmem)
    {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smemAddr(dstSmem)),
                     "n"(kCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    template <uint32_t kCols>
    __device__ __forceinline__ void tmemFreePair(uint32_t taddr)
    {
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(kCols) : "memory");
    }
    // M = 256 over the pair: rows 0..127 from this (leader) CTA's A tile, 128..255 from the peer's tile at the same
    // shared-memory offset; B = N/2 rows from each CTA; each CTA's TMEM receives its own 128 rows. Leader CTA only.
    __device__ __forceinline__ void mmaBf16Pair(uint32_t tmemD, uint64_t descA, uint64_t descB, uint32_t idesc,
                                                uint32_t accumulate)
    {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "setp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
            ::"r"(tmemD), "l"(descA), "l"(descB), "r"(idesc), "r"(accumulate)
            : "memory");
    }
    // arrive on the barrier at this shared-memory offset in BOTH CTAs once the pair's MMAs have retired
    __device__ __forceinline__ void mmaCommitPair(uint64_t* bar)
    {
        asm volatile(
            "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
            ::"r"(smemAddr(bar)), "h"((uint16_t)3)
            : "memory");
    }

    // ---------------------------------------------------------------- grouped MMA issue
    // All MMAs of one weight box (G filter rows x 4 k-steps x 1..2 M tiles) in ONE asm statement. Written in C++ the issue
    // loop recomputes every descriptor in vector registers and moves it to the uniform datapath with R2UR before each
    // UTCHMMA (~30 R2UR and ~100 instructions per filter row): ~600 cycles per filter row whatever the MMA shape, which
    // paced every layer whose 8 MMAs per row need less tensor time than that (N <= 96, single-tile items). Here the
    // operands cross to uniform registers once per box and the per-MMA descriptors are uniform adds.
    //   d0, d1: TMEM accumulators of the two M tiles (d1 only used when two != 0)
    //   a0, a1: smem descriptors of the two tiles at the box's first filter row; aRowStep = one filter row down (16-byte units)
    //   b: smem descriptor of the box's first K block; bRowStep = one K block further (16-byte units)
    //   accFirst: 0 = the very first MMA of the work item overwrites the accumulator; steps: 16-channel k-steps (1..4)
    // Only the low word of a descriptor changes between MMAs (start address, bits [0,14)); the adds are 32-bit and the
    // 64-bit operand is the register pair {low, constant high word}.
#define OPB_MMA_ROW(CG, ACC0)                                                                        \
    "mov.b64 da, {ra0, ah};\n\tmov.b64 db, {rb, bh};\n\tmov.b64 dc, {ra1, ah};\n\t"                  \
    "tcgen05.mma.cta_group::" CG ".kind::f16 [%0], da, db, %7, " ACC0 ";\n\t"                        \
    "@pt tcgen05.mma.cta_group::" CG ".kind::f16 [%1], dc, db, %7, " ACC0 ";\n\t"                    \
    "add.u32 ta, ra0, 2;\n\tadd.u32 tb, rb, 2;\n\tadd.u32 tc, ra1, 2;\n\t"                         \
    "mov.b64 da, {ta, ah};\n\tmov.b64 db, {tb, bh};\n\tmov.b64 dc, {tc, ah};\n\t"                    \
    "@p1 tcgen05.mma.cta_group::" CG ".kind::f16 [%0], da, db, %7, pT;\n\t"                          \
    "@p1t tcgen05.mma.cta_group::" CG ".kind::f16 [%1], dc, db, %7, pT;\n\t"                         \
    "add.u32 ta, ra0, 4;\n\tadd.u32 tb, rb, 4;\n\tadd.u32 tc, ra1, 4;\n\t"                         \
    "mov.b64 da, {ta, ah};\n\tmov.b64 db, {tb, bh};\n\tmov.b64 dc, {tc, ah};\n\t"                    \
    "@p2 tcgen05.mma.cta_group::" CG ".kind::f16 [%0], da, db, %7, pT;\n\t"                          \
    "@p2t tcgen05.mma.cta_group::" CG ".kind::f16 [%1], dc, db, %7, pT;\n\t"                         \
    "add.u32 ta, ra0, 6;\n\tadd.u32 tb, rb, 6;\n\tadd.u32 tc, ra1, 6;\n\t"                         \
    "mov.b64 da, {ta, ah};\n\tmov.b64 db, {tb, bh};\n\tmov.b64 dc, {tc, ah};\n\t"                    \
    "@p3 tcgen05.mma.cta_group::" CG ".kind::f16 [%0], da, db, %7, pT;\n\t"                          \
    "@p3t tcgen05.mma.cta_group::" CG ".kind::f16 [%1], dc, db, %7, pT;\n\t"
#define OPB_MMA_NEXT_ROW "add.u32 ra0, ra0, %5;\n\tadd.u32 ra1, ra1, %5;\n\tadd.u32 rb, rb, %6;\n\t"
#define OPB_MMA_PROLOGUE                                                                             \
    "{\n\t.reg .pred pf, pT, pt, p1, p2, p3, p1t, p2t, p3t;\n\t"                                     \
    ".reg .b32 ra0, ra1, rb, ah, bh, ta, tb, tc;\n\t.reg .b64 da, db, dc;\n\t"                       \
    "setp.ne.u32 pf, %8, 0;\n\tsetp.eq.u32 pT, 0, 0;\n\tsetp.ne.u32 pt, %10, 0;\n\t"                 \
    "setp.gt.u32 p1, %9, 1;\n\tsetp.gt.u32 p2, %9, 2;\n\tsetp.gt.u32 p3, %9, 3;\n\t"                 \
    "and.pred p1t, p1, pt;\n\tand.pred p2t, p2, pt;\n\tand.pred p3t, p3, pt;\n\t"                    \
    "mov.b64 {ra0, ah}, %2;\n\tmov.b64 {ra1, tc}, %3;\n\tmov.b64 {rb, bh}, %4;\n\t"
#define OPB_MMA_OPERANDS                                                                             \
    ::"r"(d0), "r"(d1), "l"(a0), "l"(a1), "l"(b), "r"(aRowStep), "r"(bRowStep), "r"(idesc), "r"(accFirst), "r"(steps), \
      "r"(two)                                                                                       \
    : "memory"
    template <int G, bool PAIR>
    __device__ __forceinline__ void mmaGroupBf16(uint32_t d0, uint32_t d1, uint64_t a0, uint64_t a1, uint64_t b,
                                                 uint32_t aRowStep, uint32_t bRowStep, uint32_t idesc, uint32_t accFirst,
                                                 uint32_t steps, uint32_t two)
    {
        static_assert(G == 1 || G == 3, "a weight box holds one filter row or the three rows of a 3x3 filter column");
        if (PAIR)
        {
            if (G == 1)
                asm volatile(OPB_MMA_PROLOGUE OPB_MMA_ROW("2", "pf") "}" OPB_MMA_OPERANDS);
            else
                asm volatile(OPB_MMA_PROLOGUE OPB_MMA_ROW("2", "pf") OPB_MMA_NEXT_ROW OPB_MMA_ROW("2", "pT") OPB_MMA_NEXT_ROW
                                 OPB_MMA_ROW("2", "pT") "}" OPB_MMA_OPERANDS);
        }
        else
        {
            if (G == 1)
                asm volatile(OPB_MMA_PROLOGUE OPB_MMA_ROW("1", "pf") "}" OPB_MMA_OPERANDS);
            else
                asm volatile(OPB_MMA_PROLOGUE OPB_MMA_ROW("1", "pf") OPB_MMA_NEXT_ROW OPB_MMA_ROW("1", "pT") OPB_MMA_NEXT_ROW
                                 OPB_MMA_ROW("1", "pT") "}" OPB_MMA_OPERANDS);
        }
    }
#undef OPB_MMA_ROW
#undef OPB_MMA_NEXT_ROW
#undef OPB_MMA_PROLOGUE
#undef OPB_MMA_OPERANDS

    // ---------------------------------------------------------------- epilogue helpers
    __device__ __forceinline__ float4 ldsF4(uint32_t saddr)
    {
        float4 v;
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr));
        return v;
    }
    __device__ __forceinline__ void stsU4(uint32_t saddr, uint4 u)
    {
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(u.x), "r"(u.y), "r"(u.z), "r"(u.w)
                     : "memory");
    }
    // (lo, hi) fp32 -> packed bf16x2, round to nearest even; lo in the low half
    __device__ __forceinline__ uint32_t packBf16(float lo, float hi)
    {
        uint32_t r;
        asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
        return r;
    }
    // the same with max(x, 0) folded into the conversion (ReLU layers)
    __device__ __forceinline__ uint32_t packBf16Relu(float lo, float hi)
    {
        uint32_t r;
        asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
        return r;
    }
    __device__ __forceinline__ uint32_t maxBf16x2(uint32_t a, uint32_t b)
    {
        uint32_t r;
        asm("max.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
        return r;
    }

    __device__ __forceinline__ void namedBarrier(uint32_t id, uint32_t threads)
    {
        asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
    }
}
}
