// Research tool (GPU): what bounds the L2 -> shared-memory operand stream of the conv kernel?
// Persistent CTAs (one per SM) issue TMA box loads from an NHWC bf16 activation tensor (or a weight matrix) through a
// ring of D shared-memory stages and nothing else; reports bytes/clk/SM and TB/s for box shape x depth x channel pitch.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o openpose_b200/lib/tma_probe tools/tma_probe.cu
#include "../openpose_b200/csrc/net/sm100.cuh"
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>
#include <vector>

using namespace opb::sm100;

struct ProbeArgs
{
    CUtensorMap map;
    int dims;          // 4 = activation boxes, 2 = weight boxes
    int depth;         // stages in flight
    int boxBytes;      // bytes one box delivers
    int stageBytes;    // smem per stage (1024-aligned)
    int boxesPerCta;
    int tilesX, tilesY, N, tileW, tileH, cBlocks;
    int rowsTotal, boxRows;   // 2-D
};

__global__ void __launch_bounds__(128, 1) probeKernel(const __grid_constant__ ProbeArgs a, unsigned long long* clocks)
{
    extern __shared__ uint8_t smemRaw[];
    uint8_t* smem = smemRaw + ((1024u - (smemAddr(smemRaw) & 1023u)) & 1023u);
    __shared__ uint64_t full[16];
    __shared__ int4 coord[512];                      // per-box coordinates, computed before the timed loop
    if (threadIdx.x == 0)
    {
        for (int i = 0; i < a.depth; ++i)
            mbarInit(&full[i], 1);
        fenceBarrierInit();
        prefetchTensorMap(&a.map);
    }
    const int perFrame = a.tilesX * a.tilesY;
    for (int b = threadIdx.x; b < a.boxesPerCta; b += blockDim.x)
    {
        // consecutive boxes of a CTA are the channel blocks of one tile, tiles are spread over CTAs
        if (a.dims == 4)
        {
            const int tile = (b / a.cBlocks) * gridDim.x + blockIdx.x;
            const int n = (tile / perFrame) % a.N, r = tile % perFrame;
            coord[b] = make_int4((b % a.cBlocks) * 64, (r % a.tilesX) * a.tileW - 1, (r / a.tilesX) * a.tileH - 1, n);
        }
        else
        {
            const int idx = b * gridDim.x + blockIdx.x;
            coord[b] = make_int4((idx % a.cBlocks) * 64, ((idx / a.cBlocks) * a.boxRows) % a.rowsTotal, 0, 0);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        const unsigned long long t0 = clock64();
        int s = 0;
        uint32_t parity = 1;                          // flips to 0 when the ring wraps the first time
        for (int b = 0; b < a.boxesPerCta + a.depth; ++b)
        {
            if (b >= a.depth)
                mbarWait(&full[s], parity);           // the box issued `depth` iterations ago has landed
            if (b < a.boxesPerCta)
            {
                const int4 c = coord[b];
                mbarExpectTx(&full[s], a.boxBytes);
                if (a.dims == 4)
                    tmaLoad4d(smem + s * a.stageBytes, &a.map, &full[s], c.x, c.y, c.z, c.w);
                else if (a.dims == 2)
                    tmaLoad2d(smem + s * a.stageBytes, &a.map, &full[s], c.x, c.y);
                else
                    asm volatile(
                        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                        ::"r"(smemAddr(smem + s * a.stageBytes)), "l"(reinterpret_cast<uint64_t>(&a.map)),
                        "r"(smemAddr(&full[s])), "r"(0), "r"(c.y), "r"(c.x / 64 * 3)
                        : "memory");
            }
            if (++s == a.depth)
            {
                s = 0;
                parity ^= 1;
            }
        }
        clocks[blockIdx.x] = clock64() - t0;
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

#define CK(x)                                                                                   \
    do                                                                                          \
    {                                                                                           \
        cudaError_t e = (x);                                                                    \
        if (e != cudaSuccess)                                                                   \
        {                                                                                       \
            printf("CUDA error %s at line %d\n", cudaGetErrorString(e), __LINE__);              \
            exit(1);                                                                            \
        }                                                                                       \
    } while (0)

int main()
{
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
    EncodeTiledFn encode = reinterpret_cast<EncodeTiledFn>(fp);
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    const double ghz = prop.clockRate * 1e-6;
    const int N = 44, H = 46, W = 82;
    const size_t maxBytes = (size_t)N * H * W * 768 * 2;
    void* buf = nullptr;
    CK(cudaMalloc(&buf, maxBytes));
    CK(cudaMemset(buf, 1, maxBytes));
    unsigned long long* dClk = nullptr;
    CK(cudaMalloc(&dClk, sizeof(unsigned long long) * sms));
    CK(cudaFuncSetAttribute(probeKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    printf("# %s, %d SMs, nominal %.2f GHz. activation tensor %dx%dx%d, bf16\n", prop.name, sms, ghz, N, H, W);
    printf("%-34s %5s %6s %9s %10s %10s %9s\n", "box", "pitch", "depth", "boxes/CTA", "us", "B/clk/SM", "TB/s");
    struct Shape
    {
        const char* name;
        int tw, th, halo;
    };
    const Shape shapes[] = {{"act 10x18 halo (8x16 tile)", 8, 16, 2}, {"act 34x6 halo (32x4 tile)", 32, 4, 2}};
    for (const Shape& sh : shapes)
        for (int pitchC : {128, 384, 288})
            for (int offC : {0, 96})
            {
                if (offC && pitchC == 128)
                    continue;
                for (int depth : {1, 2, 4, 8})
                {
                    ProbeArgs a = {};
                    const int C = pitchC == 384 ? 128 : (pitchC == 288 ? 96 : pitchC);   // the view's own channels (96: half of the 2nd chunk is out of bounds)
                    const cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
                    const cuuint64_t strides[3] = {(cuuint64_t)pitchC * 2, (cuuint64_t)pitchC * 2 * W,
                                                   (cuuint64_t)pitchC * 2 * W * H};
                    const cuuint32_t box[4] = {64, (cuuint32_t)(sh.tw + sh.halo), (cuuint32_t)(sh.th + sh.halo), 1};
                    const cuuint32_t es[4] = {1, 1, 1, 1};
                    CUresult r = encode(&a.map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4,
                                        static_cast<uint8_t*>(buf) + offC * 2, dims, strides, box, es,
                                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                        CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                    if (r != CUDA_SUCCESS)
                    {
                        printf("encode failed %d\n", (int)r);
                        continue;
                    }
                    a.dims = 4;
                    a.depth = depth;
                    a.boxBytes = box[1] * box[2] * 128;
                    a.stageBytes = (a.boxBytes + 1023) & ~1023;
                    a.tileW = sh.tw;
                    a.tileH = sh.th;
                    a.tilesX = (W + sh.tw - 1) / sh.tw;
                    a.tilesY = (H + sh.th - 1) / sh.th;
                    a.N = N;
                    a.cBlocks = (C + 63) / 64;
                    a.boxesPerCta = 200;
                    const int smemBytes = depth * a.stageBytes + 1024;
                    float best = 1e30f;
                    for (int rep = 0; rep < 4; ++rep)
                    {
                        cudaEventRecord(e0);
                        probeKernel<<<sms, 128, smemBytes>>>(a, dClk);
                        cudaEventRecord(e1);
                        CK(cudaEventSynchronize(e1));
                        float ms;
                        cudaEventElapsedTime(&ms, e0, e1);
                        if (rep && ms < best)
                            best = ms;
                    }
                    std::vector<unsigned long long> clk(sms);
                    CK(cudaMemcpy(clk.data(), dClk, sizeof(unsigned long long) * sms, cudaMemcpyDeviceToHost));
                    double avg = 0;
                    for (auto c : clk)
                        avg += (double)c / sms;
                    const double bytes = (double)a.boxesPerCta * a.boxBytes;
                    char name[96];
                    snprintf(name, sizeof(name), "%s off%d", sh.name, offC);
                    printf("%-34s %5d %6d %9d %10.1f %10.2f %9.2f\n", name, pitchC, depth, a.boxesPerCta, best * 1e3,
                           bytes / avg, bytes * sms / (best * 1e-3) * 1e-12);
                }
            }
    // weight-like 2-D boxes: 64 K-elements x rows, from a (rows x K) matrix with K = 1152
    for (int boxRows : {64, 128})
        for (int depth : {1, 2, 4, 8, 12})
        {
            ProbeArgs a = {};
            const int K = 1152, rows = 128;
            const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
            const cuuint64_t strides[1] = {(cuuint64_t)K * 2};
            const cuuint32_t box[2] = {64, (cuuint32_t)boxRows};
            const cuuint32_t es[2] = {1, 1};
            CUresult r = encode(&a.map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, dims, strides, box, es,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS)
            {
                printf("encode failed %d\n", (int)r);
                continue;
            }
            a.dims = 2;
            a.depth = depth;
            a.boxBytes = boxRows * 128;
            a.stageBytes = a.boxBytes;
            a.cBlocks = K / 64;
            a.rowsTotal = rows;
            a.boxRows = boxRows;
            a.boxesPerCta = 400;
            const int smemBytes = depth * a.stageBytes + 1024;
            float best = 1e30f;
            for (int rep = 0; rep < 4; ++rep)
            {
                cudaEventRecord(e0);
                probeKernel<<<sms, 128, smemBytes>>>(a, dClk);
                cudaEventRecord(e1);
                CK(cudaEventSynchronize(e1));
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                if (rep && ms < best)
                    best = ms;
            }
            std::vector<unsigned long long> clk(sms);
            CK(cudaMemcpy(clk.data(), dClk, sizeof(unsigned long long) * sms, cudaMemcpyDeviceToHost));
            double avg = 0;
            for (auto c : clk)
                avg += (double)c / sms;
            const double bytes = (double)a.boxesPerCta * a.boxBytes;
            char name[96];
            snprintf(name, sizeof(name), "weights 64 x %d rows (K=1152)", boxRows);
            printf("%-34s %5d %6d %9d %10.1f %10.2f %9.2f\n", name, K, depth, a.boxesPerCta, best * 1e3, bytes / avg,
                   bytes * sms / (best * 1e-3) * 1e-12);
        }
    // the same weights as 3-D boxes (64 K-elements, rows, 3 consecutive K blocks = the 3 filter rows of one column)
    for (int boxRows : {64, 128})
        for (int depth : {1, 2, 4})
        {
            ProbeArgs a = {};
            const int K = 1152, rows = 128;
            const cuuint64_t dims[3] = {64, (cuuint64_t)rows, (cuuint64_t)(K / 64)};
            const cuuint64_t strides[2] = {(cuuint64_t)K * 2, 128};
            const cuuint32_t box[3] = {64, (cuuint32_t)boxRows, 3};
            const cuuint32_t es[3] = {1, 1, 1};
            CUresult r = encode(&a.map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, buf, dims, strides, box, es,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS)
            {
                printf("encode 3-D failed %d\n", (int)r);
                continue;
            }
            a.dims = 3;
            a.depth = depth;
            a.boxBytes = boxRows * 128 * 3;
            a.stageBytes = a.boxBytes;
            a.cBlocks = K / 64 / 3;      // coord.x = 64 * 3 * (idx % 6) -> K block = 3 * (idx % 6)
            a.rowsTotal = rows;
            a.boxRows = boxRows;
            a.boxesPerCta = 200;
            const int smemBytes = depth * a.stageBytes + 1024;
            float best = 1e30f;
            for (int rep = 0; rep < 4; ++rep)
            {
                cudaEventRecord(e0);
                probeKernel<<<sms, 128, smemBytes>>>(a, dClk);
                cudaEventRecord(e1);
                CK(cudaEventSynchronize(e1));
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                if (rep && ms < best)
                    best = ms;
            }
            std::vector<unsigned long long> clk(sms);
            CK(cudaMemcpy(clk.data(), dClk, sizeof(unsigned long long) * sms, cudaMemcpyDeviceToHost));
            double avg = 0;
            for (auto c : clk)
                avg += (double)c / sms;
            const double bytes = (double)a.boxesPerCta * a.boxBytes;
            char name[96];
            snprintf(name, sizeof(name), "weights 64 x %d rows x 3 K blocks", boxRows);
            printf("%-34s %5d %6d %9d %10.1f %10.2f %9.2f\n", name, K, depth, a.boxesPerCta, best * 1e3, bytes / avg,
                   bytes * sms / (best * 1e-3) * 1e-12);
        }
    return 0;
}
