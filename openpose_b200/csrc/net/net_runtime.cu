// NetB200: the Caffe-free op::Net backend. Loads prototxt + caffemodel, plans the graph once (conv + bias +
// activation fused, concat by channel offset / multi-segment reads, weights repacked to bf16 K-major), and runs
// the forward pass as a stream of hand-written sm_100a kernels (conv_sm100.cu, aux_kernels.cu).
//
// Replaces op::NetCaffe (reference: src/openpose/net/netCaffe.cpp:45-99 ctor checks, :154-210 initialization,
// :212-261 forwardPass, :263-278 getOutputBlobArray) and the Caffe inference subset underneath it
// (3rdparty/caffe/src/caffe/net.cpp:518-535 ForwardFromTo; conv/relu/prelu/pooling/concat layers).
// Exposed through the C ABI declared in include/opb200.h (opb200_net_*).
#include "caffe_format.h"
#include "net_kernels.h"
#include <cuda_bf16.h>
#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <memory>
#include <set>
#include <string>
#include <vector>

namespace opb
{
    // ------------------------------------------------------------------------------------------------
    // tensor maps
    // ------------------------------------------------------------------------------------------------
    typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                      const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                      CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

    static EncodeTiledFn encodeTiled()
    {
        static EncodeTiledFn fn = nullptr;
        if (!fn)
        {
            void* p = nullptr;
            cudaDriverEntryPointQueryResult q;
            if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
                q == cudaDriverEntryPointSuccess)
                fn = reinterpret_cast<EncodeTiledFn>(p);
        }
        return fn;
    }

    opb200_status makeActivationMap(CUtensorMap* m, const void* base, int C, int W, int H, int N, int strideC,
                                    int boxC, int TW, int TH, bool swizzle128)
    {
        EncodeTiledFn fn = encodeTiled();
        if (!fn)
            OPB_FAIL(OPB200_ERR_CUDA, "cuTensorMapEncodeTiled is unavailable (driver too old?)");
        const cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
        const cuuint64_t strides[3] = {(cuuint64_t)strideC * 2, (cuuint64_t)strideC * 2 * W,
                                       (cuuint64_t)strideC * 2 * W * H};
        const cuuint32_t box[4] = {(cuuint32_t)boxC, (cuuint32_t)TW, (cuuint32_t)TH, 1};
        const cuuint32_t es[4] = {1, 1, 1, 1};
        const CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, es,
                              CU_TENSOR_MAP_INTERLEAVE_NONE,
                              swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS)
        {
            setError("cuTensorMapEncodeTiled(activation C=%d W=%d H=%d N=%d stride=%d box=%d,%d,%d) failed: %d", C, W,
                     H, N, strideC, boxC, TW, TH, (int)r);
            return OPB200_ERR_CUDA;
        }
        return OPB200_OK;
    }

    // Weights [rows][K] seen as (64 K-elements, rows, K / 64 blocks): one box = boxRows x 64 elements of boxBlocks
    // consecutive K blocks (the filter rows ky of one filter column), landing as boxBlocks swizzled (boxRows x 128 B)
    // tiles one after the other. TMA time per box has a floor of ~0.16 us whatever its size (tools/tma_probe.cu), so
    // one 3-block box instead of three 1-block boxes triples the weight stream's ceiling.
    opb200_status makeWeightMap(CUtensorMap* m, const void* base, int K, int rows, int boxRows, int boxBlocks)
    {
        EncodeTiledFn fn = encodeTiled();
        if (!fn)
            OPB_FAIL(OPB200_ERR_CUDA, "cuTensorMapEncodeTiled is unavailable (driver too old?)");
        const cuuint64_t dims[3] = {64, (cuuint64_t)rows, (cuuint64_t)(K / 64)};
        const cuuint64_t strides[2] = {(cuuint64_t)K * 2, 128};
        const cuuint32_t box[3] = {64, (cuuint32_t)boxRows, (cuuint32_t)boxBlocks};
        const cuuint32_t es[3] = {1, 1, 1};
        const CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, es,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS)
        {
            setError("cuTensorMapEncodeTiled(weights K=%d rows=%d box=%dx%d) failed: %d", K, rows, boxRows, boxBlocks,
                     (int)r);
            return OPB200_ERR_CUDA;
        }
        return OPB200_OK;
    }

    opb200_status makeFlatStoreMap(CUtensorMap* m, const void* base, int C, int W, int H, int N, int pitchW, int strideC,
                                   int boxC, int TW, int TH)
    {
        EncodeTiledFn fn = encodeTiled();
        if (!fn)
            OPB_FAIL(OPB200_ERR_CUDA, "cuTensorMapEncodeTiled is unavailable (driver too old?)");
        const cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)N, (cuuint64_t)H};
        const cuuint64_t strides[3] = {(cuuint64_t)strideC * 2, (cuuint64_t)strideC * 2 * pitchW,
                                       (cuuint64_t)strideC * 2 * pitchW * N};
        const cuuint32_t box[4] = {(cuuint32_t)boxC, (cuuint32_t)TW, 1, (cuuint32_t)TH};
        const cuuint32_t es[4] = {1, 1, 1, 1};
        const CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, es,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS)
        {
            setError("cuTensorMapEncodeTiled(flat store C=%d W=%d H=%d N=%d pitch=%d stride=%d) failed: %d", C, W, H, N,
                     pitchW, strideC, (int)r);
            return OPB200_ERR_CUDA;
        }
        return OPB200_OK;
    }

    // ------------------------------------------------------------------------------------------------
    // plan
    // ------------------------------------------------------------------------------------------------
    struct View   // a channel range of a device buffer (bf16 NHWC)
    {
        int buf = -1, cOff = 0, C = 0;
    };
    struct Blob
    {
        int C = 0, level = 0;
        bool isInput = false;
        std::vector<View> segs;   // >1: virtual concat (read as several TMA segments)
    };
    struct Buffer
    {
        int C = 0, level = 0;     // allocated channels (multiple of 8)
        size_t offset = 0;        // into the arena, set by reshape
    };
    enum OpKind
    {
        OP_PATCH,
        OP_CONV,
        OP_POOL,
        OP_CONVERT,
        OP_FIRSTCONV   // first layer straight from the fp32 NCHW input (mma.sync kernel in aux_kernels.cu)
    };
    struct Op
    {
        OpKind kind = OP_CONV;
        std::string name;
        int level = 0, outLevel = 0;
        std::vector<View> in;
        View out;
        // conv
        int cin = 0, cout = 0, k = 1, pad = 0, bn = 128, nTilesN = 1, chunksTotal = 0, kTotal = 0;
        int bnEff = 128;                    // Cout tile of the current shape: bn, or a divisor of it at small batches (reshapeNet)
        int segChunks[kConvMaxSegments] = {0, 0, 0, 0};
        std::vector<__nv_bfloat16> hostW;   // [nTilesN*bn][kTotal]
        std::vector<float> hostBias, hostSlope;
        std::vector<uint32_t> hostFrag;     // OP_FIRSTCONV: mma.sync B fragments [n-tile][k-step][lane][2]
        __nv_bfloat16* dW = nullptr;
        float *dBias = nullptr, *dSlope = nullptr;
        int out32Off = -1;
        bool storeBf16 = true;
        bool poolFused = false;     // a 2x2/2 MAX pooling that only this conv feeds runs in its epilogue
        View poolOut;               //   ... writing this view at level poolLevel
        int poolLevel = 0;
        double flopsPerPixel = 0;   // algorithmic: 2 * cout * cin * k * k
        // pool
        int stride = 1;
        // patch (first layer im2col): k, pad reuse the fields above; inC = network input channels
        int inC = 0;
        ConvParams cp;
    };

    struct Level
    {
        int parent = -1, k = 1, stride = 1, pad = 0;   // level 0 = network input resolution
    };

    struct Net
    {
        int gpu = 0;
        std::string lastBlobName;
        NetDef def;
        std::map<std::string, int> blobIndex;
        std::vector<Blob> blobs;
        std::vector<Buffer> buffers;
        std::vector<Level> levels;
        std::vector<Op> ops;
        int inputC = 3;
        int outBlob = -1, outC = 0, outLevel = 0;
        // shape-dependent state
        int N = 0, H = 0, W = 0;
        std::vector<int> lvH, lvW;
        std::vector<int> lvFlat;   // per level: 1 = frame-interleaved layout [y][n][x of W + sep][c] (conv_params.h)
        std::vector<int> lvSep;    // zero separator columns after every frame of a flat level = the widest horizontal padding of its convs
        uint8_t* arena = nullptr;
        size_t arenaBytes = 0;
        float* dInput = nullptr;
        float* dOut32 = nullptr;
        size_t inputBytes = 0, outBytes = 0;
        cudaGraphExec_t graphExec = nullptr;
        const float* graphInput = nullptr;
        cudaStream_t graphStream = nullptr;
        bool useGraph = true;
        double flopsPerForward = 0;
        int gridCap = kNumSMsB200;

        ~Net()
        {
            if (graphExec)
                cudaGraphExecDestroy(graphExec);
            if (arena)
                cudaFree(arena);
            for (auto& o : ops)
            {
                if (o.dW)
                    cudaFree(o.dW);
                if (o.dBias)
                    cudaFree(o.dBias);
                if (o.dSlope)
                    cudaFree(o.dSlope);
            }
        }
    };

    static int round8(int c) { return (c + 7) / 8 * 8; }

    static int newBuffer(Net& net, int C, int level)
    {
        Buffer b;
        b.C = round8(C);
        b.level = level;
        net.buffers.push_back(b);
        return (int)net.buffers.size() - 1;
    }

    static bool overlaps(const View& a, const View& b)
    {
        return a.buf == b.buf && a.cOff < b.cOff + b.C && b.cOff < a.cOff + a.C;
    }

    static opb200_status planFail(const std::string& msg)
    {
        setError("%s", msg.c_str());
        return OPB200_ERR_UNSUPPORTED;
    }

    static opb200_status planNet(Net& net, const WeightMap& wm)
    {
        const NetDef& def = net.def;
        // ---- input blob ---------------------------------------------------------------------------
        std::string inputName;
        if (!def.inputs.empty())
        {
            inputName = def.inputs[0];
            if (def.inputDim.size() >= 2)
                net.inputC = def.inputDim[1];
        }
        for (const auto& L : def.layers)
            if (L.type == "Input" && !L.tops.empty() && inputName.empty())
            {
                inputName = L.tops[0];
                if (L.inputShape.size() >= 2)
                    net.inputC = L.inputShape[1];
            }
        if (inputName.empty())
            return planFail("prototxt: the net declares no input blob");
        net.levels.push_back(Level());
        {
            Blob b;
            b.C = net.inputC;
            b.isInput = true;
            net.blobIndex[inputName] = 0;
            net.blobs.push_back(b);
        }
        // ---- pass 1: channel counts of every top, producer type ----------------------------------
        std::map<std::string, int> chan;
        std::map<std::string, std::string> producerType;
        chan[inputName] = net.inputC;
        producerType[inputName] = "Input";
        for (const auto& L : def.layers)
        {
            if (L.type == "Input" || L.tops.empty())
                continue;
            int c = 0;
            if (L.type == "Convolution")
                c = L.numOutput;
            else if (L.type == "Concat")
            {
                for (const auto& b : L.bottoms)
                {
                    if (!chan.count(b))
                        return planFail("prototxt: layer " + L.name + " reads unknown blob " + b);
                    c += chan[b];
                }
            }
            else
            {
                if (L.bottoms.empty() || !chan.count(L.bottoms[0]))
                    return planFail("prototxt: layer " + L.name + " reads an unknown blob");
                c = chan[L.bottoms[0]];
            }
            chan[L.tops[0]] = c;
            if (!(L.bottoms.size() == 1 && L.bottoms[0] == L.tops[0]))   // in-place layers keep the producer
                producerType[L.tops[0]] = L.type;
        }
        // ---- pass 2: co-location claims: a concat whose members are all fresh conv/pool outputs gets one
        //      buffer and its members are written straight into their channel slots ----------------------
        struct Claim
        {
            std::string concat;
            int off;
        };
        std::map<std::string, Claim> claims;
        std::set<std::string> owningConcats;
        for (const auto& L : def.layers)
        {
            if (L.type != "Concat")
                continue;
            bool ok = true;
            int off = 0;
            std::set<std::string> seen;
            for (const auto& b : L.bottoms)
            {
                const std::string& pt = producerType[b];
                if (claims.count(b) || seen.count(b) || (pt != "Convolution" && pt != "Pooling") || off % 8 != 0)
                    ok = false;
                seen.insert(b);
                off += chan[b];
            }
            if (!ok)
                continue;
            off = 0;
            for (const auto& b : L.bottoms)
            {
                claims[b] = Claim{L.tops[0], off};
                off += chan[b];
            }
            owningConcats.insert(L.tops[0]);
        }
        std::map<std::string, int> concatBuffer;   // owning concat top -> buffer id (created on first use)
        auto outputView = [&](const std::string& top, int C, int level) -> View {
            View v;
            v.C = C;
            auto it = claims.find(top);
            if (it == claims.end())
            {
                v.buf = newBuffer(net, C, level);
                v.cOff = 0;
                return v;
            }
            auto cb = concatBuffer.find(it->second.concat);
            if (cb == concatBuffer.end())
            {
                const int id = newBuffer(net, chan[it->second.concat], level);
                cb = concatBuffer.insert({it->second.concat, id}).first;
            }
            v.buf = cb->second;
            v.cOff = it->second.off;
            return v;
        };

        // ---- pass 3: ops ----------------------------------------------------------------------------
        std::vector<bool> consumed(def.layers.size(), false);
        for (size_t li = 0; li < def.layers.size(); ++li)
        {
            const LayerDef& L = def.layers[li];
            if (consumed[li] || L.type == "Input")
                continue;
            if (L.tops.size() != 1)
                return planFail("layer " + L.name + ": exactly one top is supported");
            std::vector<const Blob*> ins;
            for (const auto& b : L.bottoms)
            {
                auto it = net.blobIndex.find(b);
                if (it == net.blobIndex.end())
                    return planFail("layer " + L.name + " reads blob " + b + " before it is produced");
                ins.push_back(&net.blobs[it->second]);
            }
            Blob top;
            if (L.type == "Convolution")
            {
                if (ins.size() != 1)
                    return planFail("layer " + L.name + ": convolution needs one bottom");
                if (L.group != 1 || L.dilation != 1 || L.stride != 1 || L.kernel % 2 != 1 || L.pad != L.kernel / 2)
                    return planFail("layer " + L.name +
                                    ": only stride-1, group-1, undilated, odd-kernel 'same' convolutions are supported");
                const Blob src = *ins[0];
                auto w = wm.find(L.name);
                if (w == wm.end() || w->second.empty())
                    return planFail("caffemodel has no weights for layer " + L.name);
                const int cin = src.C, cout = L.numOutput, k = L.kernel;
                const BlobData& W = w->second[0];
                if (W.data.size() != (size_t)cout * cin * k * k)
                    return planFail("caffemodel: weight blob of layer " + L.name + " has the wrong size");
                if (L.biasTerm && (w->second.size() < 2 || w->second[1].data.size() != (size_t)cout))
                    return planFail("caffemodel: bias blob of layer " + L.name + " has the wrong size");
                Op op;
                op.kind = OP_CONV;
                op.name = L.name;
                op.level = op.outLevel = src.level;
                op.cin = cin;
                op.cout = cout;
                op.flopsPerPixel = 2.0 * cout * cin * k * k;
                bool patch = false;
                const bool direct = src.isInput && cin * k * k <= 32 && cout <= 64 && cout % 8 == 0 &&
                                    std::getenv("OPB200_NO_FIRSTCONV") == nullptr;
                if (direct)
                {
                    op.kind = OP_FIRSTCONV;
                    op.inC = cin;
                    op.k = k;
                    op.pad = L.pad;
                }
                else if (src.isInput)
                {
                    // first layer: fp32 NCHW input -> bf16 patch tensor; conv becomes 1x1 when Cin*k*k <= 64
                    patch = cin * k * k <= kConvBlockK;
                    Op pre;
                    pre.kind = OP_PATCH;
                    pre.name = L.name + "/input";
                    pre.level = pre.outLevel = src.level;
                    pre.inC = cin;
                    pre.k = patch ? k : 1;
                    pre.pad = patch ? L.pad : 0;
                    pre.out.C = patch ? cin * k * k : cin;
                    pre.out.buf = newBuffer(net, pre.out.C, src.level);
                    op.in.push_back(pre.out);
                    net.ops.push_back(pre);
                    op.k = patch ? 1 : k;
                    op.pad = patch ? 0 : L.pad;
                }
                else
                {
                    op.in = src.segs;
                    op.k = k;
                    op.pad = L.pad;
                }
                if (op.in.size() > (size_t)kConvMaxSegments)
                    return planFail("layer " + L.name + ": more than 4 concatenated inputs");
                op.bn = cout <= 32 ? 32 : (cout <= 64 ? 64 : (cout <= 96 ? 96 : 128));
                op.nTilesN = (cout + op.bn - 1) / op.bn;
                op.chunksTotal = 0;
                for (size_t s = 0; s < op.in.size(); ++s)
                {
                    op.segChunks[s] = (op.in[s].C + kConvBlockK - 1) / kConvBlockK;
                    op.chunksTotal += op.segChunks[s];
                }
                const int taps = op.k * op.k;
                op.kTotal = taps * op.chunksTotal * kConvBlockK;
                const int rows = op.nTilesN * op.bn;
                op.hostW.assign((size_t)rows * op.kTotal, __float2bfloat16(0.f));
                op.hostBias.assign(rows, 0.f);
                op.hostSlope.assign(rows, 1.f);   // identity unless an activation layer follows
                if (direct)
                {
                    // B fragment of mma.m16n8k16 (16x8, "col"): lane (g = lane/4, q = lane%4) holds
                    // {B[2q][g], B[2q+1][g]} and {B[2q+8][g], B[2q+9][g]}, B[kidx][n] = W[n][kidx], kidx = c*k*k + ky*k + kx
                    const int K = cin * k * k;
                    auto wAt = [&](int n, int j) { return (n < cout && j < K) ? W.data[(size_t)n * K + j] : 0.f; };
                    auto pack = [&](float lo, float hi) {
                        const __nv_bfloat16 a = __float2bfloat16(lo), b = __float2bfloat16(hi);
                        uint16_t ua, ub;
                        std::memcpy(&ua, &a, 2);
                        std::memcpy(&ub, &b, 2);
                        return (uint32_t)ua | ((uint32_t)ub << 16);
                    };
                    op.hostFrag.assign(8 * 2 * 32 * 2, 0u);
                    for (int nt = 0; nt < 8; ++nt)
                        for (int s2 = 0; s2 < 2; ++s2)
                            for (int lane = 0; lane < 32; ++lane)
                            {
                                const int g = lane >> 2, q = lane & 3, n = nt * 8 + g, j = 16 * s2 + 2 * q;
                                uint32_t* f = &op.hostFrag[((size_t)(nt * 2 + s2) * 32 + lane) * 2];
                                f[0] = pack(wAt(n, j), wAt(n, j + 1));
                                f[1] = pack(wAt(n, j + 8), wAt(n, j + 9));
                            }
                    op.hostW.clear();
                }
                else if (patch)
                {
                    for (int co = 0; co < cout; ++co)
                        for (int j = 0; j < cin * k * k; ++j)
                            op.hostW[(size_t)co * op.kTotal + j] = __float2bfloat16(W.data[(size_t)co * cin * k * k + j]);
                }
                else
                {
                    int segStart = 0, chunkStart = 0;
                    for (size_t s = 0; s < op.in.size(); ++s)
                    {
                        for (int c = 0; c < op.in[s].C; ++c)
                        {
                            const int ci = segStart + c;
                            // K order of the kernel: chunk-major, then filter column kx, then filter row ky
                            const int chunk = chunkStart + c / kConvBlockK, lane = c % kConvBlockK;
                            for (int ky = 0; ky < op.k; ++ky)
                                for (int kx = 0; kx < op.k; ++kx)
                                {
                                    const size_t kb = ((size_t)chunk * op.k + kx) * op.k + ky;
                                    const size_t col = kb * kConvBlockK + lane;
                                    const int t = ky * op.k + kx;   // Caffe: W[co][ci][ky][kx]
                                    for (int co = 0; co < cout; ++co)
                                        op.hostW[(size_t)co * op.kTotal + col] =
                                            __float2bfloat16(W.data[((size_t)co * cin + ci) * taps + t]);
                                }
                        }
                        segStart += op.in[s].C;
                        chunkStart += op.segChunks[s];
                    }
                    if (segStart != cin)
                        return planFail("layer " + L.name + ": input channel bookkeeping mismatch");
                }
                if (L.biasTerm)
                    for (int co = 0; co < cout; ++co)
                        op.hostBias[co] = w->second[1].data[co];
                // fuse the in-place activation that follows (ReLU / PReLU)
                for (size_t lj = li + 1; lj < def.layers.size(); ++lj)
                {
                    const LayerDef& A = def.layers[lj];
                    const bool inPlace = A.bottoms.size() == 1 && A.tops.size() == 1 && A.bottoms[0] == L.tops[0] &&
                                         A.tops[0] == L.tops[0];
                    if (!inPlace)
                    {
                        bool reads = false;
                        for (const auto& b : A.bottoms)
                            reads = reads || b == L.tops[0];
                        if (reads)
                            break;      // a consumer: no activation in between
                        continue;
                    }
                    if (A.type == "ReLU")
                        std::fill(op.hostSlope.begin(), op.hostSlope.end(), 0.f);
                    else if (A.type == "PReLU")
                    {
                        auto pw = wm.find(A.name);
                        if (pw == wm.end() || pw->second.empty())
                            return planFail("caffemodel has no slopes for layer " + A.name);
                        const auto& sl = pw->second[0].data;
                        if (sl.size() != (size_t)cout && sl.size() != 1)
                            return planFail("caffemodel: PReLU blob of layer " + A.name + " has the wrong size");
                        for (int co = 0; co < cout; ++co)
                            op.hostSlope[co] = sl.size() == 1 ? sl[0] : sl[co];
                    }
                    else
                        return planFail("layer " + A.name + ": unsupported in-place layer type " + A.type);
                    consumed[lj] = true;
                    break;
                }
                op.out = outputView(L.tops[0], cout, src.level);
                top.C = cout;
                top.level = src.level;
                top.segs.push_back(op.out);
                net.ops.push_back(std::move(op));
            }
            else if (L.type == "Pooling")
            {
                if (ins.size() != 1 || ins[0]->isInput || ins[0]->segs.size() != 1)
                    return planFail("layer " + L.name + ": pooling needs one buffer-backed bottom");
                if (L.poolMethod != 0 || L.pad != 0)
                    return planFail("layer " + L.name + ": only unpadded MAX pooling is supported");
                Level lv;
                lv.parent = ins[0]->level;
                lv.k = L.kernel;
                lv.stride = L.stride;
                net.levels.push_back(lv);
                Op op;
                op.kind = OP_POOL;
                op.name = L.name;
                op.level = ins[0]->level;
                op.outLevel = (int)net.levels.size() - 1;
                op.in = ins[0]->segs;
                op.k = L.kernel;
                op.stride = L.stride;
                op.out = outputView(L.tops[0], ins[0]->C, op.outLevel);
                if (op.in[0].C % 8 != 0 || op.in[0].cOff % 8 != 0)
                    return planFail("layer " + L.name + ": pooled blob must have a multiple of 8 channels");
                top.C = ins[0]->C;
                top.level = op.outLevel;
                top.segs.push_back(op.out);
                net.ops.push_back(std::move(op));
            }
            else if (L.type == "Concat")
            {
                if (L.axis != 1)
                    return planFail("layer " + L.name + ": only channel concatenation (axis 1) is supported");
                top.level = ins.empty() ? 0 : ins[0]->level;
                for (const Blob* b : ins)
                {
                    if (b->isInput || b->level != top.level)
                        return planFail("layer " + L.name + ": concat members must be same-size intermediate blobs");
                    top.C += b->C;
                }
                if (owningConcats.count(L.tops[0]))
                {
                    View v;
                    v.buf = concatBuffer.at(L.tops[0]);
                    v.cOff = 0;
                    v.C = top.C;
                    top.segs.push_back(v);
                }
                else
                    for (const Blob* b : ins)
                        top.segs.insert(top.segs.end(), b->segs.begin(), b->segs.end());
            }
            else
                return planFail("layer " + L.name + ": unsupported layer type " + L.type);
            net.blobIndex[L.tops[0]] = (int)net.blobs.size();
            net.blobs.push_back(top);
        }

        // ---- output blob: members written in fp32 NCHW straight from the producing conv's epilogue ------
        auto ob = net.blobIndex.find(net.lastBlobName);
        if (ob == net.blobIndex.end())
        {
            setError("The output blob is a nullptr. Did you use the same name than the prototxt? (Used: %s).",
                     net.lastBlobName.c_str());
            return OPB200_ERR_INVALID;
        }
        net.outBlob = ob->second;
        const Blob outB = net.blobs[net.outBlob];
        if (outB.isInput)
            return planFail("the output blob cannot be the network input");
        net.outC = outB.C;
        net.outLevel = outB.level;
        int cOff = 0;
        for (const View& seg : outB.segs)
        {
            int left = seg.C, segOff = seg.cOff;
            while (left > 0)   // a segment may span several co-located conv outputs
            {
                bool found = false;
                for (auto& op : net.ops)
                    if (op.kind == OP_CONV && op.out.buf == seg.buf && op.out.cOff == segOff && op.out.C <= left)
                    {
                        op.out32Off = cOff;
                        cOff += op.out.C;
                        segOff += op.out.C;
                        left -= op.out.C;
                        found = true;
                        break;
                    }
                if (!found)
                {
                    Op cv;
                    cv.kind = OP_CONVERT;
                    cv.name = net.lastBlobName + "/convert";
                    cv.level = cv.outLevel = outB.level;
                    View v = seg;
                    v.cOff = segOff;
                    v.C = left;
                    cv.in.push_back(v);
                    cv.out32Off = cOff;
                    cOff += left;
                    left = 0;
                    net.ops.push_back(cv);
                }
            }
        }
        // ---- which conv outputs does anybody read in bf16? ------------------------------------------------
        for (auto& op : net.ops)
        {
            if (op.kind != OP_CONV)
                continue;
            bool read = false;
            for (const auto& other : net.ops)
                for (const View& v : other.in)
                    read = read || overlaps(v, op.out);
            op.storeBf16 = read || op.out32Off < 0;
        }
        // ---- fold 2x2 / stride 2 poolings into the epilogue of the conv that feeds them ----------------------
        // (conv -> in-place ReLU -> pool with no other reader of the conv output: the full-resolution blob is
        // then never written or read back; kernel: conv_sm100.cu, poolFused)
        static const bool fusePools = std::getenv("OPB200_NO_POOL_FUSION") == nullptr;
        for (size_t pi = 0; fusePools && pi < net.ops.size(); ++pi)
        {
            const Op& pool = net.ops[pi];
            if (pool.kind != OP_POOL || pool.k != 2 || pool.stride != 2 || pool.in.size() != 1)
                continue;
            int producer = -1, readers = 0;
            for (size_t ci = 0; ci < net.ops.size(); ++ci)
            {
                const Op& c = net.ops[ci];
                if (c.kind == OP_CONV && c.out.buf == pool.in[0].buf && c.out.cOff == pool.in[0].cOff &&
                    c.out.C == pool.in[0].C && c.out32Off < 0)
                    producer = (int)ci;
                for (const View& v : c.in)
                    readers += overlaps(v, pool.in[0]) ? 1 : 0;
            }
            if (producer < 0 || readers != 1 || (size_t)producer > pi)
                continue;
            Op& conv = net.ops[producer];
            conv.poolFused = true;
            conv.poolOut = pool.out;
            conv.poolLevel = pool.outLevel;
            conv.storeBf16 = false;
            net.ops.erase(net.ops.begin() + pi);
            --pi;
        }
        return OPB200_OK;
    }

    static opb200_status uploadWeights(Net& net)
    {
        for (auto& op : net.ops)
        {
            if (op.kind != OP_CONV && op.kind != OP_FIRSTCONV)
                continue;
            if (op.kind == OP_FIRSTCONV)
            {
                OPB_CUDA(cudaMalloc(&op.dW, op.hostFrag.size() * sizeof(uint32_t)));
                OPB_CUDA(cudaMemcpy(op.dW, op.hostFrag.data(), op.hostFrag.size() * sizeof(uint32_t),
                                    cudaMemcpyHostToDevice));
            }
            else
            {
                OPB_CUDA(cudaMalloc(&op.dW, op.hostW.size() * sizeof(__nv_bfloat16)));
                OPB_CUDA(cudaMemcpy(op.dW, op.hostW.data(), op.hostW.size() * sizeof(__nv_bfloat16),
                                    cudaMemcpyHostToDevice));
            }
            OPB_CUDA(cudaMalloc(&op.dBias, op.hostBias.size() * sizeof(float)));
            OPB_CUDA(cudaMalloc(&op.dSlope, op.hostSlope.size() * sizeof(float)));
            OPB_CUDA(cudaMemcpy(op.dBias, op.hostBias.data(), op.hostBias.size() * sizeof(float), cudaMemcpyHostToDevice));
            OPB_CUDA(cudaMemcpy(op.dSlope, op.hostSlope.data(), op.hostSlope.size() * sizeof(float),
                                cudaMemcpyHostToDevice));
            std::vector<__nv_bfloat16>().swap(op.hostW);
        }
        return OPB200_OK;
    }

    // Host-only part of a reshape: the size of every level for an (N, H, W) input and which levels are stored
    // frame-interleaved, with how many separator columns (also behind opb200_net_describe_shape, for the CPU tests).
    static void layoutLevels(Net& net, int N, int H, int W)
    {
        net.lvH.assign(net.levels.size(), 0);
        net.lvW.assign(net.levels.size(), 0);
        net.lvH[0] = H;
        net.lvW[0] = W;
        for (size_t l = 1; l < net.levels.size(); ++l)
        {
            const Level& lv = net.levels[l];
            // Caffe pooling output size, pad 0 (pooling_layer.cpp:91-97): ceil((in - k) / stride) + 1
            auto pooled = [&](int in) {
                return in < lv.k ? 1 : (int)std::ceil((double)(in - lv.k) / lv.stride) + 1;
            };
            net.lvH[l] = pooled(net.lvH[lv.parent]);
            net.lvW[l] = pooled(net.lvW[lv.parent]);
        }
        // Flat levels (conv_params.h): only where every reader and writer of the level's buffers is the conv kernel
        // (TMA loads, TMA stores, fused pooled stores) or the fp32 converter, and where it saves M tiles.
        const bool flatEnabled = !(std::getenv("OPB200_CONV_FLAT") && std::atoi(std::getenv("OPB200_CONV_FLAT")) == 0);
        net.lvFlat.assign(net.levels.size(), 0);
        net.lvSep.assign(net.levels.size(), 1);
        for (size_t l = 0; l < net.levels.size(); ++l)
        {
            const int w = net.lvW[l], h = net.lvH[l];
            // The separator columns are the horizontal zero padding between neighbouring frames, so there must be as many
            // as the widest padding of the level's convs: 1 for 3x3, 3 for the 7x7 layers of COCO / hand / face. (Round 1
            // and most of round 2 always used one: with pad 3 a 7x7 filter then read two columns of the NEIGHBOURING frame
            // -- wrong outputs for those graphs whenever the batch was large enough for the flat layout to be chosen.)
            int sep = 1;
            for (const auto& op : net.ops)
                if (op.kind == OP_CONV && op.level == (int)l && op.pad > sep)
                    sep = op.pad;
            net.lvSep[l] = sep;
            const long long tilesPlain = (long long)N * ((w + kConvTileW - 1) / kConvTileW);
            const long long tilesFlat = ((long long)N * (w + sep) + kConvTileW - 1) / kConvTileW;
            bool ok = flatEnabled && N > 1 && tilesFlat < tilesPlain && h > 0;
            for (const auto& op : net.ops)
            {
                const bool touches = op.level == (int)l || op.outLevel == (int)l || (op.poolFused && op.poolLevel == (int)l);
                if (!touches)
                    continue;
                if (op.kind == OP_CONV)
                {
                    if (op.level == (int)l && op.poolFused)
                        ok = false;   // a flat level whose conv also pools: pooled windows would cross frame borders
                }
                else if (op.kind != OP_CONVERT)
                    ok = false;       // patch / first conv / stand-alone pooling address pixels as [n][y][x]
            }
            net.lvFlat[l] = ok ? 1 : 0;
        }
    }

    // ------------------------------------------------------------------------------------------------
    // reshape: level sizes, arena, tensor maps (reference: reshapeNetCaffe, netCaffe.cpp:105-122)
    // ------------------------------------------------------------------------------------------------
    // `own` is the net's stream (created non-blocking, so the legacy default stream does not order against it): the
    // previous shape's work on it and on the caller's stream `user` (which may still read the old output blob) is
    // drained before the graph and the arena are torn down, and the zero fill of the new arena is enqueued on `own`,
    // ahead of everything the following forward puts there.
    static opb200_status reshapeNet(Net& net, int N, int H, int W, cudaStream_t own, cudaStream_t user)
    {
        OPB_CUDA(cudaStreamSynchronize(own));
        OPB_CUDA(cudaStreamSynchronize(user));
        if (net.graphExec)
        {
            cudaGraphExecDestroy(net.graphExec);
            net.graphExec = nullptr;
        }
        layoutLevels(net, N, H, W);
        size_t total = 0;
        auto carve = [&](size_t bytes) {
            const size_t at = total;
            total += (bytes + 1023) / 1024 * 1024;
            return at;
        };
        for (auto& b : net.buffers)
        {
            const int w = net.lvW[b.level] + (net.lvFlat[b.level] ? net.lvSep[b.level] : 0);
            b.offset = carve((size_t)N * net.lvH[b.level] * w * b.C * 2);
        }
        net.inputBytes = (size_t)N * net.inputC * H * W * 4;
        net.outBytes = (size_t)N * net.outC * net.lvH[net.outLevel] * net.lvW[net.outLevel] * 4;
        const size_t inOff = carve(net.inputBytes), outOff = carve(net.outBytes);
        if (total > net.arenaBytes)
        {
            if (net.arena)
                cudaFree(net.arena);
            net.arena = nullptr;
            net.arenaBytes = 0;
            OPB_CUDA(cudaMalloc(&net.arena, total));
            net.arenaBytes = total;
        }
        OPB_CUDA(cudaMemsetAsync(net.arena, 0, total, own));
        net.dInput = reinterpret_cast<float*>(net.arena + inOff);
        net.dOut32 = reinterpret_cast<float*>(net.arena + outOff);
        net.N = N;
        net.H = H;
        net.W = W;
        net.flopsPerForward = 0;
        auto viewPtr = [&](const View& v) {
            return net.arena + net.buffers[v.buf].offset + (size_t)v.cOff * 2;
        };
        // Serpentine tile order: every other conv launch walks its tiles last to first, so a layer starts on the part
        // of its input the previous launch wrote last (still in L2) instead of the part written first (evicted when a
        // batch's activations exceed the L2). OPB200_CONV_SERPENTINE=0 turns it off.
        static const bool serpentine = !(std::getenv("OPB200_CONV_SERPENTINE") && std::atoi(std::getenv("OPB200_CONV_SERPENTINE")) == 0);
        int convIndex = 0;
        for (auto& op : net.ops)
        {
            if (op.kind == OP_FIRSTCONV)
                net.flopsPerForward += op.flopsPerPixel * (double)N * net.lvH[op.level] * net.lvW[op.level];
            if (op.kind != OP_CONV)
                continue;
            const int h = net.lvH[op.level], w = net.lvW[op.level];
            ConvParams& p = op.cp;
            std::memset(&p, 0, sizeof(p));
            p.reverse = (serpentine && (convIndex++ & 1) == 0) ? 1 : 0;   // the first-layer kernel runs first to last
            const bool flat = net.lvFlat[op.level] != 0;
            const int sep = net.lvSep[op.level];
            const int kw = flat ? N * (w + sep) : w, kn = flat ? 1 : N;   // the image the kernel sees
            p.flatPitch = flat ? w + sep : 0;
            p.flatW = w;
            p.flatN = N;
            p.tilesX = (kw + kConvTileW - 1) / kConvTileW;
            p.tilesY = (h + kConvTileH - 1) / kConvTileH;
            p.numMTiles = kn * p.tilesX * p.tilesY;
            p.H = h;
            p.W = kw;
            p.N = kn;
            p.k = op.k;
            p.pad = op.pad;
            // 3x3: one halo box per chunk serves all nine taps; larger kernels keep one box per filter column
            // (a 7x7 halo box pair would not leave room for the weight pipeline)
            static const bool haloBox = std::getenv("OPB200_CONV_NO_HALO") == nullptr;
            p.kxSpan = (haloBox && op.k == 3) ? op.k : 1;
            p.boxW = kConvTileW + p.kxSpan - 1;
            p.nSeg = (int)op.in.size();
            p.chunksTotal = op.chunksTotal;
            // Small batches: a 46x82 level has 33 M tiles per frame, i.e. 17 tile pairs (9 quads in pair mode) for 148 SMs.
            // When the default tiling leaves most SMs idle, work items shrink -- to ONE M tile per CTA, then to narrower
            // Cout tiles (a divisor of the weight layout's tile, so weights / bias tables are shared) -- until about three
            // quarters of the SMs have one. CTA pairs stay where they pay: a cta_group::2 MMA (M = 256) retires in the time a
            // cta_group::1 MMA (M = 128) needs (timeline of a 128->128 layer at batch 1: 64 against 171 clk per MMA).
            // OPB200_CONV_SMALL=0 disables.
            static const bool smallOn = !(std::getenv("OPB200_CONV_SMALL") && std::atoi(std::getenv("OPB200_CONV_SMALL")) == 0);
            static const int minBn = std::getenv("OPB200_CONV_SMALL_MINBN") ? std::atoi(std::getenv("OPB200_CONV_SMALL_MINBN")) : 32;
            op.bnEff = op.bn;
            p.tilesPerItem = 2;
            int smallPair = -1;   // -1: the default pair rule decides
            {
                const int wantCtas = (kNumSMsB200 * 3) / 4;
                const bool pairShape = op.k >= 3 && op.chunksTotal >= 2;
                auto ctas = [&](int bn, int tpi, bool pair) {
                    const int perGroup = (pair ? 2 : 1) * tpi;
                    return (long long)((p.numMTiles + perGroup - 1) / perGroup) * ((op.cout + bn - 1) / bn) * (pair ? 2 : 1);
                };
                const bool defaultPair = op.bn >= 96 && pairShape;
                if (smallOn && ctas(op.bn, 2, defaultPair) < wantCtas)
                {
                    p.tilesPerItem = 1;
                    const int cands[3] = {op.bn, op.bn == 128 ? 64 : 32, 32};
                    for (int ci = 0; ci < 3; ++ci)
                    {
                        if (op.bn % cands[ci] != 0 || cands[ci] > op.bnEff || cands[ci] < minBn)
                            continue;
                        if (pairShape && op.bn >= 64 && cands[ci] < 64)
                            continue;   // rather 68 CTAs issuing pair MMAs than 132 issuing single-CTA ones
                        op.bnEff = cands[ci];
                        smallPair = (pairShape && op.bnEff >= 64) ? 1 : 0;
                        if (ctas(op.bnEff, 1, smallPair == 1) >= wantCtas)
                            break;
                    }
                    if (smallPair < 0)
                        smallPair = defaultPair ? 1 : 0;
                }
            }
            p.nTilesN = (op.cout + op.bnEff - 1) / op.bnEff;
            p.cout = op.cout;
            p.bias = op.dBias;
            p.slope = op.dSlope;
            p.storeBf16 = op.storeBf16 ? 1 : 0;
            p.actMode = 1;
            for (int co = 0; co < op.cout; ++co)
                if (op.hostSlope[co] != 0.f)
                    p.actMode = 0;
            if (op.out32Off >= 0)
                p.actMode = 0;   // the fast ReLU path clamps in the bf16 conversion only: fp32 outputs take the general one
            p.out32 = op.out32Off >= 0 ? net.dOut32 : nullptr;
            p.out32C = net.outC;
            p.out32Off = op.out32Off >= 0 ? op.out32Off : 0;
            if (op.chunksTotal > kConvMaxChunks)
                OPB_FAIL(OPB200_ERR_UNSUPPORTED, "conv: more than 2048 input channels");
            for (int s = 0, chunk = 0; s < p.nSeg; ++s)
            {
                p.segChunks[s] = op.segChunks[s];
                for (int c = 0; c < op.segChunks[s]; ++c, ++chunk)
                {
                    const int valid = std::min(kConvBlockK, op.in[s].C - c * kConvBlockK);
                    p.chunkSteps[chunk] = (uint8_t)((valid + 15) / 16);
                }
                const opb200_status st = makeActivationMap(&p.aMap[s], viewPtr(op.in[s]), op.in[s].C, kw, h, kn,
                                                           net.buffers[op.in[s].buf].C, kConvBlockK, p.boxW,
                                                           kConvTileH + op.k - 1, true);
                if (st != OPB200_OK)
                    return st;
            }
            // CTA-pair kernel (cta_group::2, conv_sm100.cu): halves the weight bytes each SM pulls from L2. Measured
            // per shape (profiles/r01_net_layer_table_pair_all_b44.txt): +15-25 % on 3x3 layers with >= 2 input chunks and
            // BN >= 96, slower on 1x1 layers, on the 64-wide trunk layers and on single-chunk layers (short main
            // loops, where the cross-CTA hand-offs cost more than the traffic saves).
            // OPB200_CONV_PAIR: 0 = never, 2 = every layer with BN >= 64, unset / 1 = the rule above.
            static const int pairEnv = std::getenv("OPB200_CONV_PAIR") ? std::atoi(std::getenv("OPB200_CONV_PAIR")) : 1;
            p.pairMode = (pairEnv == 2 && op.bn >= 64) || (pairEnv == 1 && op.bn >= 96 && op.k >= 3 && op.chunksTotal >= 2);
            if (smallPair >= 0 && pairEnv != 0)
                p.pairMode = smallPair;
            // weight boxes hold the k filter rows of one filter column for 3x3 layers (OPB200_CONV_BGROUP=0: one K block per box)
            static const bool bGroupOn = !(std::getenv("OPB200_CONV_BGROUP") && std::atoi(std::getenv("OPB200_CONV_BGROUP")) == 0);
            p.bGroup = (bGroupOn && op.k == 3) ? 3 : 1;
            opb200_status st = makeWeightMap(&p.bMap, op.dW, op.kTotal, op.nTilesN * op.bn, p.pairMode ? op.bnEff / 2 : op.bnEff,
                                             p.bGroup);
            if (st != OPB200_OK)
                return st;
            p.poolFused = op.poolFused ? 1 : 0;
            p.poolOutFlat = (op.poolFused && net.lvFlat[op.poolLevel]) ? 1 : 0;
            if (op.poolFused && p.poolOutFlat)   // pooled blob on a flat level: per-frame store map
                st = makeFlatStoreMap(&p.oMap, viewPtr(op.poolOut), op.poolOut.C, net.lvW[op.poolLevel],
                                      net.lvH[op.poolLevel], N, net.lvW[op.poolLevel] + net.lvSep[op.poolLevel],
                                      net.buffers[op.poolOut.buf].C, 32, kConvTileW / 2, kConvTileH / 2);
            else if (op.poolFused)   // the pooled blob, one (32 ch, 4, 8) box per tile
                st = makeActivationMap(&p.oMap, viewPtr(op.poolOut), op.poolOut.C, net.lvW[op.poolLevel],
                                       net.lvH[op.poolLevel], N, net.buffers[op.poolOut.buf].C, 32, kConvTileW / 2,
                                       kConvTileH / 2, false);
            else   // (flat level: one image of width N * (W + sep); the epilogue writes zeros into the separator columns)
                st = makeActivationMap(&p.oMap, viewPtr(op.out), op.out.C, kw, h, kn, net.buffers[op.out.buf].C, 32,
                                       kConvTileW, kConvTileH, false);
            if (st != OPB200_OK)
                return st;
            net.flopsPerForward += op.flopsPerPixel * (double)N * h * w;
        }
        return OPB200_OK;
    }

    // ------------------------------------------------------------------------------------------------
    // forward
    // ------------------------------------------------------------------------------------------------
    struct NetHandle
    {
        Net net;
        cudaStream_t own = nullptr;
        cudaEvent_t evIn = nullptr, evOut = nullptr;
        int launchesPerForward = 0;
    };

    static opb200_status enqueueOps(Net& net, cudaStream_t st, int* launches)
    {
        auto viewPtr = [&](const View& v) { return net.arena + net.buffers[v.buf].offset + (size_t)v.cOff * 2; };
        int n = 0;
        for (auto& op : net.ops)
        {
            opb200_status s = OPB200_OK;
            const int h = net.lvH[op.level], w = net.lvW[op.level];
            switch (op.kind)
            {
            case OP_PATCH:
                s = launchInputPatch(viewPtr(op.out), net.dInput, net.N, op.inC, h, w, op.k, op.pad,
                                     net.buffers[op.out.buf].C, st);
                break;
            case OP_CONV: s = launchConv(op.cp, op.bnEff, net.gridCap, st); break;
            case OP_FIRSTCONV:
            {
                bool reluOnly = true;
                for (float sl : op.hostSlope)
                    reluOnly = reluOnly && sl == 0.f;
                s = launchFirstConv(viewPtr(op.out), net.dInput, reinterpret_cast<const uint32_t*>(op.dW), op.dBias,
                                    op.dSlope, net.N, op.inC, h, w, op.k, op.pad, op.cout, net.buffers[op.out.buf].C,
                                    reluOnly, st);
                break;
            }
            case OP_POOL:
                s = launchMaxPool(viewPtr(op.out), viewPtr(op.in[0]), net.N, op.in[0].C, h, w, net.lvH[op.outLevel],
                                  net.lvW[op.outLevel], op.k, op.stride, net.buffers[op.in[0].buf].C,
                                  net.buffers[op.out.buf].C, st);
                break;
            case OP_CONVERT:
            {
                const long long sc = net.buffers[op.in[0].buf].C;
                const bool flat = net.lvFlat[op.level] != 0;
                const int pitch = w + net.lvSep[op.level];
                s = launchNhwcToNchw(net.dOut32, viewPtr(op.in[0]), net.N, op.in[0].C, h, w, (int)sc, net.outC,
                                     op.out32Off, st, flat ? pitch * sc : 0, flat ? (long long)net.N * pitch * sc : 0);
            }
                break;
            }
            if (s != OPB200_OK)
                return s;
            ++n;
        }
        if (launches)
            *launches = n;
        return OPB200_OK;
    }

    static opb200_status forwardNet(NetHandle& h, const float* input, bool onDevice, int N, int C, int H, int W,
                                    cudaStream_t user)
    {
        Net& net = h.net;
        OPB_CUDA(cudaSetDevice(net.gpu));
        if (N != net.N || H != net.H || W != net.W)
        {
            const opb200_status s = reshapeNet(net, N, H, W, h.own, user);
            if (s != OPB200_OK)
                return s;
        }
        // fork from the caller's stream onto the net's own (capturable) stream
        OPB_CUDA(cudaEventRecord(h.evIn, user));
        OPB_CUDA(cudaStreamWaitEvent(h.own, h.evIn, 0));
        if (input != net.dInput)   // (opb200_net_input_buffer: the caller wrote the input in place)
            OPB_CUDA(cudaMemcpyAsync(net.dInput, input, net.inputBytes,
                                     onDevice ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, h.own));
        if (net.useGraph)
        {
            if (!net.graphExec)
            {
                // warm launch outside capture (sets function attributes, surfaces launch errors early)
                opb200_status s = enqueueOps(net, h.own, &h.launchesPerForward);
                if (s != OPB200_OK)
                    return s;
                OPB_CUDA(cudaStreamSynchronize(h.own));
                cudaGraph_t graph = nullptr;
                OPB_CUDA(cudaStreamBeginCapture(h.own, cudaStreamCaptureModeThreadLocal));
                s = enqueueOps(net, h.own, nullptr);
                const cudaError_t e = cudaStreamEndCapture(h.own, &graph);
                if (s != OPB200_OK || e != cudaSuccess)
                {
                    if (graph)
                        cudaGraphDestroy(graph);   // a failed capture must not leak what it recorded
                    if (s != OPB200_OK)
                        return s;
                    OPB_CUDA(e);
                }
                const cudaError_t ei = cudaGraphInstantiate(&net.graphExec, graph, 0);
                cudaGraphDestroy(graph);
                OPB_CUDA(ei);
            }
            OPB_CUDA(cudaGraphLaunch(net.graphExec, h.own));
        }
        else
        {
            const opb200_status s = enqueueOps(net, h.own, &h.launchesPerForward);
            if (s != OPB200_OK)
                return s;
        }
        OPB_CUDA(cudaEventRecord(h.evOut, h.own));
        OPB_CUDA(cudaStreamWaitEvent(user, h.evOut, 0));
        return OPB200_OK;
    }
}

// ----------------------------------------------------------------------------------------------------
// C ABI (include/opb200.h)
// ----------------------------------------------------------------------------------------------------
using namespace opb;

struct opb200_net
{
    NetHandle h;
};

static bool fileExists(const char* p)
{
    FILE* f = p ? std::fopen(p, "rb") : nullptr;
    if (f)
        std::fclose(f);
    return f != nullptr;
}

extern "C" opb200_status opb200_net_create(opb200_net** out, const char* prototxt, const char* caffemodel, int gpu_id,
                                           const char* last_blob_name)
{
    OPB_REQUIRE(out && prototxt && caffemodel, "net_create: null argument");
    *out = nullptr;
    // same wording as NetCaffe's constructor checks (reference: src/openpose/net/netCaffe.cpp:55-66)
    static const char* hint =
        ".\nPossible causes:\n\t1. Not downloading the OpenPose trained models.\n\t2. Not running OpenPose from the "
        "root directory (i.e., where the `model` folder is located, but do not move the `model` folder!).\n\t3. Using "
        "paths with spaces.";
    if (!fileExists(prototxt))
    {
        setError("Prototxt file not found: %s%s", prototxt, hint);
        return OPB200_ERR_IO;
    }
    if (!fileExists(caffemodel))
    {
        setError("Caffe trained model file not found: %s%s", caffemodel, hint);
        return OPB200_ERR_IO;
    }
    std::unique_ptr<opb200_net> n(new opb200_net());
    Net& net = n->h.net;
    net.gpu = gpu_id;
    net.lastBlobName = (last_blob_name && *last_blob_name) ? last_blob_name : "net_output";
    net.useGraph = std::getenv("OPB200_NO_GRAPH") == nullptr;
    if (const char* g = std::getenv("OPB200_GRID_CAP"))
        net.gridCap = std::max(1, std::atoi(g));
    std::string text, err;
    if (!readFile(prototxt, text) || !parsePrototxt(text, net.def, err))
    {
        setError("%s (%s)", err.empty() ? "prototxt: cannot read file" : err.c_str(), prototxt);
        return OPB200_ERR_IO;
    }
    WeightMap wm;
    if (!readCaffemodel(caffemodel, wm, err))
    {
        setError("%s", err.c_str());
        return OPB200_ERR_IO;
    }
    opb200_status s = planNet(net, wm);
    if (s != OPB200_OK)
        return s;
    OPB_CUDA(cudaSetDevice(gpu_id));
    s = uploadWeights(net);
    if (s != OPB200_OK)
        return s;
    OPB_CUDA(cudaStreamCreateWithFlags(&n->h.own, cudaStreamNonBlocking));
    OPB_CUDA(cudaEventCreateWithFlags(&n->h.evIn, cudaEventDisableTiming));
    OPB_CUDA(cudaEventCreateWithFlags(&n->h.evOut, cudaEventDisableTiming));
    *out = n.release();
    return OPB200_OK;
}

extern "C" void opb200_net_destroy(opb200_net* n)
{
    if (!n)
        return;
    cudaSetDevice(n->h.net.gpu);
    if (n->h.own)
    {
        cudaStreamSynchronize(n->h.own);
        cudaStreamDestroy(n->h.own);
    }
    if (n->h.evIn)
        cudaEventDestroy(n->h.evIn);
    if (n->h.evOut)
        cudaEventDestroy(n->h.evOut);
    delete n;
}

extern "C" opb200_status opb200_net_input_buffer(opb200_net* n, const int shape[4], float** dev_ptr, void* stream)
{
    OPB_REQUIRE(n && shape && dev_ptr, "net_input_buffer: null argument");
    Net& net = n->h.net;
    if (shape[0] <= 0 || shape[2] <= 0 || shape[3] <= 0)
    {
        setError("The Array inputData cannot be empty.");
        return OPB200_ERR_INVALID;
    }
    OPB_REQUIRE(shape[1] == net.inputC, "The Array inputData must have 4 dimensions: [batch size, 3 (RGB), height, width].");
    OPB_CUDA(cudaSetDevice(net.gpu));
    if (shape[0] != net.N || shape[2] != net.H || shape[3] != net.W)
    {
        const opb200_status s = reshapeNet(net, shape[0], shape[2], shape[3], n->h.own, static_cast<cudaStream_t>(stream));
        if (s != OPB200_OK)
            return s;
        // the zero fill of a fresh arena runs on the net's stream: the caller's writes on `stream` come after it
        OPB_CUDA(cudaEventRecord(n->h.evOut, n->h.own));
        OPB_CUDA(cudaStreamWaitEvent(static_cast<cudaStream_t>(stream), n->h.evOut, 0));
    }
    *dev_ptr = net.dInput;
    return OPB200_OK;
}

extern "C" opb200_status opb200_net_forward(opb200_net* n, const float* input, int input_on_device,
                                            const int shape[4], void* stream)
{
    OPB_REQUIRE(n != nullptr, "net_forward: null net");
    // same checks and wording as NetCaffe::forwardPass (reference: src/openpose/net/netCaffe.cpp:218-222)
    if (!input || !shape || shape[0] <= 0 || shape[2] <= 0 || shape[3] <= 0)
    {
        setError("The Array inputData cannot be empty.");
        return OPB200_ERR_INVALID;
    }
    if (shape[1] != n->h.net.inputC)
    {
        setError("The Array inputData must have 4 dimensions: [batch size, 3 (RGB), height, width].");
        return OPB200_ERR_INVALID;
    }
    return forwardNet(n->h, input, input_on_device != 0, shape[0], shape[1], shape[2], shape[3],
                      static_cast<cudaStream_t>(stream));
}

extern "C" opb200_status opb200_net_output(opb200_net* n, const float** dev_ptr, int shape[4])
{
    OPB_REQUIRE(n && dev_ptr && shape, "net_output: null argument");
    Net& net = n->h.net;
    OPB_REQUIRE(net.N > 0, "net_output: forward has not run yet");
    *dev_ptr = net.dOut32;
    shape[0] = net.N;
    shape[1] = net.outC;
    shape[2] = net.lvH[net.outLevel];
    shape[3] = net.lvW[net.outLevel];
    return OPB200_OK;
}

extern "C" opb200_status opb200_net_info(opb200_net* n, opb200_net_info_t* info)
{
    OPB_REQUIRE(n && info, "net_info: null argument");
    Net& net = n->h.net;
    int convs = 0;
    for (const auto& op : net.ops)
        convs += (op.kind == OP_CONV || op.kind == OP_FIRSTCONV);
    info->conv_layers = convs;
    info->kernel_launches_per_forward = (int)net.ops.size();
    info->flops_per_forward = net.flopsPerForward;
    info->arena_bytes = (double)net.arenaBytes;
    info->output_channels = net.outC;
    return OPB200_OK;
}

// Test hook: copy any intermediate blob (bf16 NHWC views) to a host fp32 NCHW array. Synchronises the device.
extern "C" opb200_status opb200_net_read_blob(opb200_net* n, const char* name, float* host_nchw, size_t capacity,
                                              int shape[4])
{
    OPB_REQUIRE(n && name && shape, "net_read_blob: null argument");
    Net& net = n->h.net;
    auto it = net.blobIndex.find(name);
    if (it == net.blobIndex.end() || net.blobs[it->second].isInput)
    {
        setError("net_read_blob: no intermediate blob named %s", name);
        return OPB200_ERR_INVALID;
    }
    OPB_REQUIRE(net.N > 0, "net_read_blob: forward has not run yet");
    const Blob& b = net.blobs[it->second];
    const int h = net.lvH[b.level], w = net.lvW[b.level];
    shape[0] = net.N;
    shape[1] = b.C;
    shape[2] = h;
    shape[3] = w;
    const size_t count = (size_t)net.N * b.C * h * w;
    if (!host_nchw)
        return OPB200_OK;   // shape query
    OPB_REQUIRE(capacity >= count, "net_read_blob: host buffer too small");
    OPB_CUDA(cudaSetDevice(net.gpu));
    OPB_CUDA(cudaDeviceSynchronize());
    float* tmp = nullptr;
    OPB_CUDA(cudaMalloc(&tmp, count * sizeof(float)));
    int cOff = 0;
    opb200_status s = OPB200_OK;
    for (const View& v : b.segs)
    {
        const long long sc = net.buffers[v.buf].C;
        const bool flat = net.lvFlat[b.level] != 0;
        const int pitch = w + net.lvSep[b.level];
        s = launchNhwcToNchw(tmp, net.arena + net.buffers[v.buf].offset + (size_t)v.cOff * 2, net.N, v.C, h, w,
                             (int)sc, b.C, cOff, nullptr, flat ? pitch * sc : 0,
                             flat ? (long long)net.N * pitch * sc : 0);
        if (s != OPB200_OK)
            break;
        cOff += v.C;
    }
    if (s == OPB200_OK)
    {
        const cudaError_t e = cudaMemcpy(host_nchw, tmp, count * sizeof(float), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess)
        {
            setError("CUDA error %s in net_read_blob", cudaGetErrorName(e));
            s = OPB200_ERR_CUDA;
        }
    }
    cudaFree(tmp);
    return s;
}

// Host-only: parse + plan, no CUDA call. Writes one line per planned op into `text` (for tests and debugging):
//   conv <name> cin cout k bn tilesN segs=[buf:off:C,...] out=buf:off:C out32=<off|-1> bf16=<0|1>
extern "C" opb200_status opb200_net_describe_shape(const char* prototxt, const char* caffemodel, const char* last_blob_name,
                                                   int n, int h, int w, char* text, size_t capacity)
{
    OPB_REQUIRE(prototxt && caffemodel && text && capacity > 0, "net_describe_shape: null argument");
    OPB_REQUIRE(n > 0 && h > 0 && w > 0, "net_describe_shape: empty input shape");
    Net net;
    net.lastBlobName = (last_blob_name && *last_blob_name) ? last_blob_name : "net_output";
    std::string src, err;
    if (!readFile(prototxt, src) || !parsePrototxt(src, net.def, err))
    {
        setError("%s (%s)", err.empty() ? "prototxt: cannot read file" : err.c_str(), prototxt);
        return OPB200_ERR_IO;
    }
    WeightMap wm;
    if (!readCaffemodel(caffemodel, wm, err))
    {
        setError("%s", err.c_str());
        return OPB200_ERR_IO;
    }
    const opb200_status s = planNet(net, wm);
    if (s != OPB200_OK)
        return s;
    layoutLevels(net, n, h, w);
    std::string out;
    char line[256];
    for (size_t l = 0; l < net.levels.size(); ++l)
    {
        int maxPad = 0, convs = 0;
        for (const auto& op : net.ops)
            if (op.kind == OP_CONV && op.level == (int)l)
            {
                ++convs;
                maxPad = std::max(maxPad, op.pad);
            }
        std::snprintf(line, sizeof(line), "level %d h=%d w=%d convs=%d max_pad=%d flat=%d sep=%d\n", (int)l, net.lvH[l], net.lvW[l],
                      convs, maxPad, net.lvFlat[l], net.lvFlat[l] ? net.lvSep[l] : 0);
        out += line;
    }
    OPB_REQUIRE(out.size() + 1 <= capacity, "net_describe_shape: text buffer too small");
    std::memcpy(text, out.c_str(), out.size() + 1);
    return OPB200_OK;
}

extern "C" opb200_status opb200_net_describe(const char* prototxt, const char* caffemodel, const char* last_blob_name,
                                             char* text, size_t capacity)
{
    OPB_REQUIRE(prototxt && caffemodel && text && capacity > 0, "net_describe: null argument");
    Net net;
    net.lastBlobName = (last_blob_name && *last_blob_name) ? last_blob_name : "net_output";
    std::string src, err;
    if (!readFile(prototxt, src) || !parsePrototxt(src, net.def, err))
    {
        setError("%s (%s)", err.empty() ? "prototxt: cannot read file" : err.c_str(), prototxt);
        return OPB200_ERR_IO;
    }
    WeightMap wm;
    if (!readCaffemodel(caffemodel, wm, err))
    {
        setError("%s", err.c_str());
        return OPB200_ERR_IO;
    }
    const opb200_status s = planNet(net, wm);
    if (s != OPB200_OK)
        return s;
    std::string out;
    char line[1024];
    static const char* kinds[] = {"patch", "conv", "pool", "convert", "conv"};
    for (const auto& op : net.ops)
    {
        std::string segs;
        for (const View& v : op.in)
        {
            std::snprintf(line, sizeof(line), "%s%d:%d:%d", segs.empty() ? "" : ",", v.buf, v.cOff, v.C);
            segs += line;
        }
        const View& ov = op.poolFused ? op.poolOut : op.out;
        std::snprintf(line, sizeof(line), "%s %s cin=%d cout=%d k=%d bn=%d tilesN=%d segs=[%s] out=%d:%d:%d out32=%d bf16=%d pool=%d\n",
                      kinds[op.kind], op.name.c_str(), op.cin, op.cout, op.k, op.bn, op.nTilesN, segs.c_str(),
                      ov.buf, ov.cOff, ov.C, op.out32Off, op.storeBf16 ? 1 : 0, op.poolFused ? 1 : 0);
        out += line;
    }
    std::snprintf(line, sizeof(line), "output %s C=%d buffers=%d levels=%d\n", net.lastBlobName.c_str(), net.outC,
                  (int)net.buffers.size(), (int)net.levels.size());
    out += line;
    if (out.size() + 1 > capacity)
        OPB_FAIL(OPB200_ERR_INVALID, "net_describe: text buffer too small");
    std::memcpy(text, out.c_str(), out.size() + 1);
    return OPB200_OK;
}
