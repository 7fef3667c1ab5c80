// Heat-map export: the channel-group selection and range remap of op::PoseExtractorNet::getHeatMapsCopy
// (reference: src/openpose/pose/poseExtractorNet.cpp:106-243) as one kernel. The reference copies each selected group
// device -> host with a blocking cudaMemcpy and then walks it with scalar host loops (fastTruncate, * and +,
// positiveIntRound); here one launch reads the selected channels once and writes the remapped values once -- into
// device memory or straight into mapped pinned host memory (UVA), so an export is a single pass over the data with
// no staging copy. Algorithmic bytes: 8 * selectedChannels * H * W per frame (4 read + 4 written).
#include "../common.cuh"

namespace opb
{
    constexpr int kExpThreads = 256;
    // op::ScaleMode (include/openpose/core/enumClasses.hpp:6-17)
    enum : int
    {
        kZeroToOne = 3,
        kZeroToOneFixedAspect = 4,
        kPlusMinusOne = 5,
        kPlusMinusOneFixedAspect = 6,
        kUnsignedChar = 7,
        kNoScale = 8
    };

    // fastTruncate (include/openpose/utilities/fastMath.hpp:72-88): the reference's compare-and-select, so a NaN
    // passes through as it does there (fminf / fmaxf would drop it).
    __device__ __forceinline__ float truncate(float v, float lo)
    {
        const float a = (lo > v ? lo : v);
        return (1.f < a ? 1.f : a);
    }
    // positiveIntRound = (int)(a + 0.5f). The operand is clamped, so only a NaN is out of range; the host's
    // float -> int conversion (x86 cvttss2si) then yields INT_MIN where the device's cvt would yield 0.
    __device__ __forceinline__ float roundPos(float a)
    {
        const float r = add(a, 0.5f);
        return r != r ? -2147483648.f : (float)(int)r;
    }

    // kind 0: body parts / background (range [0,1]); kind 1: PAFs (range [-1,1]). poseExtractorNet.cpp:139-151,
    // 213-227.
    template <int KIND>
    __device__ __forceinline__ float remap(float v, int mode)
    {
        if (mode == kNoScale)
            return v;
        if (KIND == 0)
        {
            const float t = truncate(v, 0.f);
            if (mode == kPlusMinusOne || mode == kPlusMinusOneFixedAspect)
                return sub(mul(t, 2.f), 1.f);
            if (mode == kUnsignedChar)
                return roundPos(mul(t, 255.f));
            return t;
        }
        const float t = truncate(v, -1.f);
        if (mode == kZeroToOne || mode == kZeroToOneFixedAspect)
            return add(mul(t, 0.5f), 0.5f);
        if (mode == kUnsignedChar)
            return roundPos(add(mul(t, 128.5f), 128.5f));
        return t;
    }

    struct ExportArgs
    {
        float* dst;
        const float* heat;
        int hw, heatChannels, outChannels;
        int partsOut, bkgOut;       // how many selected channels come from each leading group
        int bkgSrc, pafSrc;         // first source channel of the background / PAF groups
        int mode;
    };

    template <bool VEC>
    __global__ void __launch_bounds__(kExpThreads) exportKernel(const ExportArgs a)
    {
        const int c = blockIdx.y, f = blockIdx.z;
        // selected channel -> source channel and value range
        int srcC, kind = 0;
        if (c < a.partsOut)
            srcC = c;
        else if (c < a.partsOut + a.bkgOut)
            srcC = a.bkgSrc + (c - a.partsOut);
        else
        {
            srcC = a.pafSrc + (c - a.partsOut - a.bkgOut);
            kind = 1;
        }
        const float* s = a.heat + ((size_t)f * a.heatChannels + srcC) * a.hw;
        float* d = a.dst + ((size_t)f * a.outChannels + c) * a.hw;
        const int i = blockIdx.x * kExpThreads + threadIdx.x;
        if (VEC)
        {
            if (i * 4 >= a.hw)
                return;
            float4 v = __ldcs(reinterpret_cast<const float4*>(s) + i);
            if (kind == 0)
                v = make_float4(remap<0>(v.x, a.mode), remap<0>(v.y, a.mode), remap<0>(v.z, a.mode),
                                remap<0>(v.w, a.mode));
            else
                v = make_float4(remap<1>(v.x, a.mode), remap<1>(v.y, a.mode), remap<1>(v.z, a.mode),
                                remap<1>(v.w, a.mode));
            __stcs(reinterpret_cast<float4*>(d) + i, v);
        }
        else
        {
            if (i >= a.hw)
                return;
            d[i] = kind == 0 ? remap<0>(s[i], a.mode) : remap<1>(s[i], a.mode);
        }
    }
}

extern "C" int opb200_heatmaps_export_channels(int parts, int has_bkg, int paf_channels, int types_mask)
{
    return ((types_mask & OPB200_HEATMAP_PARTS) ? parts : 0) + ((types_mask & OPB200_HEATMAP_BACKGROUND) ? (has_bkg ? 1 : 0) : 0) +
           ((types_mask & OPB200_HEATMAP_PAFS) ? paf_channels : 0);
}

extern "C" opb200_status opb200_heatmaps_export(float* dst, const float* heat, const int heat_shape[4], int parts,
                                                int has_bkg, int paf_channels, int types_mask, int scale_mode,
                                                void* stream)
{
    using namespace opb;
    OPB_REQUIRE(heat && heat_shape, "heatmaps_export: null argument");
    const int n = heat_shape[0], ch = heat_shape[1], h = heat_shape[2], w = heat_shape[3];
    OPB_REQUIRE(n > 0 && h > 0 && w > 0 && parts > 0 && paf_channels >= 0, "heatmaps_export: empty heat-maps");
    OPB_REQUIRE(parts + (has_bkg ? 1 : 0) + paf_channels <= ch, "heatmaps_export: heat-maps have fewer channels than the model");
    // any other op::ScaleMode value takes the reference's final else branch (plain fastTruncate)
    OPB_REQUIRE(scale_mode >= 0 && scale_mode <= kNoScale, "heatmaps_export: not an op::ScaleMode value");
    // poseExtractorNet.cpp:196-200
    OPB_REQUIRE(!(types_mask & OPB200_HEATMAP_BACKGROUND) || has_bkg,
                "You enabled `--heatmaps_add_bkg` for a model that does not contain one. Please, remove this flag for "
                "this model.");
    ExportArgs a;
    a.dst = dst;
    a.heat = heat;
    a.hw = h * w;
    a.heatChannels = ch;
    a.partsOut = (types_mask & OPB200_HEATMAP_PARTS) ? parts : 0;
    a.bkgOut = (types_mask & OPB200_HEATMAP_BACKGROUND) ? 1 : 0;
    a.bkgSrc = parts;
    a.pafSrc = parts + (has_bkg ? 1 : 0);
    a.outChannels = opb200_heatmaps_export_channels(parts, has_bkg, paf_channels, types_mask);
    a.mode = scale_mode;
    if (a.outChannels == 0)
        return OPB200_OK; // no heat-map type selected: the reference returns an empty Array
    OPB_REQUIRE(dst, "heatmaps_export: null destination");
    OPB_REQUIRE(n <= 65535 && a.outChannels <= 65535, "heatmaps_export: too many frames / channels for one launch");
    const bool vec = (a.hw % 4 == 0) && ((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(heat)) % 16 == 0);
    const int items = vec ? a.hw / 4 : a.hw;
    const dim3 grid((items + kExpThreads - 1) / kExpThreads, a.outChannels, n);
    if (vec)
        exportKernel<true><<<grid, kExpThreads, 0, static_cast<cudaStream_t>(stream)>>>(a);
    else
        exportKernel<false><<<grid, kExpThreads, 0, static_cast<cudaStream_t>(stream)>>>(a);
    OPB_LAUNCH_CHECK();
    return OPB200_OK;
}
