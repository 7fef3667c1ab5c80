// TEST INFRASTRUCTURE (oracle) -- never linked into the product.
//
// Plain-C entry surface (refcuda_*) over the reference's own, UNMODIFIED CUDA sources
//   /root/reference/src/openpose/net/resizeAndMergeBase.cu   (resizeAndMergeGpu, resizeAndPadRbgGpu)
//   /root/reference/src/openpose/net/nmsBase.cu              (nmsGpu)
//   /root/reference/src/openpose/net/maximumBase.cu          (maximumGpu)
//   /root/reference/src/openpose/net/bodyPartConnectorBase.cu (connectBodyPartsGpu: pafScoreKernel + host assembly)
//   /root/reference/src/openpose/gpu/cuda.cu, cuda.cpp        (reorderAndNormalize, cudaCheck)
// compiled by oracle/Makefile (target `refcuda`) with nvcc's defaults (-fmad=true, as the reference's CUDA build has
// them) into oracle/_ref/libopref_cuda.so. The GPU tests use it as the parity oracle for the CUDA-build semantics and
// bench.py times it on the same box as the GPU-vs-GPU baseline of the post-processing stages.
// Every data pointer is a DEVICE pointer unless the name says host. The reference launches on the legacy default
// stream; nothing here synchronises except where the reference itself does (cudaMemcpy, thrust).
// Nothing here restates reference arithmetic.
#include <array>
#include <cstdio>
#include <cstring>
#include <vector>
#include <cuda_runtime.h>
#include <openpose/core/array.hpp>
#include <openpose/core/point.hpp>
#include <openpose/gpu/cuda.hpp>
#include <openpose/net/bodyPartConnectorBase.hpp>
#include <openpose/net/maximumBase.hpp>
#include <openpose/net/nmsBase.hpp>
#include <openpose/net/resizeAndMergeBase.hpp>
#include <openpose/pose/poseParameters.hpp>

namespace
{
    thread_local char gErr[2048] = "";
    int fail(const std::exception& e)
    {
        std::snprintf(gErr, sizeof(gErr), "%s", e.what());
        cudaGetLastError();
        return 1;
    }
    std::array<int, 4> arr4(const int* s) { return {{s[0], s[1], s[2], s[3]}}; }
}

extern "C"
{
    const char* refcuda_last_error() { return gErr; }

    // op::resizeAndMergeGpu<float> (resizeAndMergeBase.cu:275-528). srcs: host array of device pointers.
    int refcuda_resize_merge(float* target, const float* const* srcs, const int* srcShapes, const float* scaleRatios,
                             int nScales, const int* targetShape)
    {
        try
        {
            std::vector<const float*> sourcePtrs(srcs, srcs + nScales);
            std::vector<std::array<int, 4>> sourceSizes;
            for (int i = 0; i < nScales; i++)
                sourceSizes.push_back(arr4(srcShapes + 4 * i));
            const std::vector<float> ratios(scaleRatios, scaleRatios + nScales);
            op::resizeAndMergeGpu(target, sourcePtrs, arr4(targetShape), sourceSizes, ratios);
            return 0;
        }
        catch (const std::exception& e) { return fail(e); }
    }

    // op::nmsGpu<float> (nmsBase.cu:249-351). kernel: int map of the source's full shape (nmsCaffe.cpp:123).
    int refcuda_nms(float* target, int* kernel, const float* source, float threshold, const int* targetShape,
                    const int* sourceShape, float offsetX, float offsetY)
    {
        try
        {
            op::nmsGpu(target, kernel, source, threshold, arr4(targetShape), arr4(sourceShape),
                       op::Point<float>{offsetX, offsetY});
            return 0;
        }
        catch (const std::exception& e) { return fail(e); }
    }

    // op::maximumGpu<float> (maximumBase.cu:63-120).
    int refcuda_maximum(float* target, const float* source, const int* targetShape, const int* sourceShape)
    {
        try
        {
            op::maximumGpu(target, source, arr4(targetShape), arr4(sourceShape));
            return 0;
        }
        catch (const std::exception& e) { return fail(e); }
    }

    // op::connectBodyPartsGpu<float> (bodyPartConnectorBase.cu:148-250), fed as BodyPartConnectorCaffe::Forward_gpu
    // feeds it (bodyPartConnectorCaffe.cpp:332-404): peaks on the host AND the device, the two index tables on the
    // device, pair-score cube (pairs, maxPeaks, maxPeaks) host + device.
    // pairScoresHostOut (optional) receives the cube the kernel produced; keypointsOut / scoresOut have room for
    // `capacity` people; *people = the number found (may exceed capacity, then only `capacity` are copied).
    int refcuda_connect(const float* heatMapGpu, const float* peaksHost, const float* peaksGpu, int poseModel,
                        int width, int height, int maxPeaks, float interMinAboveThreshold, float interThreshold,
                        int minSubsetCnt, float minSubsetScore, float defaultNmsThreshold, float scaleFactor,
                        int maximizePositives, float* pairScoresGpu, const unsigned int* bodyPartPairsGpu,
                        const unsigned int* mapIdxGpu, float* pairScoresHostOut, float* keypointsOut,
                        float* scoresOut, int capacity, int* people)
    {
        try
        {
            const auto model = (op::PoseModel)poseModel;
            const int pairs = (int)(op::getPosePartPairs(model).size() / 2);
            const int parts = (int)op::getPoseNumberBodyParts(model);
            op::Array<float> pairScoresCpu({pairs, maxPeaks, maxPeaks});
            op::Array<float> poseKeypoints, poseScores;
            op::connectBodyPartsGpu(poseKeypoints, poseScores, heatMapGpu, peaksHost, model,
                                    op::Point<int>{width, height}, maxPeaks, interMinAboveThreshold, interThreshold,
                                    minSubsetCnt, minSubsetScore, defaultNmsThreshold, scaleFactor,
                                    maximizePositives != 0, pairScoresCpu, pairScoresGpu, bodyPartPairsGpu,
                                    mapIdxGpu, peaksGpu);
            if (pairScoresHostOut)
                std::memcpy(pairScoresHostOut, pairScoresCpu.getConstPtr(), sizeof(float) * pairScoresCpu.getVolume());
            const int n = poseKeypoints.empty() ? 0 : poseKeypoints.getSize(0);
            *people = n;
            const int m = n < capacity ? n : capacity;
            if (m > 0)
            {
                std::memcpy(keypointsOut, poseKeypoints.getConstPtr(), sizeof(float) * (size_t)m * parts * 3);
                std::memcpy(scoresOut, poseScores.getConstPtr(), sizeof(float) * (size_t)m);
            }
            return 0;
        }
        catch (const std::exception& e) { return fail(e); }
    }

    // op::reorderAndNormalize<float> (gpu/cuda.cu:11-25,36-60): u8 HWC -> float CHW, x/256 - 0.5.
    int refcuda_reorder_and_normalize(float* target, const unsigned char* src, int width, int height, int channels)
    {
        try
        {
            op::reorderAndNormalize(target, src, width, height, channels);
            return 0;
        }
        catch (const std::exception& e) { return fail(e); }
    }

    // op::resizeAndPadRbgGpu<float> (resizeAndMergeBase.cu:56-78,531-551): float CHW planes -> net input.
    int refcuda_resize_and_pad_rbg(float* target, const float* src, int widthSource, int heightSource,
                                   int widthTarget, int heightTarget, float scaleFactor)
    {
        try
        {
            op::resizeAndPadRbgGpu(target, src, widthSource, heightSource, widthTarget, heightTarget, scaleFactor);
            op::cudaCheck(__LINE__, __FUNCTION__, __FILE__);
            return 0;
        }
        catch (const std::exception& e) { return fail(e); }
    }
}
