// TEST HARNESS ONLY -- link support for tests/cpp/extractor_main.cpp in an image without OpenCV C++.
// A real OpenPose build links the reference's own utilities/openCv.cpp, openCvPrivate.cpp, net/resizeAndMergeBase.cpp,
// net/maximumBase.cpp (all need OpenCV); here only the symbols the unmodified extractor objects reference are supplied:
//   * op::resizeGetScaleFactor -- live on the path (poseExtractorCaffe.cpp:306-310): restated from
//     src/openpose/utilities/openCv.cpp:182-195 (two divisions and a min);
//   * resizeFixedAspectRatio / keepRoiInside / uCharCvMatToFloatPtr -- only named by the TOP_DOWN_REFINEMENT branch
//     (a compile-time false) and by the hand/face croppers: abort if ever reached;
//   * resizeAndMergeCpu / maximumCpu -- the layers' Forward_cpu bodies, not reached with a GPU: report an error.
#include <array>
#include <cstdlib>
#include <vector>
#include <opencv2/core/core.hpp>
#include <openpose/core/common.hpp>
#include <openpose/net/maximumBase.hpp>
#include <openpose/net/resizeAndMergeBase.hpp>
#include <openpose/utilities/fastMath.hpp>

namespace op
{
    double resizeGetScaleFactor(const Point<int>& initialSize, const Point<int>& targetSize)
    {
        const auto ratioWidth = (targetSize.x - 1) / (double)(initialSize.x - 1);
        const auto ratioHeight = (targetSize.y - 1) / (double)(initialSize.y - 1);
        return fastMin(ratioWidth, ratioHeight);
    }

    void keepRoiInside(Rectangle<int>&, const int, const int) { std::abort(); }

    void resizeFixedAspectRatio(cv::Mat&, const cv::Mat&, const double, const Point<int>&, const int,
                                const cv::Scalar&)
    {
        std::abort();
    }

    template <typename T>
    void resizeAndMergeCpu(T*, const std::vector<const T*>&, const std::array<int, 4>&,
                           const std::vector<std::array<int, 4>>&, const std::vector<T>&)
    {
        error("resizeAndMergeCpu is not part of this harness (GPU path only).", __LINE__, __FUNCTION__, __FILE__);
    }
    template <typename T>
    void maximumCpu(T*, const T* const, const std::array<int, 4>&, const std::array<int, 4>&)
    {
        error("maximumCpu is not part of this harness (GPU path only).", __LINE__, __FUNCTION__, __FILE__);
    }
    template void resizeAndMergeCpu(float*, const std::vector<const float*>&, const std::array<int, 4>&,
                                    const std::vector<std::array<int, 4>>&, const std::vector<float>&);
    template void resizeAndMergeCpu(double*, const std::vector<const double*>&, const std::array<int, 4>&,
                                    const std::vector<std::array<int, 4>>&, const std::vector<double>&);
    template void maximumCpu(float*, const float* const, const std::array<int, 4>&, const std::array<int, 4>&);
    template void maximumCpu(double*, const double* const, const std::array<int, 4>&, const std::array<int, 4>&);
}
