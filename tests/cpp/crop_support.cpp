// TEST HARNESS ONLY -- link support for tests/cpp/crop_main.cpp (the reference's hand / face extractors over the B200
// backend) in an image without OpenCV C++: op::uCharCvMatToFloatPtr restated from src/openpose/utilities/openCv.cpp:57-130
// (HWC u8 -> CHW float, then v / 256 - 0.5 for normalize == 1: its scalar branch; the AVX branch computes the same FMA,
// exact for 8-bit v). A real OpenPose build links the reference's own openCv.cpp.
#include <opencv2/core/core.hpp>
#include <openpose/core/common.hpp>
#include <openpose/core/matrix.hpp>

namespace op
{
    void uCharCvMatToFloatPtr(float* floatPtrImage, const Matrix& matImage, const int normalize)
    {
        const cv::Mat cvImage = OP_OP2CVCONSTMAT(matImage);
        const int width = cvImage.cols, height = cvImage.rows, channels = cvImage.channels();
        const auto* const originFramePtr = cvImage.data;
        for (auto c = 0; c < channels; c++)
            for (auto y = 0; y < height; y++)
                for (auto x = 0; x < width; x++)
                    floatPtrImage[(c * height + y) * width + x] = float(originFramePtr[(y * width + x) * channels + c]);
        if (normalize == 1)
        {
            const auto ratio = 1.f / 256.f;
            for (auto pixel = 0; pixel < width * height * channels; ++pixel)
                floatPtrImage[pixel] = floatPtrImage[pixel] * ratio - 0.5f;
        }
        else if (normalize != 0)
            error("Harness: only normalize = 0 / 1 are restated.", __LINE__, __FUNCTION__, __FILE__);
    }
}
