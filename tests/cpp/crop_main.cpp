// Runs the reference's UNMODIFIED hand and face extractors -- src/openpose/hand/handExtractorCaffe.cpp, handExtractorNet.cpp,
// face/faceExtractorCaffe.cpp, faceExtractorNet.cpp, core/matrix.cpp and the *Caffe layer classes, compiled from
// /root/reference with -DUSE_CAFFE -DUSE_CUDA (oracle/Makefile, target `cropextractor`) -- over the B200 backend
// (integration/netCaffeB200.cpp, postB200.cpp, arrayCpuGpuB200.cpp, libopb200.so). cv::Mat / cv::warpAffine come from the
// functional stand-in tests/cpp/standin_cv (warp = the cv2-pinned oracle restatement).
//
// usage: crop_test <request.bin> <out prefix>
// request.bin: int32 kind (0 hand, 1 face), people, numberScales, heatMapTypesMask, heatMapScaleMode, frameH, frameW;
//   float32 rangeScales; string modelFolder (int32 length + bytes); u8 frame (H, W, 3) BGR; float32 rectangles
//   (hand: people x 2 x 4, face: people x 4) as x, y, width, height.
// outputs <prefix>.<name>.bin (int32 ndims, dims, float32 data): hand -> left, right (people, 21, 3) [+ heatleft,
//   heatright]; face -> face (people, 70, 3) [+ heatface].
#include <array>
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>
#include <opencv2/core/core.hpp>
#include <openpose/core/array.hpp>
#include <openpose/core/matrix.hpp>
#include <openpose/face/faceExtractorCaffe.hpp>
#include <openpose/hand/handExtractorCaffe.hpp>

namespace
{
    template <typename T>
    T get(std::ifstream& in)
    {
        T v;
        in.read(reinterpret_cast<char*>(&v), sizeof(T));
        if (!in)
            throw std::runtime_error("request file truncated");
        return v;
    }
    void putArray(const std::string& path, const op::Array<float>& a)
    {
        std::ofstream out(path, std::ios::binary);
        const std::vector<int> dims = a.empty() ? std::vector<int>{0} : a.getSize();
        const int32_t nd = (int32_t)dims.size();
        out.write(reinterpret_cast<const char*>(&nd), 4);
        for (const auto d : dims)
        {
            const int32_t v = d;
            out.write(reinterpret_cast<const char*>(&v), 4);
        }
        if (!a.empty())
            out.write(reinterpret_cast<const char*>(a.getConstPtr()), sizeof(float) * a.getVolume());
    }
}

int main(int argc, char** argv)
{
    if (argc != 3)
    {
        std::fprintf(stderr, "usage: %s <request.bin> <out prefix>\n", argv[0]);
        return 2;
    }
    try
    {
        std::ifstream in(argv[1], std::ios::binary);
        if (!in)
            throw std::runtime_error("cannot open request file");
        const std::string prefix = argv[2];
        const int kind = get<int32_t>(in), people = get<int32_t>(in), numberScales = get<int32_t>(in);
        const int typesMask = get<int32_t>(in);
        const auto scaleMode = (op::ScaleMode)get<int32_t>(in);
        const int frameH = get<int32_t>(in), frameW = get<int32_t>(in);
        const float rangeScales = get<float>(in);
        const auto len = get<int32_t>(in);
        std::string modelFolder((size_t)len, '\0');
        in.read(&modelFolder[0], len);
        cv::Mat frame(frameH, frameW, CV_8UC3);
        in.read(reinterpret_cast<char*>(frame.data), (size_t)frameH * frameW * 3);
        std::vector<float> rects((size_t)people * (kind == 0 ? 8 : 4));
        in.read(reinterpret_cast<char*>(rects.data()), sizeof(float) * rects.size());
        if (!in)
            throw std::runtime_error("request file truncated");
        std::vector<op::HeatMapType> heatMapTypes;
        if (typesMask & 1) heatMapTypes.push_back(op::HeatMapType::Parts);
        if (typesMask & 2) heatMapTypes.push_back(op::HeatMapType::Background);
        if (typesMask & 4) heatMapTypes.push_back(op::HeatMapType::PAFs);
        const op::Matrix inputData = OP_CV2OPMAT(frame);
        const op::Point<int> netSize{368, 368};
        if (kind == 0)
        {
            std::vector<std::array<op::Rectangle<float>, 2>> handRectangles((size_t)people);
            for (int p = 0; p < people; p++)
                for (int h = 0; h < 2; h++)
                {
                    const float* r = &rects[(size_t)(p * 2 + h) * 4];
                    handRectangles[p][h] = op::Rectangle<float>{r[0], r[1], r[2], r[3]};
                }
            op::HandExtractorCaffe extractor{netSize, netSize, modelFolder, 0, numberScales, rangeScales, heatMapTypes,
                                             scaleMode, false};
            extractor.initializationOnThread();
            extractor.forwardPass(handRectangles, inputData);
            const auto keypoints = extractor.getHandKeypoints();
            putArray(prefix + ".left.bin", keypoints[0]);
            putArray(prefix + ".right.bin", keypoints[1]);
            const auto heatMaps = extractor.getHeatMaps();
            putArray(prefix + ".heatleft.bin", heatMaps[0]);
            putArray(prefix + ".heatright.bin", heatMaps[1]);
        }
        else
        {
            std::vector<op::Rectangle<float>> faceRectangles((size_t)people);
            for (int p = 0; p < people; p++)
                faceRectangles[p] = op::Rectangle<float>{rects[p * 4], rects[p * 4 + 1], rects[p * 4 + 2], rects[p * 4 + 3]};
            op::FaceExtractorCaffe extractor{netSize, netSize, modelFolder, 0, heatMapTypes, scaleMode, false};
            extractor.initializationOnThread();
            extractor.forwardPass(faceRectangles, inputData);
            putArray(prefix + ".face.bin", extractor.getFaceKeypoints());
            putArray(prefix + ".heatface.bin", extractor.getHeatMaps());
        }
        std::printf("crop extractor ok: kind %d, %d people\n", kind, people);
        return 0;
    }
    catch (const std::exception& e)
    {
        std::fprintf(stderr, "crop extractor error: %s\n", e.what());
        return 1;
    }
}
