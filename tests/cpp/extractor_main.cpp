// Runs the reference's UNMODIFIED body extractor -- src/openpose/pose/poseExtractorCaffe.cpp, pose/poseExtractorNet.cpp,
// net/{resizeAndMerge,nms,bodyPartConnector,maximum}Caffe.cpp, compiled from /root/reference with -DUSE_CAFFE -DUSE_CUDA
// (oracle/Makefile, target `extractor`) -- over the B200 backend: integration/netCaffeB200.cpp (op::NetCaffe),
// integration/postB200.cpp (the four *Gpu functions), integration/arrayCpuGpuB200.cpp (op::ArrayCpuGpu) + libopb200.so.
// tests/test_extractor_cpp_gpu.py writes the request file, runs this binary and checks every output.
//
// usage: extractor_test <request.bin> <out prefix>
// request.bin (little endian): int32 poseModel, enableNet, nScales, inputW, inputH, heatMapTypesMask, heatMapScaleMode,
//   addPartCandidates, maximizePositives, repeats; then 3 strings (int32 length + bytes): modelFolder, protoTxtPath,
//   caffeModelPath; per scale: float64 scaleInputToNetInput, int32 n,c,h,w, float32 data; if !enableNet: int32 n,c,h,w +
//   float32 poseNetOutput.
// outputs <prefix>.<name>.bin: int32 ndims, dims..., float32 data -- keypoints, scores, heatblob (the resized heat-map
//   blob via getHeatMapCpuConstPtr), peaks (getCandidatesCpuConstPtr), heatmaps (getHeatMapsCopy), candidates
//   (getCandidatesCopy, flattened as per-part counts + triplets), scale (getScaleNetToOutput).
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <stdexcept>
#include <string>
#include <vector>
#include <openpose/core/array.hpp>
#include <openpose/pose/poseExtractorCaffe.hpp>
#include <openpose/pose/poseParameters.hpp>

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
    std::string getString(std::ifstream& in)
    {
        const auto n = get<int32_t>(in);
        std::string s((size_t)n, '\0');
        if (n > 0)
            in.read(&s[0], n);
        return s;
    }
    op::Array<float> getArray(std::ifstream& in)
    {
        std::vector<int> dims(4);
        for (auto& d : dims)
            d = get<int32_t>(in);
        op::Array<float> a(dims);
        in.read(reinterpret_cast<char*>(a.getPtr()), sizeof(float) * a.getVolume());
        if (!in)
            throw std::runtime_error("request file truncated");
        return a;
    }
    void put(const std::string& path, const std::vector<int>& dims, const float* data)
    {
        std::ofstream out(path, std::ios::binary);
        const int32_t nd = (int32_t)dims.size();
        out.write(reinterpret_cast<const char*>(&nd), 4);
        size_t volume = 1;
        for (const auto d : dims)
        {
            const int32_t v = d;
            out.write(reinterpret_cast<const char*>(&v), 4);
            volume *= (size_t)d;
        }
        if (volume > 0 && data != nullptr)
            out.write(reinterpret_cast<const char*>(data), sizeof(float) * volume);
    }
    void putArray(const std::string& path, const op::Array<float>& a)
    {
        put(path, a.empty() ? std::vector<int>{0} : a.getSize(), a.empty() ? nullptr : a.getConstPtr());
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
        const auto poseModel = (op::PoseModel)get<int32_t>(in);
        const bool enableNet = get<int32_t>(in) != 0;
        const int nScales = get<int32_t>(in);
        const op::Point<int> inputDataSize{get<int32_t>(in), 0};
        const int inputH = get<int32_t>(in);
        const int typesMask = get<int32_t>(in);
        const auto scaleMode = (op::ScaleMode)get<int32_t>(in);
        const bool addPartCandidates = get<int32_t>(in) != 0;
        const bool maximizePositives = get<int32_t>(in) != 0;
        const int repeats = get<int32_t>(in);
        const auto modelFolder = getString(in);
        const auto protoTxtPath = getString(in);
        const auto caffeModelPath = getString(in);
        std::vector<op::HeatMapType> heatMapTypes;
        if (typesMask & 1) heatMapTypes.push_back(op::HeatMapType::Parts);
        if (typesMask & 2) heatMapTypes.push_back(op::HeatMapType::Background);
        if (typesMask & 4) heatMapTypes.push_back(op::HeatMapType::PAFs);
        std::vector<op::Array<float>> inputNetData;
        std::vector<double> scaleInputToNetInputs;
        for (int i = 0; i < nScales; i++)
        {
            scaleInputToNetInputs.push_back(get<double>(in));
            inputNetData.push_back(getArray(in));
        }
        const op::Array<float> poseNetOutput = enableNet ? op::Array<float>{} : getArray(in);

        // The reference's own class, as WPoseExtractor / the tutorials drive it
        op::PoseExtractorCaffe extractor{poseModel, modelFolder, 0, heatMapTypes, scaleMode, addPartCandidates,
                                         maximizePositives, protoTxtPath, caffeModelPath, 0.f, enableNet, false};
        extractor.initializationOnThread();
        for (int r = 0; r < repeats; r++)
            extractor.forwardPass(inputNetData, op::Point<int>{inputDataSize.x, inputH}, scaleInputToNetInputs,
                                  poseNetOutput);

        putArray(prefix + ".keypoints.bin", extractor.getPoseKeypoints());
        putArray(prefix + ".scores.bin", extractor.getPoseScores());
        const auto heatSize = extractor.getHeatMapSize();
        put(prefix + ".heatblob.bin", heatSize, extractor.getHeatMapCpuConstPtr());
        const int parts = (int)op::getPoseNumberBodyParts(poseModel);
        put(prefix + ".peaks.bin", {parts, (int)op::getPoseMaxPeaks() + 1, 3}, extractor.getCandidatesCpuConstPtr());
        putArray(prefix + ".heatmaps.bin", extractor.getHeatMapsCopy());
        const auto candidates = extractor.getCandidatesCopy();
        std::vector<float> flat;
        flat.push_back((float)candidates.size());
        for (const auto& part : candidates)
        {
            flat.push_back((float)part.size());
            for (const auto& c : part)
                flat.insert(flat.end(), c.begin(), c.end());
        }
        put(prefix + ".candidates.bin", {(int)flat.size()}, flat.data());
        const float scale = extractor.getScaleNetToOutput();
        put(prefix + ".scale.bin", {1}, &scale);
        std::printf("extractor ok: %d people\n", extractor.getPoseKeypoints().empty() ? 0
                                                                                      : extractor.getPoseKeypoints().getSize(0));
        return 0;
    }
    catch (const std::exception& e)
    {
        std::fprintf(stderr, "extractor error: %s\n", e.what());
        return 1;
    }
}
