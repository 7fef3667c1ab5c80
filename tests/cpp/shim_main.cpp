// GPU test driver for the reference-side bindings in integration/: built against the REFERENCE's own headers
// (oracle/Makefile target `shim`, needs /root/reference) and run on the GPU box by tests/test_shim_gpu.py.
//   shim_test <prototxt> <caffemodel> <input.bin> <N> <H> <W> <out_prefix>
// Runs op::NetB200 -> op::resizeAndMergeGpu -> op::nmsGpu -> op::connectBodyPartsGpu exactly as
// PoseExtractorCaffe::forwardPass orders them (src/openpose/pose/poseExtractorCaffe.cpp:200-634), using
// op::Array / op::ArrayCpuGpu for every blob, and dumps net_output, peaks, keypoints and scores of frame 0.
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <vector>
#include <openpose/core/arrayCpuGpu.hpp>
#include <openpose/net/bodyPartConnectorBase.hpp>
#include <openpose/net/nmsBase.hpp>
#include <openpose/net/resizeAndMergeBase.hpp>
#include <openpose/pose/poseParameters.hpp>
#include "../../integration/netB200.hpp"

static void dump(const std::string& path, const float* p, size_t n)
{
    std::ofstream f(path, std::ios::binary);
    f.write(reinterpret_cast<const char*>(p), (std::streamsize)(n * sizeof(float)));
}

int main(int argc, char** argv)
{
    if (argc != 8)
    {
        std::fprintf(stderr, "usage: %s prototxt caffemodel input.bin N H W out_prefix\n", argv[0]);
        return 2;
    }
    try
    {
        const int N = std::atoi(argv[4]), H = std::atoi(argv[5]), W = std::atoi(argv[6]);
        const std::string prefix = argv[7];
        op::Array<float> input({N, 3, H, W});
        {
            std::ifstream f(argv[3], std::ios::binary);
            f.read(reinterpret_cast<char*>(input.getPtr()), (std::streamsize)(input.getVolume() * sizeof(float)));
            if (!f)
                op::error("cannot read the input file", __LINE__, __FUNCTION__, __FILE__);
        }
        // error path first: NetCaffe's message for a wrong blob name must come through op::error
        bool threw = false;
        try
        {
            op::NetB200 bad{argv[1], argv[2], 0, false, "no_such_blob"};
            bad.initializationOnThread();
        }
        catch (const std::exception& e)
        {
            threw = std::string(e.what()).find("The output blob is a nullptr") != std::string::npos;
        }
        if (!threw)
        {
            std::fprintf(stderr, "expected op::error for an unknown output blob\n");
            return 3;
        }
        std::shared_ptr<op::Net> net = std::make_shared<op::NetB200>(argv[1], argv[2], 0, false);
        net->initializationOnThread();
        net->forwardPass(input);
        const auto netOut = net->getOutputBlobArray();                    // (N, 78, H/8, W/8)
        dump(prefix + "_net.bin", netOut->cpu_data(), (size_t)netOut->count());
        std::printf("net_output %d %d %d %d\n", netOut->shape(0), netOut->shape(1), netOut->shape(2), netOut->shape(3));

        const auto poseModel = op::PoseModel::BODY_25;
        const int parts = (int)op::getPoseNumberBodyParts(poseModel);
        const int maxPeaks = (int)op::getPoseMaxPeaks();
        // frame 0 only: the reference's layers merge batch 1 (resizeAndMergeCaffe.cpp:75-82)
        op::ArrayCpuGpu<float> heat{1, netOut->shape(1), H, W};
        op::ArrayCpuGpu<float> peaks{1, parts, maxPeaks + 1, 3};
        const std::array<int, 4> heatSize{1, netOut->shape(1), H, W};
        const std::array<int, 4> srcSize{1, netOut->shape(1), netOut->shape(2), netOut->shape(3)};
        op::resizeAndMergeGpu(heat.mutable_gpu_data(), std::vector<const float*>{netOut->gpu_data()}, heatSize,
                              std::vector<std::array<int, 4>>{srcSize}, std::vector<float>{1.f});
        const float scaleNetToOutput = 1.95913f;
        const float offset = float(0.5 / double(scaleNetToOutput));
        op::nmsGpu(peaks.mutable_gpu_data(), (int*)nullptr, heat.gpu_data(), 0.05f,
                   std::array<int, 4>{1, parts, maxPeaks + 1, 3}, heatSize, op::Point<float>{offset, offset});
        dump(prefix + "_peaks.bin", peaks.cpu_data(), (size_t)peaks.count());
        op::Array<float> poseKeypoints, poseScores;
        op::connectBodyPartsGpu(poseKeypoints, poseScores, heat.gpu_data(), peaks.cpu_data(), poseModel,
                                op::Point<int>{W, H}, maxPeaks, 0.95f, 0.05f, 3, 0.4f, 0.05f, scaleNetToOutput, false,
                                op::Array<float>{}, (float*)nullptr, (const unsigned int*)nullptr,
                                (const unsigned int*)nullptr, peaks.gpu_data());
        const int people = poseKeypoints.empty() ? 0 : poseKeypoints.getSize(0);
        std::printf("people %d\n", people);
        if (people > 0)
        {
            dump(prefix + "_kp.bin", poseKeypoints.getConstPtr(), poseKeypoints.getVolume());
            dump(prefix + "_sc.bin", poseScores.getConstPtr(), poseScores.getVolume());
        }
        return 0;
    }
    catch (const std::exception& e)
    {
        std::fprintf(stderr, "shim_test failed: %s\n", e.what());
        return 1;
    }
}
