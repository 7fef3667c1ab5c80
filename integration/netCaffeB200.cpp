// op::NetCaffe implemented over the B200 backend: compile this file INSTEAD of the reference's
// src/openpose/net/netCaffe.cpp and every construction site of NetCaffe (pose/poseExtractorCaffe.cpp:82-86,
// hand/handExtractorCaffe.cpp:34-35, face/faceExtractorCaffe.cpp:32) runs on libopb200.so with NO source change --
// the class declaration stays the reference's own include/openpose/net/netCaffe.hpp:9-32 (pimpl `ImplNetCaffe`).
// The alternative is the one-line swap NetCaffe -> NetB200 (integration/netB200.hpp), like the commented NetOpenCv
// lines of poseExtractorCaffe.cpp:87-91. tests/cpp/extractor_main.cpp runs the reference's unmodified
// PoseExtractorCaffe through this file.
#include <openpose/net/netCaffe.hpp>
#include <openpose/utilities/errorAndLog.hpp>
#include "netB200.hpp"

namespace op
{
    struct NetCaffe::ImplNetCaffe
    {
        NetB200 mNet;

        ImplNetCaffe(const std::string& caffeProto, const std::string& caffeTrainedModel, const int gpuId,
                     const bool enableGoogleLogging, const std::string& lastBlobName) :
            mNet{caffeProto, caffeTrainedModel, gpuId, enableGoogleLogging, lastBlobName}
        {
        }
    };

    NetCaffe::NetCaffe(const std::string& caffeProto, const std::string& caffeTrainedModel, const int gpuId,
                       const bool enableGoogleLogging, const std::string& lastBlobName) :
        upImpl{new ImplNetCaffe{caffeProto, caffeTrainedModel, gpuId, enableGoogleLogging, lastBlobName}}
    {
    }

    NetCaffe::~NetCaffe()
    {
    }

    void NetCaffe::initializationOnThread()
    {
        try
        {
            upImpl->mNet.initializationOnThread();
        }
        catch (const std::exception& e)
        {
            error(e.what(), __LINE__, __FUNCTION__, __FILE__);
        }
    }

    void NetCaffe::forwardPass(const Array<float>& inputData) const
    {
        try
        {
            upImpl->mNet.forwardPass(inputData);
        }
        catch (const std::exception& e)
        {
            error(e.what(), __LINE__, __FUNCTION__, __FILE__);
        }
    }

    std::shared_ptr<ArrayCpuGpu<float>> NetCaffe::getOutputBlobArray() const
    {
        try
        {
            return upImpl->mNet.getOutputBlobArray();
        }
        catch (const std::exception& e)
        {
            error(e.what(), __LINE__, __FUNCTION__, __FILE__);
            return nullptr;
        }
    }
}
