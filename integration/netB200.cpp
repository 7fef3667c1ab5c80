// op::NetB200: thin C++ shim, all work happens behind the C ABI (include/opb200.h). Error convention of the
// reference: every body is try { ... } catch -> op::error(e.what(), __LINE__, __FUNCTION__, __FILE__)
// (reference: src/openpose/net/netCaffe.cpp:125-148,154-210,212-261,263-278).
#include "netB200.hpp"
#include <opb200.h>
#include <openpose/core/arrayCpuGpu.hpp>
#include <openpose/utilities/errorAndLog.hpp>

namespace op
{
    struct NetB200::ImplNetB200
    {
        const int mGpuId;
        const std::string mCaffeProto;
        const std::string mCaffeTrainedModel;
        const std::string mLastBlobName;
        opb200_net* pNet = nullptr;
        // The reference hands out a non-owning ArrayCpuGpu view of the live output blob; here the view is
        // re-pointed (set_gpu_data + Reshape) after every forward pass, so cached shared_ptrs keep working.
        std::shared_ptr<ArrayCpuGpu<float>> spOutputBlob;

        ImplNetB200(const std::string& caffeProto, const std::string& caffeTrainedModel, const int gpuId,
                    const std::string& lastBlobName) :
            mGpuId{gpuId}, mCaffeProto{caffeProto}, mCaffeTrainedModel{caffeTrainedModel}, mLastBlobName{lastBlobName}
        {
        }
    };

    static void checkStatus(const opb200_status status, const int line, const std::string& function,
                            const std::string& file)
    {
        if (status != OPB200_OK)
            error(opb200_last_error(), line, function, file);
    }

    NetB200::NetB200(const std::string& caffeProto, const std::string& caffeTrainedModel, const int gpuId,
                     const bool enableGoogleLogging, const std::string& lastBlobName) :
        upImpl{new ImplNetB200{caffeProto, caffeTrainedModel, gpuId, lastBlobName}}
    {
        UNUSED(enableGoogleLogging); // no glog on this path
    }

    NetB200::~NetB200()
    {
        if (upImpl && upImpl->pNet)
            opb200_net_destroy(upImpl->pNet);
    }

    void NetB200::initializationOnThread()
    {
        try
        {
            // File-existence checks (same messages as NetCaffe's constructor), prototxt + caffemodel parsing,
            // graph planning, bf16 weight upload: all inside opb200_net_create.
            checkStatus(opb200_net_create(&upImpl->pNet, upImpl->mCaffeProto.c_str(),
                                          upImpl->mCaffeTrainedModel.c_str(), upImpl->mGpuId,
                                          upImpl->mLastBlobName.c_str()),
                        __LINE__, __FUNCTION__, __FILE__);
            upImpl->spOutputBlob = std::make_shared<ArrayCpuGpu<float>>();
        }
        catch (const std::exception& e)
        {
            error(e.what(), __LINE__, __FUNCTION__, __FILE__);
        }
    }

    void NetB200::forwardPass(const Array<float>& inputData) const
    {
        try
        {
            // Sanity checks (reference: netCaffe.cpp:218-222)
            if (inputData.empty())
                error("The Array inputData cannot be empty.", __LINE__, __FUNCTION__, __FILE__);
            if (inputData.getNumberDimensions() != 4 || inputData.getSize(1) != 3)
                error("The Array inputData must have 4 dimensions: [batch size, 3 (RGB), height, width].",
                      __LINE__, __FUNCTION__, __FILE__);
            const int shape[4] = {inputData.getSize(0), inputData.getSize(1), inputData.getSize(2),
                                  inputData.getSize(3)};
            // Host input, legacy default stream: results are stream-ordered before any later default-stream
            // work of the caller (resizeAndMergeGpu, cudaMemcpy ...), as with Caffe's ForwardFrom(0).
            checkStatus(opb200_net_forward(upImpl->pNet, inputData.getConstPtr(), 0, shape, nullptr),
                        __LINE__, __FUNCTION__, __FILE__);
            const float* devicePtr = nullptr;
            int outShape[4];
            checkStatus(opb200_net_output(upImpl->pNet, &devicePtr, outShape), __LINE__, __FUNCTION__, __FILE__);
            upImpl->spOutputBlob->Reshape(outShape[0], outShape[1], outShape[2], outShape[3]);
            upImpl->spOutputBlob->set_gpu_data(const_cast<float*>(devicePtr));
        }
        catch (const std::exception& e)
        {
            error(e.what(), __LINE__, __FUNCTION__, __FILE__);
        }
    }

    std::shared_ptr<ArrayCpuGpu<float>> NetB200::getOutputBlobArray() const
    {
        try
        {
            return upImpl->spOutputBlob;
        }
        catch (const std::exception& e)
        {
            error(e.what(), __LINE__, __FUNCTION__, __FILE__);
            return nullptr;
        }
    }
}
