// op::NetB200 -- drop-in op::Net backend over libopb200.so (C ABI: include/opb200.h).
// Same constructor signature as op::NetCaffe (reference: include/openpose/net/netCaffe.hpp:12-13), so the three
// construction sites (src/openpose/pose/poseExtractorCaffe.cpp:82-86, hand/handExtractorCaffe.cpp:34-35,
// face/faceExtractorCaffe.cpp:32) become one-line swaps, exactly like the commented NetOpenCv alternative.
#ifndef OPENPOSE_NET_NET_B200_HPP
#define OPENPOSE_NET_NET_B200_HPP

#include <openpose/core/common.hpp>
#include <openpose/net/net.hpp>

namespace op
{
    class OP_API NetB200 : public Net
    {
    public:
        NetB200(const std::string& caffeProto, const std::string& caffeTrainedModel, const int gpuId = 0,
                const bool enableGoogleLogging = true, const std::string& lastBlobName = "net_output");

        virtual ~NetB200();

        void initializationOnThread();

        void forwardPass(const Array<float>& inputNetData) const;

        std::shared_ptr<ArrayCpuGpu<float>> getOutputBlobArray() const;

    private:
        struct ImplNetB200;
        std::unique_ptr<ImplNetB200> upImpl;

        DELETE_COPY(NetB200);
    };
}

#endif // OPENPOSE_NET_NET_B200_HPP
