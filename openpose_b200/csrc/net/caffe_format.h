// Product-side readers for the two file formats at the op::Net seam (no protobuf, no Caffe):
//   * prototxt: protobuf text format, the subset Caffe deploy files use
//   * caffemodel: protobuf wire format of caffe.NetParameter
// Field numbers follow 3rdparty/caffe/src/caffe/proto/caffe.proto of the reference tree
// (NetParameter :65-100, LayerParameter :327-408, BlobProto :6-22, BlobShape :1-4).
#pragma once
#include <map>
#include <string>
#include <vector>

namespace opb
{
    struct LayerDef
    {
        std::string name, type;
        std::vector<std::string> bottoms, tops;
        // Convolution (ConvolutionParameter caffe.proto:593-624)
        int numOutput = 0, kernel = 1, pad = 0, stride = 1, group = 1, dilation = 1;
        bool biasTerm = true;
        // Pooling: 0 = MAX, 1 = AVE, 2 = STOCHASTIC
        int poolMethod = 0;
        // Concat
        int axis = 1;
        // PReLU
        bool channelShared = false;
        // Input layer shape (input_param { shape { dim ... } })
        std::vector<int> inputShape;
    };

    struct NetDef
    {
        std::string name;
        std::vector<std::string> inputs;   // legacy top-level "input:" fields
        std::vector<int> inputDim;         // legacy top-level "input_dim:" fields
        std::vector<LayerDef> layers;
    };

    struct BlobData
    {
        std::vector<int> shape;
        std::vector<float> data;
    };
    using WeightMap = std::map<std::string, std::vector<BlobData>>;   // layer name -> blobs

    // Both return false and fill `err` on failure.
    bool parsePrototxt(const std::string& text, NetDef& out, std::string& err);
    bool readCaffemodel(const std::string& path, WeightMap& out, std::string& err);
    bool readFile(const std::string& path, std::string& out);
}
