// Stand-in for <caffe/blob.hpp> when OpenPose is built with -DUSE_CAFFE but WITHOUT Caffe (the NetB200 backend).
//
// The reference's *Caffe layer classes and extractors (src/openpose/net/{resizeAndMerge,nms,bodyPartConnector,maximum}Caffe.cpp,
// pose/poseExtractorCaffe.cpp, hand/handExtractorCaffe.cpp, face/faceExtractorCaffe.cpp) include this header behind
// USE_CAFFE but use no caffe:: symbol: their blobs are op::ArrayCpuGpu<T> (backed by integration/arrayCpuGpuB200.cpp).
// What they do lean on is what Caffe's headers pull in transitively:
//   * <cmath> (std::round, resizeAndMergeCaffe.cpp:80-81),
//   * the CUDA runtime API (cudaMalloc / cudaMemcpy / cudaFree, bodyPartConnectorCaffe.cpp:346-383),
//   * the NOT_IMPLEMENTED macro of caffe/common.hpp (the layers' Backward_* bodies; glog's LOG(FATAL) there, an
//     exception here -- never reached at inference).
#ifndef OPB200_STANDIN_CAFFE_BLOB_HPP
#define OPB200_STANDIN_CAFFE_BLOB_HPP
#include <cmath>
#include <stdexcept>
#ifdef USE_CUDA
    #include <cuda.h>
    #include <cuda_runtime.h>
#endif
#ifndef NOT_IMPLEMENTED
    #define NOT_IMPLEMENTED throw std::runtime_error("Not Implemented Yet")
#endif
#endif // OPB200_STANDIN_CAFFE_BLOB_HPP
