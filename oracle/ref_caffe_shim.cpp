// TEST INFRASTRUCTURE (oracle) -- never linked into the product.
//
// Pins the convolution arithmetic of the net oracle (oracle/ref_net.py, a torch restatement) on the reference's own
// bundled Caffe code: 3rdparty/caffe/src/caffe/util/im2col.cpp is compiled UNMODIFIED (stand-in headers for glog /
// cblas / caffe/common.hpp under oracle/caffe_standin) and driven exactly as
// BaseConvolutionLayer::forward_cpu_gemm + forward_cpu_bias do (3rdparty/caffe/src/caffe/layers/base_conv_layer.cpp:
// 258-281): col = im2col_cpu(input); output = weights (Cout x Cin*kh*kw) * col (Cin*kh*kw x H*W); output += bias * ones.
// The GEMM itself is an external BLAS in Caffe (Atlas / OpenBLAS / MKL chosen at build time, summation order
// unspecified), so it is a plain float triple loop here. MAX pooling restates PoolingLayer::Forward_cpu
// (layers/pooling_layer.cpp:141-175 with the ceil-mode shape rule of :91-111), which cannot be compiled alone (Layer /
// Blob / protobuf).
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <vector>
#include "caffe/util/im2col.hpp"

namespace caffe
{
    // link support: col2im_cpu in the same object names caffe_set (util/math_functions.cpp:56-66, not compilable alone:
    // boost RNG); never called by the oracle
    template <typename Dtype>
    void caffe_set(const int N, const Dtype alpha, Dtype* Y)
    {
        for (int i = 0; i < N; ++i)
            Y[i] = alpha;
    }
    template void caffe_set<float>(const int, const float, float*);
    template void caffe_set<double>(const int, const double, double*);
}

extern "C"
{
    // one image: input (cin, h, w), weights (cout, cin, k, k), bias (cout) -> output (cout, h, w); stride 1, pad k/2
    void refcaffe_conv(float* output, const float* input, const float* weights, const float* bias, int cin, int cout,
                       int h, int w, int k, int pad)
    {
        const int outH = h + 2 * pad - k + 1, outW = w + 2 * pad - k + 1;   // conv_layer.cpp:8-22
        const int kernelDim = cin * k * k, spatial = outH * outW;
        std::vector<float> col((size_t)kernelDim * spatial);
        const float* colBuff = input;
        if (!(k == 1 && pad == 0))                                          // is_1x1_: the input is its own column buffer
        {
            caffe::im2col_cpu<float>(input, cin, h, w, k, k, pad, pad, 1, 1, 1, 1, col.data());
            colBuff = col.data();
        }
        for (int m = 0; m < cout; ++m)                                      // caffe_cpu_gemm(NoTrans, NoTrans, M, N, K, 1, W, col, 0, out)
        {
            float* out = output + (size_t)m * spatial;
            std::fill(out, out + spatial, 0.f);
            for (int kk = 0; kk < kernelDim; ++kk)
            {
                const float wv = weights[(size_t)m * kernelDim + kk];
                const float* c = colBuff + (size_t)kk * spatial;
                for (int n = 0; n < spatial; ++n)
                    out[n] += wv * c[n];
            }
            for (int n = 0; n < spatial; ++n)                               // forward_cpu_bias: += bias[m] * 1
                out[n] += bias[m] * 1.f;
        }
    }

    // MAX pooling, one image (c, h, w) -> (c, ph, pw); returns ph * 65536 + pw
    int refcaffe_max_pool(float* top, const float* bottom, int channels, int h, int w, int kernel, int stride, int pad)
    {
        int ph = (int)std::ceil((float)(h + 2 * pad - kernel) / stride) + 1;   // pooling_layer.cpp:91-94
        int pw = (int)std::ceil((float)(w + 2 * pad - kernel) / stride) + 1;
        if (pad)                                                               // :95-105
        {
            if ((ph - 1) * stride >= h + pad)
                --ph;
            if ((pw - 1) * stride >= w + pad)
                --pw;
        }
        if (top == nullptr)
            return ph * 65536 + pw;
        for (int c = 0; c < channels; ++c)
        {
            const float* b = bottom + (size_t)c * h * w;
            float* t = top + (size_t)c * ph * pw;
            std::fill(t, t + (size_t)ph * pw, -FLT_MAX);
            for (int y = 0; y < ph; ++y)
                for (int x = 0; x < pw; ++x)
                {
                    int hstart = y * stride - pad, wstart = x * stride - pad;
                    const int hend = std::min(hstart + kernel, h), wend = std::min(wstart + kernel, w);
                    hstart = std::max(hstart, 0);
                    wstart = std::max(wstart, 0);
                    for (int yy = hstart; yy < hend; ++yy)
                        for (int xx = wstart; xx < wend; ++xx)
                            if (b[yy * w + xx] > t[y * pw + x])
                                t[y * pw + x] = b[yy * w + xx];
                }
        }
        return ph * 65536 + pw;
    }
}
