// TEST HARNESS ONLY. A small FUNCTIONAL stand-in for the parts of OpenCV's C++ API the reference's hand / face
// extractors and core/matrix.cpp use (this image has no OpenCV C++; a real OpenPose build has the real library):
// cv::Mat as a reference-counted row-major buffer (eye, at<T>, clone, ROI view, the shape queries of matrix.cpp),
// cv::Size / Rect / Scalar, and cv::warpAffine for 8-bit 3-channel images with INTER_LINEAR | WARP_INVERSE_MAP and
// BORDER_CONSTANT -- implemented by oracle/post_oracle.c:opo_warp_affine_u8, the restatement of OpenCV's fixed-point
// bilinear warp that tests/test_crop_cpu.py holds bit-equal to cv2 4.13.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <vector>

#define CV_8U 0
#define CV_32F 5
#define CV_64F 6
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn) - 1) << 3))
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_INTER_LINEAR 1
#define CV_WARP_INVERSE_MAP 16

extern "C" void opo_warp_affine_u8(unsigned char* dst, const unsigned char* frame, const double* m, int hs, int ws, int ht,
                                   int wt);

namespace cv
{
    enum { INTER_LINEAR = 1, WARP_INVERSE_MAP = 16, BORDER_CONSTANT = 0 };
    struct Size
    {
        int width, height;
    };
    struct Rect
    {
        int x, y, width, height;
    };
    struct Scalar
    {
        double v[4];
        Scalar(double a = 0, double b = 0, double c = 0, double d = 0) : v{a, b, c, d} {}
    };

    class Mat
    {
    public:
        int rows = 0, cols = 0;
        unsigned char* data = nullptr;
        struct SizeProxy
        {
            const Mat* m;
            int operator[](int i) const { return i == 0 ? m->rows : m->cols; }
        } size{this};

        Mat() {}
        Mat(int r, int c, int type) { create(r, c, type); }
        Mat(int r, int c, int type, void* d) : rows(r), cols(c), data((unsigned char*)d), mType(type), mStep((size_t)c * elemSizeOf(type)) {}
        Mat(const Mat& m, const Rect& roi)
            : rows(roi.height), cols(roi.width), data(m.data + (size_t)roi.y * m.mStep + (size_t)roi.x * m.elemSize()), mType(m.mType),
              mStep(m.mStep), mBuf(m.mBuf), mSub(true)
        {
        }
        Mat(const Mat& o) : rows(o.rows), cols(o.cols), data(o.data), mType(o.mType), mStep(o.mStep), mBuf(o.mBuf), mSub(o.mSub) {}
        Mat& operator=(const Mat& o)
        {
            rows = o.rows; cols = o.cols; data = o.data; mType = o.mType; mStep = o.mStep; mBuf = o.mBuf; mSub = o.mSub;
            return *this;
        }
        void create(int r, int c, int type)
        {
            rows = r; cols = c; mType = type; mStep = (size_t)c * elemSizeOf(type); mSub = false;
            mBuf = std::make_shared<std::vector<unsigned char>>((size_t)r * mStep, (unsigned char)0);
            data = mBuf->data();
        }
        static Mat eye(int r, int c, int type)
        {
            Mat m(r, c, type);
            for (int i = 0; i < std::min(r, c); ++i)
            {
                if (type == CV_64F) m.at<double>(i, i) = 1.0;
                else if (type == CV_32F) m.at<float>(i, i) = 1.f;
                else m.data[(size_t)i * m.mStep + (size_t)i * m.elemSize()] = 1;
            }
            return m;
        }
        template <typename T> T& at(int r, int c) { return *reinterpret_cast<T*>(data + (size_t)r * mStep + (size_t)c * sizeof(T)); }
        template <typename T> const T& at(int r, int c) const { return *reinterpret_cast<const T*>(data + (size_t)r * mStep + (size_t)c * sizeof(T)); }
        Mat clone() const
        {
            Mat m(rows, cols, mType);
            for (int r = 0; r < rows; ++r)
                std::memcpy(m.data + (size_t)r * m.mStep, data + (size_t)r * mStep, (size_t)cols * elemSize());
            return m;
        }
        void copyTo(Mat& o) const { o = clone(); }
        Mat& setTo(const Scalar& s)
        {
            if (depth() != CV_8U)
                throw std::runtime_error("stand-in cv::Mat::setTo: 8-bit only");
            for (int r = 0; r < rows; ++r)
                std::memset(data + (size_t)r * mStep, (int)s.v[0], (size_t)cols * elemSize());
            return *this;
        }
        bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
        int type() const { return mType; }
        int depth() const { return mType & 7; }
        int channels() const { return (mType >> 3) + 1; }
        int dims = 2;
        size_t elemSize() const { return elemSizeOf(mType); }
        size_t elemSize1() const { return depthSize(mType & 7); }
        size_t step1(int i = 0) const { return i == 0 ? mStep / elemSize1() : channels(); }
        size_t total() const { return (size_t)rows * cols; }
        bool isContinuous() const { return mStep == (size_t)cols * elemSize(); }
        bool isSubmatrix() const { return mSub; }
        int checkVector(int elemChannels, int d = -1, bool requireContinuous = true) const
        {
            return ((d < 0 || depth() == d) && (isContinuous() || !requireContinuous) && channels() == elemChannels) ? (int)total() : -1;
        }

    private:
        static size_t depthSize(int d) { return d == CV_64F ? 8 : (d == CV_32F ? 4 : (d >= 2 && d <= 3 ? 2 : (d == 4 ? 4 : 1))); }
        static size_t elemSizeOf(int type) { return depthSize(type & 7) * (size_t)((type >> 3) + 1); }
        int mType = 0;
        size_t mStep = 0;
        std::shared_ptr<std::vector<unsigned char>> mBuf;
        bool mSub = false;
    };

    // 8-bit, 3 channels, INTER_LINEAR | WARP_INVERSE_MAP, BORDER_CONSTANT (0): the one call the extractors make
    inline void warpAffine(const Mat& src, Mat& dst, const Mat& M, Size dsize, int flags, int borderMode, const Scalar& border)
    {
        if (src.type() != CV_8UC3 || !src.isContinuous() || M.type() != CV_64F || M.rows != 2 || M.cols != 3 ||
            flags != (INTER_LINEAR | WARP_INVERSE_MAP) || borderMode != BORDER_CONSTANT || border.v[0] != 0 || border.v[1] != 0 ||
            border.v[2] != 0)
            throw std::runtime_error("stand-in cv::warpAffine: only the hand / face extractors' call is implemented");
        const double m[6] = {M.at<double>(0, 0), M.at<double>(0, 1), M.at<double>(0, 2),
                             M.at<double>(1, 0), M.at<double>(1, 1), M.at<double>(1, 2)};
        dst.create(dsize.height, dsize.width, CV_8UC3);
        opo_warp_affine_u8(dst.data, src.data, m, src.rows, src.cols, dsize.height, dsize.width);
    }
}
