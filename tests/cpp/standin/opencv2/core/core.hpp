// TEST HARNESS ONLY. Type-only stand-in for <opencv2/core/core.hpp> so the reference's pose/poseExtractorCaffe.cpp
// compiles in this image (no OpenCV C++ here; a real OpenPose build has the real one). The only code that names
// these types is the TOP_DOWN_REFINEMENT branch (poseExtractorCaffe.cpp:13,405-470), a compile-time `false`.
#pragma once
#include <algorithm> // the real header brings <algorithm> / <cmath> along; poseExtractorCaffe.cpp:299,388 rely on that
#include <cmath>
#define CV_32FC1 5
namespace cv
{
    struct Rect
    {
        int x, y, width, height;
    };
    struct Scalar
    {
        double v[4];
        Scalar(double a = 0, double b = 0, double c = 0, double d = 0) : v{a, b, c, d} {}
    };
    struct Mat
    {
        Mat() {}
        Mat(int, int, int, const void*) {}
        Mat(const Mat&, const Rect&) {}
    };
    enum { BORDER_CONSTANT = 0 };
}
