// TEST INFRASTRUCTURE (oracle). Type-only stand-in for <opencv2/imgproc/imgproc.hpp> so the reference's utilities/keypoint.cpp
// compiles without OpenCV: only renderKeypointsCpu uses these names and the oracle never calls it.
#pragma once
#include <cmath>
#define CV_32FC3 21
namespace cv
{
    struct Mat
    {
        unsigned char* data = nullptr;
        Mat() {}
        int channels() const { return 3; }
        int dims = 2;
        int rows = 0, cols = 0;
        struct { int p[3]; int operator[](int i) const { return p[i]; } } size{};
        Mat(int, int, int, void* d) : data((unsigned char*)d) {}
    };
    struct Scalar
    {
        double v[4];
        Scalar(double a = 0, double b = 0, double c = 0, double d = 0) : v{a, b, c, d} {}
    };
    struct Point
    {
        int x, y;
        Point(int a = 0, int b = 0) : x(a), y(b) {}
    };
    inline void line(Mat&, Point, Point, const Scalar&, int, int, int) {}
    inline void circle(Mat&, Point, int, const Scalar&, int, int, int) {}
}
