// TEST INFRASTRUCTURE (oracle). Type-only stand-in for <opencv2/highgui/highgui.hpp> so the reference's filestream/fileStream.cpp
// compiles without OpenCV: only its image / cv::FileStorage functions use these names and the oracle never calls them
// (it drives savePeopleJson, which is plain std::ofstream code).
#pragma once
#define CV_MAJOR_VERSION 4
#include <string>
#include <vector>
namespace cv
{
    struct Mat
    {
        bool empty() const { return true; }
    };
    struct FileNode
    {
        void operator>>(Mat&) const {}
    };
    struct FileStorage
    {
        enum { READ = 0, WRITE = 1 };
        FileStorage(const std::string&, int) {}
        bool isOpened() const { return false; }
        void release() {}
        FileNode operator[](const std::string&) const { return FileNode(); }
    };
    inline FileStorage& operator<<(FileStorage& fs, const std::string&) { return fs; }
    inline FileStorage& operator<<(FileStorage& fs, const Mat&) { return fs; }
    inline bool imwrite(const std::string&, const Mat&, const std::vector<int>&) { return false; }
    inline Mat imread(const std::string&, int) { return Mat(); }
}
