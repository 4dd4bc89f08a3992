// Minimal cv::Mat stand-in for the oracle builds of the reference's headers/sources (OpenCV is not installed here).
// Only what those sources touch: rows/cols, type(), at<T>(r,c), ptr<T>(), and convertTo for the one conversion
// the reference performs (CV_32S depth -> CV_16U with saturation, cuda_icp/scene/common.cpp:23).
#pragma once
#include <cstddef>
#include <cstdint>
#include <memory>
#include <vector>
#ifndef CV_32F
#define CV_16U 2
#define CV_32S 4
#define CV_32F 5
#endif
typedef unsigned short ushort;
namespace cv {
struct Mat {
    int rows = 0, cols = 0;
    float* p = nullptr;          // element storage (any type; `t` says which)
    int t = CV_32F;
    std::shared_ptr<std::vector<unsigned char>> own;  // set when the Mat owns its storage (convertTo result)
    int type() const { return t; }
    template <class T> T& at(int r, int c) { return ((T*)p)[(size_t)r * cols + c]; }
    template <class T> const T& at(int r, int c) const { return ((const T*)p)[(size_t)r * cols + c]; }
    template <class T> T* ptr(int r = 0) { return ((T*)p) + (size_t)r * cols; }
    template <class T> const T* ptr(int r = 0) const { return ((const T*)p) + (size_t)r * cols; }
    void convertTo(Mat& dst, int rtype) const {
        // saturate_cast<ushort>(int): clamp to [0, 65535]
        dst.rows = rows; dst.cols = cols; dst.t = rtype;
        dst.own = std::make_shared<std::vector<unsigned char>>((size_t)rows * cols * 2);
        unsigned short* o = (unsigned short*)dst.own->data();
        const int32_t* in = (const int32_t*)p;
        for (size_t i = 0; i < (size_t)rows * cols; ++i) o[i] = (unsigned short)(in[i] < 0 ? 0 : (in[i] > 65535 ? 65535 : in[i]));
        dst.p = (float*)o;
    }
};
}
