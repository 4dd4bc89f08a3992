// Minimal cv::Mat stand-in: only what the reference's model.h touches (type(), at<float>(r,c)).
#pragma once
#include <cstddef>
#ifndef CV_32F
#define CV_32F 5
#endif
namespace cv {
struct Mat {
    int rows = 0, cols = 0; float* p = nullptr;
    int type() const { return CV_32F; }
    template <class T> T& at(int r, int c) { return ((T*)p)[r * cols + c]; }
    template <class T> const T& at(int r, int c) const { return ((const T*)p)[r * cols + c]; }
};
}
