// Minimal stand-in for <opencv2/core.hpp> (OpenCV headers are not in the build container): just
// enough of cv::Mat for tests/test_cpp_host.py to COMPILE and run the cv::Mat overloads of
// include/TSDFfusion.hpp (-DTSDF_B200_WITH_OPENCV).  Test infrastructure only.
#pragma once
#include <cstring>
#include <vector>
#define CV_8U 0
#define CV_32F 5
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn)-1) << 3))
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
namespace cv {
class Mat {
 public:
  Mat() : rows(0), cols(0), data(nullptr), type_(0) {}
  Mat(int r, int c, int type, void* ext) : rows(r), cols(c), data(static_cast<unsigned char*>(ext)), type_(type) {}
  Mat(int r, int c, int type) : rows(r), cols(c), type_(type) {
    own_.resize((size_t)r * c * elem_size());
    data = own_.data();
  }
  int type() const { return type_; }
  bool isContinuous() const { return true; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  void convertTo(Mat& dst, int rtype) const {  // only the CV_32F -> CV_32F case is needed
    (void)rtype;
    dst = Mat(rows, cols, type_);
    std::memcpy(dst.data, data, own_size());
  }
  template <typename T> T& at(int r, int c) { return reinterpret_cast<T*>(data)[(size_t)r * cols + c]; }
  int rows, cols;
  unsigned char* data;
 private:
  size_t elem_size() const { return ((type_ & 7) == CV_32F ? 4 : 1) * ((type_ >> 3) + 1); }
  size_t own_size() const { return (size_t)rows * cols * elem_size(); }
  int type_;
  std::vector<unsigned char> own_;
};
}  // namespace cv
