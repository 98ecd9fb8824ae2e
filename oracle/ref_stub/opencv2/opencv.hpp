// Stand-in for <opencv2/opencv.hpp>, used ONLY to compile the reference's hot-path
// sources into oracle/_ref (OpenCV is not installed in this image).  The reference
// touches OpenCV solely inside Image::Load/Save and ColorImage::Load/Save
// (include/vulcan/image.h:99-133,230-266, src/image.cu:213-273), which the parity and
// bench harness never call: every function here aborts if it is ever reached.
#pragma once
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <string>

#define CV_LOAD_IMAGE_ANYDEPTH 2
#define CV_LOAD_IMAGE_ANYCOLOR 4
#define CV_RGB2GRAY 7
#define CV_GRAY2RGB 8
#define CV_BGR2RGB 4
#define CV_RGB2BGR 4
#define CV_8UC1 0
#define CV_8UC3 16
#define CV_32FC1 5
#define CV_32FC3 21

namespace cv {

[[noreturn]] inline void stub_abort(const char* what) {
  std::fprintf(stderr, "opencv stub: %s is not available in this build\n", what);
  std::abort();
}

struct Mat {
  unsigned char* data = nullptr;
  int cols = 0, rows = 0;
  Mat() {}
  Mat(int, int, int, void*) { stub_abort("cv::Mat(rows, cols, type, data)"); }
  int channels() const { return 1; }
  size_t total() const { return (size_t)cols * rows; }
  Mat clone() const { stub_abort("cv::Mat::clone"); }
  void convertTo(Mat&, int, double = 1, double = 0) const { stub_abort("cv::Mat::convertTo"); }
};

inline Mat imread(const std::string&, int) { stub_abort("cv::imread"); }
inline bool imwrite(const std::string&, const Mat&) { stub_abort("cv::imwrite"); }
inline void cvtColor(const Mat&, Mat&, int) { stub_abort("cv::cvtColor"); }

}  // namespace cv
