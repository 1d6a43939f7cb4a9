// highgui / imgproc entry points node.cpp calls (imread, resize, imwrite, waitKey, destroyAllWindows), declared only:
// the compile check never links.  TEST INFRASTRUCTURE.
#pragma once
#include <string>
#include <opencv2/core/core.hpp>
namespace cv {
enum { IMREAD_GRAYSCALE = 0 };
Mat imread(const std::string& path, int flags);
bool imwrite(const std::string& path, const Mat& img);
void resize(const Mat& src, Mat& dst, Size size);
int waitKey(int delay);
void destroyAllWindows();
}  // namespace cv
