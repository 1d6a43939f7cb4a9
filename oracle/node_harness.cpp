// oracle/node_harness.cpp — TEST INFRASTRUCTURE, not product code.
//
// Lets the reference's UNMODIFIED node.cpp (compiled where it lies, together with its unmodified src/{Solver,SGM,BM,cost}.cpp)
// link and run on a box without ROS, OpenCV or a camera, so that its disparity -> point cloud loop (node.cpp:119-143, which
// lives inside main()) can be the oracle of sgm_point_cloud.  This file only DEFINES the I/O entry points node.cpp calls and
// the stand-in headers merely declare: imread / resize / imwrite / waitKey, the sky detector, read_calib, the three publishers,
// num2str / get_cur_ms.  imread feeds a raw 8-bit image named by the environment; publish_float32 / publish_pointcloud write
// what node.cpp hands them to files.
//   NODE_LEFT_RAW, NODE_RIGHT_RAW   raw uint8 images of ROS_PARAM__sgm_node_img_h x ROS_PARAM__sgm_node_img_w (see ros_stubs/ros/ros.h)
//   NODE_OUT                        prefix of the outputs: <prefix>_disp.f32, <prefix>_xyz.f64, <prefix>_int.u8
//   NODE_SKY_ROWS                   rows (working resolution) of synthetic sky at the top of both masks; unset / 0 = empty masks
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "inc/global.h"
#include "inc/roshelper.h"
#include "inc/utils.h"
#include "sky_detector/imageSkyDetector.h"
#include "LKRefine/LKSubPixel.h"

static int env_int(const char* name, int fallback) {
    const char* v = std::getenv(name);
    return v ? std::atoi(v) : fallback;
}
static std::string env_str(const char* name) {
    const char* v = std::getenv(name);
    return v ? std::string(v) : std::string();
}

// ---- src/utils.cpp (not compiled: it needs highgui / VideoCapture) ----
double get_cur_ms() {
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}
std::string num2str(int i) { char b[32]; std::snprintf(b, sizeof(b), "%06d", i); return b; }
std::string num2strbeta(int i) { char b[32]; std::snprintf(b, sizeof(b), "%02d", i); return b; }
CamIntrinsics read_calib(std::string) { return CamIntrinsics{718.856f, 718.856f, 607.1928f, 185.2157f}; }  // KITTI-like, fixed

// ---- LKRefine (needs Eigen; created by Solver's ctor, never invoked: src/SGM.cpp:823-824) ----
namespace {
class NoopLK : public LKSubPixel {
public:
    NoopLK(int h, int w, int s, int d) : LKSubPixel(h, w, s, d) {}
    void LKRefine(const cv::Mat&, const cv::Mat&, cv::Mat&) override {}
};
}  // namespace
LKSubPixel::LKSubPixel(int, int, int, int) {}
LKSubPixel::~LKSubPixel() {}
std::shared_ptr<LKSubPixel> LKSubPixel::create(int h, int w, int s, int d) { return std::make_shared<NoopLK>(h, w, s, d); }

// ---- highgui / imgproc ----
namespace cv {
Mat imread(const std::string& path, int) {
    const int h = env_int("ROS_PARAM__sgm_node_img_h", 360), w = env_int("ROS_PARAM__sgm_node_img_w", 1240);
    const std::string raw = env_str(path.find("image_0") != std::string::npos ? "NODE_LEFT_RAW" : "NODE_RIGHT_RAW");
    Mat m(h, w, CV_8UC1);
    FILE* f = std::fopen(raw.c_str(), "rb");
    if (!f || std::fread(m.data, 1, (size_t)h * w, f) != (size_t)h * w) {
        std::fprintf(stderr, "node_harness: cannot read %s\n", raw.c_str());
        std::exit(3);
    }
    std::fclose(f);
    return m;
}
bool imwrite(const std::string&, const Mat&) { return true; }
void resize(const Mat& src, Mat& dst, Size size) {
    if (src.rows != size.height || src.cols != size.width) {
        std::fprintf(stderr, "node_harness: resize is the identity only\n");
        std::exit(3);
    }
    Mat c = src.clone();
    dst = c;
}
int waitKey(int) { return 0; }
void destroyAllWindows() {}
}  // namespace cv

// ---- sky detector: synthetic masks (node.cpp:82-89) ----
namespace sky_detector {
void SkyAreaDetector::detect(const cv::Mat& img, const std::string, cv::Mat& sky_label, int scale) {
    const int rows = env_int("NODE_SKY_ROWS", 0);
    if (rows <= 0) {
        sky_label = cv::Mat();
        return;
    }
    sky_label = cv::Mat::zeros(img.rows / scale, img.cols / scale, CV_8UC1);
    for (int i = 0; i < rows && i < sky_label.rows; ++i) std::memset(sky_label.ptr<uchar>(i), 255, (size_t)sky_label.cols);
}
}  // namespace sky_detector

// ---- src/roshelper.cpp: the publishers capture what node.cpp hands them ----
static void dump(const std::string& suffix, const void* p, size_t bytes) {
    FILE* f = std::fopen((env_str("NODE_OUT") + suffix).c_str(), "wb");
    if (!f || std::fwrite(p, 1, bytes, f) != bytes) std::exit(4);
    std::fclose(f);
}
void publish_float32(image_transport::Publisher, const cv::Mat& img) {
    std::vector<float> v((size_t)img.rows * img.cols);
    for (int i = 0; i < img.rows; ++i) std::memcpy(v.data() + (size_t)i * img.cols, img.ptr<float>(i), (size_t)img.cols * sizeof(float));
    dump("_disp.f32", v.data(), v.size() * sizeof(float));
}
void publish_rgb(image_transport::Publisher, const cv::Mat&) {}
void publish_pointcloud(ros::Publisher, const std::vector<cv::Point3d>& pts, const std::vector<uchar>& intensity) {
    std::vector<double> xyz(pts.size() * 3);
    for (size_t k = 0; k < pts.size(); ++k) {
        xyz[3 * k] = pts[k].x;
        xyz[3 * k + 1] = pts[k].y;
        xyz[3 * k + 2] = pts[k].z;
    }
    dump("_xyz.f64", xyz.data(), xyz.size() * sizeof(double));
    dump("_int.u8", intensity.data(), intensity.size());
}
