// Solver.h — drop-in replacement for the reference's inc/Solver.h (hilbertw/stereo_matching, inc/Solver.h:1-73).
//
// Same public surface (ctor (h, w, s, d), the two process() overloads, show_disp(), get_disp(), SolverPtr, the tuning
// constants), so node.cpp:49,93,104,107 and the ROS launch file compile and run unchanged.  What is different is what
// sits behind it: the reference keeps a float cost volume, census tables and eight path volumes on the host and walks
// them with OpenMP loops; here the object owns an sgm_engine handle (include/sgm_b200.h) and every stage of process()
// runs as sm_100a CUDA kernels.  There is no CPU matching path in this class: if the engine cannot be created (no CUDA
// device) the constructor reports the error and aborts, which is also what the reference's asserts do on bad arguments
// (src/Solver.cpp:6-10).
//
// Install: copy inc/{Solver,SGM,BM,cost}.h and src/{Solver,SGM,BM,cost}.cpp of this directory over the reference's files
// of the same names, add include/ of this repository to the include path and link libsgm_b200.so.
#pragma once

#include "global.h"  // the package's umbrella include (OpenCV, ROS, `using namespace cv`): the reference's own file, unchanged
#include "utils.h"
#include "cost.h"

#include <vector>

#include "sgm_b200.h"

// Matching parameters.  Names and values are the reference's (inc/Solver.h:10-20); they are handed to the engine through
// sgm_params, so editing them here changes the device pipeline exactly like editing them changed the CPU one.
const int WIN_H = 7;              // census window, rows
const int WIN_W = 9;              // census window, columns
const int COST_WIN_H = 3;         // recursive cost filter, vertical window
const int COST_WIN_W = 5;         // recursive cost filter, horizontal window
const float UNIQUE_RATIO = 0.7;   // uniqueness: best / second-best above this and |dd| > gap -> invalid
const bool WEIGHTED_COST = 0;     // kept for source compatibility; weighted windows were never enabled in the reference
const float LR_CHECK_DIS = 1;     // left-right consistency tolerance
const int MEDIAN_FILTER_H = 5;    // hole-filling median window
const int MEDIAN_FILTER_W = 5;
const int SPECKLE_SIZE = 1000;    // components up to this many pixels are removed
const int SPECKLE_DIS = 2;        // |a - b| < this joins two pixels into one component


class Solver
{
public:
    explicit Solver(int h, int w, int s, int d);
    virtual ~Solver();

    Solver(const Solver&) = delete;
    Solver& operator=(Solver&) = delete;

    // One stereo pair in, disparity ready in get_disp() on return (node.cpp:93).  Images are CV_8UC1, full
    // resolution (h x w as given to the constructor); masks are CV_8UC1 at working resolution with 255 = sky.
    virtual void process(Mat &img_l, Mat &img_r) = 0;
    virtual void process(Mat &img_l, Mat &img_r, Mat &sky_mask, Mat &sky_mask_beta) = 0;

    // Left image stacked over the colour-mapped disparity, as the reference draws it (src/Solver.cpp:55-93).
    virtual void show_disp(Mat &debug_view);
    // CV_32FC1, working resolution, invalid pixels == max_disp + 1 (node.cpp:104,129).  Valid until the next process().
    virtual const Mat& get_disp() const { return filtered_disp; }

    // Addition to the reference surface: the disparity -> point cloud loop of node.cpp:119-143 as one device call
    // (same arithmetic, same raster order).  Returns the number of points.
    size_t get_point_cloud(const CamIntrinsics &cam, double baseline, float max_range, std::vector<cv::Point3d> &pts,
                           std::vector<uchar> &pixel);

protected:
    int img_h, img_w;               // working resolution (h / s, w / s)
    int scale;
    int max_disp, invalid_disp;

    Mat img_l, img_r;               // the caller's images of the last process() (headers only; not copied)
    Mat disp, disp_beta;            // CV_8UC1 integer disparities of the left / right view (right: SGM only)
    Mat filtered_disp;              // CV_32FC1 final result
    Mat filtered_disp_beta;         // CV_32FC1 sub-pixel disparity of the right view (SGM only)
    Mat colored_disp;               // CV_8UC3 scratch of show_disp()
    Mat sky_mask, sky_mask_beta;    // masks handed to the last 4-argument process() (kept between frames like the reference)

    sgm_engine *engine;             // the device pipeline (NULL until a derived constructor called open_engine)
    bool two_views;                 // the pipeline also computes the right ("beta") view (SGM; BM is left-only)

protected:
    // Creates the device pipeline for `algo` (sgm_algo) with the constants above; called once by SGM / BM constructors.
    void open_engine(int algo, int paths, int full_h, int full_w);
    // Runs one frame on the device and refreshes disp / disp_beta / filtered_disp / filtered_disp_beta.
    void run_engine(Mat &l, Mat &r);
    // Hooks the reference offered for copying externally computed results into the Mats (src/Solver.cpp:376-400).
    void fetch_disparity(uchar *d);
    void fetch_disparity(float *d);
    // Colour map of filtered_disp into colored_disp (src/Solver.cpp:652-707).
    void colormap();
};

typedef std::shared_ptr<Solver> SolverPtr;
