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

#if defined(__has_include)
#if __has_include("../LKRefine/LKSubPixel.h")
#include "../LKRefine/LKSubPixel.h"  // the package's own sub-pixel refiner (inc/Solver.h:7); never invoked on the path (src/SGM.cpp:823-824)
#define SGM_B200_HAVE_LKREFINE 1
#endif
#endif

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

    // Host MIRRORS of device state, under the reference's names (inc/Solver.h:47-49).  The reference computes into these
    // arrays on the CPU; here the volumes live in device memory and these stay NULL until one of the build_* / fetch_*
    // helpers below is called, which copies what the device computed for the LAST frame (sgm_get_tap).  Nothing on the
    // path reads them.
    uint64_t *cost_table_l, *cost_table_r;  // census tables [img_h * img_w]
    float *cost;                            // cost volume [img_h * img_w * max_disp] (allocated on first use: 4 * W * H * D bytes)
    float *weight;                          // window weights; WEIGHTED_COST is 0 in the reference and here: always NULL

#ifdef SGM_B200_HAVE_LKREFINE
    LKSubPixelPtr lkptr;                    // created like the reference does (src/Solver.cpp:51), never invoked
#endif

    sgm_engine *engine;             // the device pipeline (NULL until a derived constructor called open_engine)
    bool two_views;                 // the pipeline also computes the right ("beta") view (SGM; BM is left-only)

protected:
    // Creates the device pipeline for `algo` (sgm_algo) with the constants above; called once by SGM / BM constructors.
    void open_engine(int algo, int paths, int full_h, int full_w);
    // Runs one frame on the device and refreshes disp / disp_beta / filtered_disp / filtered_disp_beta.
    void run_engine(Mat &l, Mat &r);

    // The reference's protected stage helpers (inc/Solver.h:55-69), kept by name.  In the reference a subclass calls them in
    // sequence to BUILD the result on the CPU; here process() has already run every stage on the device, so they REFRESH the
    // mirrors above from the last frame instead (and the in-place filters have nothing left to do):
    void build_cost_table();            // cost_table_l / cost_table_r <- census tables of the last frame
    void build_dsi();                   // cost <- the (filtered) cost volume of the left view
    void build_dsi_from_table();        //   "
    void build_dsi_from_table_beta();   // cost <- the cost volume of the right view (SGM only)
    void cost_horizontal_filter(int win_size);  // no-op: the volume fetched above is already filtered (fused on the device)
    void cost_vertical_filter(int win_size);    // no-op
    float find_dsi_mean_max();          // prints max and mean of the `cost` mirror, returns the mean (src/Solver.cpp:251-270)
    float find_table_mean_max();        // the same over the cost_table_r mirror, divided like the reference does (:273-293)
    void fetch_cost(float *p);          // src/Solver.cpp:371-374: copies an external volume into `cost`
    void compute_subpixel(const Mat &disp, Mat &filtered_disp);  // filtered_disp <- the device's sub-pixel image for `disp` (left or beta)
    void post_filter();                 // no-op: median and speckle filter ran on the device; filtered_disp holds their result
    // Hooks the reference offered for copying externally computed results into the Mats (src/Solver.cpp:376-400).
    void fetch_disparity(uchar *d);
    void fetch_disparity(float *d);
    // Colour map of filtered_disp into colored_disp (src/Solver.cpp:652-707).
    void colormap();
};

typedef std::shared_ptr<Solver> SolverPtr;
