// Stand-in for the reference package's inc/global.h when the host shells are built OUTSIDE the catkin package (tests on
// a box without ROS / C++ OpenCV).  In the real package the reference's own global.h is used unchanged.
// TEST INFRASTRUCTURE.
#pragma once
#include <stdint.h>
#include <unistd.h>

#include <fstream>
#include <iostream>
#include <memory>
#include <sstream>

#include <opencv2/core/core.hpp>  // oracle/shim stand-in (or the real OpenCV when present)
#include <opencv2/imgproc.hpp>

using namespace cv;
