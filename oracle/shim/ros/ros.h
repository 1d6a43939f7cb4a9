// Empty stand-in: lets the reference headers (inc/global.h, LKRefine/LKSubPixel.h) be
// included on a box without ROS / Eigen / OpenCV highgui. Test infrastructure only.
#pragma once
