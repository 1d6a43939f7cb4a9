// <opencv2/opencv.hpp> for sky_detector/imageSkyDetector.h in the node.cpp compile check.  TEST INFRASTRUCTURE.
#pragma once
#include <vector>
#include <opencv2/core/core.hpp>
#include <opencv2/highgui/highgui.hpp>
