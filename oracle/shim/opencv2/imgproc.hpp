// Stand-in for <opencv2/imgproc.hpp>: GaussianBlur / cvtColor live in the core shim.
// Test infrastructure only.
#pragma once
#include "core/core.hpp"
