// see ros/ros.h in this directory.  TEST INFRASTRUCTURE.
#pragma once
namespace sensor_msgs {
struct PointCloud {};
}  // namespace sensor_msgs
