// see ros/ros.h in this directory.  TEST INFRASTRUCTURE.
#pragma once
#include <string>
#include <ros/ros.h>
namespace image_transport {
class Publisher {};
class ImageTransport {
public:
    explicit ImageTransport(const ros::NodeHandle&) {}
    Publisher advertise(const std::string&, int) { return Publisher(); }
};
}  // namespace image_transport
