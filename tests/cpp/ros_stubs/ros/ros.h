// Minimal ROS stand-in: to COMPILE-CHECK the reference's unmodified node.cpp against this repository's Solver / SGM / BM
// headers on a box without ROS (tests/test_host_shell.py), and to LINK AND RUN it with the reference's own solver as the
// oracle of the point-cloud loop (oracle/node_harness.cpp).  Integer parameters can be set through the environment:
// nh.param("/sgm_node/img_w", ...) reads ROS_PARAM__sgm_node_img_w.  TEST INFRASTRUCTURE.
#pragma once
#include <cstdlib>
#include <string>
namespace ros {
inline void init(int&, char**, const std::string&) {}
inline void spin() {}
class Publisher {};
class NodeHandle {
public:
    explicit NodeHandle(const std::string& = std::string()) {}
    template <class M>
    Publisher advertise(const std::string&, int) { return Publisher(); }
    template <class T>
    bool param(const std::string&, T& value, const T& fallback) {
        value = fallback;
        return false;
    }
    bool param(const std::string& name, int& value, const int& fallback) {
        std::string key = "ROS_PARAM_" + name;
        for (char& c : key)
            if (c == '/') c = '_';
        const char* v = std::getenv(key.c_str());
        value = v ? std::atoi(v) : fallback;
        return v != nullptr;
    }
};
}  // namespace ros
