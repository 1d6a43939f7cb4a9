// Minimal ROS stand-in, only to COMPILE-CHECK the reference's unmodified node.cpp against this repository's
// Solver / SGM / BM headers on a box without ROS (tests/test_host_shell.py).  Nothing here runs.  TEST INFRASTRUCTURE.
#pragma once
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
};
}  // namespace ros
