// TEST INFRASTRUCTURE ONLY. Stand-in for <ros/ros.h> (ROS is not installed here): the few names the
// nav_core adapter uses. Parameters come from a process-wide table the test harness fills
// (ros_shim::params()), keyed by the resolved name, e.g. "~/planner/path_planner/max_iterations".
#pragma once
#include <cstdio>
#include <map>
#include <string>
#include <variant>
#include <vector>

#define ROS_ERROR(...) (std::fprintf(stderr, "[ERROR] " __VA_ARGS__), std::fputc('\n', stderr))
#define ROS_WARN(...) (std::fprintf(stderr, "[WARN] " __VA_ARGS__), std::fputc('\n', stderr))
#define ROS_INFO(...) (std::fprintf(stderr, "[INFO] " __VA_ARGS__), std::fputc('\n', stderr))

namespace ros_shim
{
using Value = std::variant<bool, int, double, std::string, std::vector<float>>;
inline std::map<std::string, Value> &params()
{
    static std::map<std::string, Value> table;
    return table;
}
}  // namespace ros_shim

namespace ros
{
struct Time
{
    double sec = 0.0;
    static Time now() { return Time{}; }
};

class NodeHandle
{
public:
    explicit NodeHandle(const std::string &ns = "") : _ns(ns) {}
    NodeHandle(const NodeHandle &parent, const std::string &ns) : _ns(parent._ns + "/" + ns) {}

    template <class T>
    bool param(const std::string &name, T &out, const T &fallback) const
    {
        auto it = ros_shim::params().find(_ns + "/" + name);
        if (it == ros_shim::params().end())
        {
            out = fallback;
            return false;
        }
        if constexpr (std::is_same_v<T, float>)
            out = static_cast<float>(std::get<double>(it->second));
        else
            out = std::get<T>(it->second);
        return true;
    }

private:
    std::string _ns;
};
}  // namespace ros
