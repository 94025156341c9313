// TEST INFRASTRUCTURE ONLY (oracle). Stand-in for the OMPL types the reference planner uses
// (motion_hybrid.h:4-7,36; astar_hybrid.cpp:84-91,297-318). OMPL is an un-vendored, un-pinned
// dependency of the reference (path_planner/CMakeLists.txt:25, README.md:16 -> Ubuntu 20.04
// libompl-dev, OMPL 1.4.2 inferred): parity for anything computed in here is UNPINNED. The SE2
// state is three doubles (x, y, yaw). The standard headers below are the ones astar_hybrid.h
// uses without including.
#pragma once
#include <memory>
#include <vector>
#include <queue>
#include <unordered_map>
#include <stdexcept>
#include <cmath>
#include <limits>
#include <algorithm>

namespace ompl
{
namespace base
{
struct State
{
    double v[3] = {0.0, 0.0, 0.0};
    double getX() const { return v[0]; }
    double getY() const { return v[1]; }
    double getYaw() const { return v[2]; }
};

class StateSpace
{
public:
    virtual ~StateSpace() = default;
    virtual double distance(const State *a, const State *b) const = 0;
    virtual void interpolate(const State *from, const State *to, double t, State *out) const = 0;
};
using StateSpacePtr = std::shared_ptr<StateSpace>;

// yaw wrap of SO2StateSpace::enforceBounds
inline double pp_wrap_yaw(double yaw)
{
    const double pi = 3.14159265358979323846;
    double v = std::fmod(yaw, 2.0 * pi);
    if (v < -pi)
        v += 2.0 * pi;
    else if (v >= pi)
        v -= 2.0 * pi;
    return v;
}
}
}
