// Batched replacement for the obstacle association of the TEB follower: the KDTree2D member
// (teb_path_follower/commons.h:199, built at teb_optimizer.h:29) and the per-pose loop
// radiusSearch + keep-first-K at teb_optimizer.cpp:294-311. Header-only over the C ABI.
//
//   ObstacleIndex idx(handle);
//   idx.setPoints(points_xy, n);            // or idx.extractFromCostmap(...) (teb_follower.cpp:140-169)
//   idx.query(poses_xy, q, radius_arg, K);   // one launch for all poses of a buildGraph round
//   idx.count(i), idx.index(i, j), idx.dist2(i, j)  -> what results[j].first / .second were
//   idx.setRobotModel(...); idx.evaluateEdges(poses_xyt, params);   // per optimiser iteration (N3)
//   idx.edgeError(i, j), idx.edgeJacobian(i, j)
//
// radius_arg is passed through exactly as the reference passes it to nanoflann (compared against
// squared distances, strict <).
#pragma once
#include "pp_b200.h"

#include <cstdint>
#include <vector>

class ObstacleIndex
{
public:
    explicit ObstacleIndex(pp_handle h) : _h(h), _n(0), _q(0), _k(0) {}

    bool setPoints(const double *points_xy, uint32_t n)
    {
        _points.assign(points_xy, points_xy + 2 * static_cast<size_t>(n));
        _n = n;
        _resident = false;
        return true;
    }
    // Candidate extraction on the GPU; the points also come back to the host because EdgeObstacle
    // keeps pointers into the point vector (teb_optimizer.cpp:305-306).
    bool extractFromCostmap(double resolution, double origin_x, double origin_y, double robot_x, double robot_y,
                            double robot_theta, double max_follow_range, double behind_ignore_range)
    {
        uint32_t n = 0;
        if (pp_obstacles_extract(_h, resolution, origin_x, origin_y, robot_x, robot_y, robot_theta, max_follow_range,
                                 behind_ignore_range, nullptr, 0, &n) != PP_OK)
            return false;
        _points.resize(2 * static_cast<size_t>(n));
        if (n && pp_obstacles_extract(_h, resolution, origin_x, origin_y, robot_x, robot_y, robot_theta, max_follow_range,
                                      behind_ignore_range, _points.data(), n, &n) != PP_OK)
            return false;
        _n = n;
        _resident = true;
        return true;
    }
    bool query(const double *poses_xy, uint32_t q, double radius_arg, uint32_t k)
    {
        _q = q;
        _k = k;
        _idx.assign(static_cast<size_t>(q) * k, 0);
        _d2.assign(static_cast<size_t>(q) * k, 0.0);
        _cnt.assign(q, 0);
        if (_n == 0 || q == 0)
            return true;
        const pp_status s = pp_radius_topk(_h, _resident ? nullptr : _points.data(), _n, poses_xy, q, radius_arg, k, _idx.data(),
                                           _d2.data(), _cnt.data());
        _resident = s == PP_OK;
        return s == PP_OK;
    }
    // N3: what EdgeObstacle::computeError (g2o_edges/edge_obstacle.hpp:19-26) and g2o's numeric
    // linearizeOplus give for every edge (pose i, kept obstacle j) of the last query(), in one launch.
    // robot_numbers as RobotModelHelper::createRobotModel takes them (robot_model.cpp:68-95).
    bool setRobotModel(const double *robot_numbers, uint32_t n)
    {
        return pp_robot_model_set(_h, robot_numbers, n, &_inscribed, &_circum) == PP_OK;
    }
    double inscribedRadius() const { return _inscribed; }
    double circumRadius() const { return _circum; }
    bool evaluateEdges(const double *poses_xyt, const pp_obstacle_edge_params &params, bool jacobians = true)
    {
        _err.assign(static_cast<size_t>(_q) * _k * 2, 0.0);
        _jac.assign(jacobians ? static_cast<size_t>(_q) * _k * 6 : 0, 0.0);
        if (_n == 0 || _q == 0)
            return true;
        return pp_obstacle_edges(_h, poses_xyt, _q, _k, &params, _err.data(), jacobians ? _jac.data() : nullptr) == PP_OK;
    }
    // _error[0..1] of the edge and its 2x3 row-major Jacobian with respect to (x, y, theta)
    const double *edgeError(uint32_t pose, uint32_t j) const { return &_err[(static_cast<size_t>(pose) * _k + j) * 2]; }
    const double *edgeJacobian(uint32_t pose, uint32_t j) const { return &_jac[(static_cast<size_t>(pose) * _k + j) * 6]; }

    uint32_t size() const { return _n; }
    const std::vector<double> &points() const { return _points; }
    uint32_t count(uint32_t pose) const { return _cnt[pose]; }
    uint32_t index(uint32_t pose, uint32_t j) const { return _idx[static_cast<size_t>(pose) * _k + j]; }
    double dist2(uint32_t pose, uint32_t j) const { return _d2[static_cast<size_t>(pose) * _k + j]; }

private:
    pp_handle _h;
    std::vector<double> _points;
    uint32_t _n, _q, _k;
    bool _resident = false;
    std::vector<uint32_t> _idx, _cnt;
    std::vector<double> _d2, _err, _jac;
    double _inscribed = 0.0, _circum = 0.0;
};
