// TEST INFRASTRUCTURE ONLY. C entry points around the reference's OWN obstacle-edge code
// (src/path_follower/src/teb_path_follower/robot_model.cpp, commons.cpp,
// g2o_edges/edge_obstacle.hpp), compiled unmodified from /root/reference by oracle/Makefile against
// the Eigen / g2o stand-ins in oracle/shim_teb/ into oracle/_ref/libpp_teb_ref.so.
#include "g2o_edges/edge_obstacle.hpp"
#include "robot_model.h"

#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

extern "C" {

// RobotModelHelper::createRobotModel (robot_model.cpp:68-95) and the radii it computes (:15-37)
int teb_ref_robot_radii(const double *numbers, int n, double *inscribed, double *circum)
{
    std::vector<double> v(numbers, numbers + n);
    std::unique_ptr<RobotModel> robot(RobotModelHelper::createRobotModel(v));
    *inscribed = robot->inscribedRadius();
    *circum = robot->circumRadius();
    return 0;
}

// One EdgeObstacle per (pose q, slot k < cnt[q]) as TebOptimizer::addObstacleEdges builds them
// (teb_optimizer.cpp:283-313), then computeError() (edge_obstacle.hpp:19-26) and the numeric
// linearizeOplus(). err_out: Q*K*2, jac_out: Q*K*6 (row-major 2x3); unused slots are zero.
// info4 = obstacle_collision_dist, margin_obstacle_collision, obstacle_inflation_dist,
// margin_obstacle_inflation (commons.h:219-222).
int teb_ref_obstacle_edges(const double *poses_xyt, uint32_t Q, const double *pts_xy, const uint32_t *idx,
                           const uint32_t *cnt, uint32_t K, const double *robot_numbers, int n_robot,
                           const double *info4, double *err_out, double *jac_out)
{
    std::vector<double> v(robot_numbers, robot_numbers + n_robot);
    std::unique_ptr<RobotModel> robot(RobotModelHelper::createRobotModel(v));
    FollowerInfo info;
    info.obstacle_collision_dist = info4[0];
    info.margin_obstacle_collision = info4[1];
    info.obstacle_inflation_dist = info4[2];
    info.margin_obstacle_inflation = info4[3];
    std::memset(err_out, 0, sizeof(double) * Q * K * 2);
    if (jac_out)
        std::memset(jac_out, 0, sizeof(double) * Q * K * 6);
    for (uint32_t q = 0; q < Q; ++q)
    {
        Vertex2E vertex(poses_xyt[3 * q], poses_xyt[3 * q + 1], poses_xyt[3 * q + 2]);
        for (uint32_t k = 0; k < cnt[q] && k < K; ++k)
        {
            const uint32_t i = idx[q * K + k];
            Point2D obstacle(pts_xy[2 * i], pts_xy[2 * i + 1]);
            EdgeObstacle edge(&obstacle);
            edge.setInfoParameters(robot.get(), &info);
            edge.setVertex(0, &vertex);
            edge.computeError();
            err_out[(q * K + k) * 2] = edge.error()[0];
            err_out[(q * K + k) * 2 + 1] = edge.error()[1];
            if (jac_out)
            {
                edge.linearizeOplus();
                std::memcpy(jac_out + (q * K + k) * 6, edge.jacobianOplusXi(), sizeof(double) * 6);
            }
        }
    }
    return 0;
}
}
