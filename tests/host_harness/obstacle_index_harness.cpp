// TEST INFRASTRUCTURE ONLY. Drives path_planner_b200/host/obstacle_index.h (the C++ mirror of the TEB
// follower's obstacle association, INTEGRATION.md sections 3 and 3c) the way TebOptimizer would.
#include "obstacle_index.h"

#include <cstring>

extern "C" int harness_obstacle_index(int device, const double *points_xy, uint32_t n, const double *poses_xyt, uint32_t q,
                                      double radius_arg, uint32_t k, const double *robot_numbers, uint32_t n_robot,
                                      uint32_t *idx_out, double *d2_out, uint32_t *cnt_out, double *err_out, double *jac_out,
                                      double *radii_out)
{
    pp_handle h = nullptr;
    if (pp_create(&h, device) != PP_OK)
        return 0;
    int ok = 1;
    {
        ObstacleIndex index(h);
        index.setPoints(points_xy, n);
        std::vector<double> xy(2 * static_cast<size_t>(q));
        for (uint32_t i = 0; i < q; ++i)
        {
            xy[2 * i] = poses_xyt[3 * i];
            xy[2 * i + 1] = poses_xyt[3 * i + 1];
        }
        ok = ok && index.query(xy.data(), q, radius_arg, k);
        ok = ok && index.setRobotModel(robot_numbers, n_robot);
        const pp_obstacle_edge_params params = {0.1, 0.05, 0.2, 0.1};  // commons.h:219-222
        ok = ok && index.evaluateEdges(poses_xyt, params);
        radii_out[0] = index.inscribedRadius();
        radii_out[1] = index.circumRadius();
        for (uint32_t i = 0; ok && i < q; ++i)
        {
            cnt_out[i] = index.count(i);
            for (uint32_t j = 0; j < index.count(i); ++j)
            {
                idx_out[i * k + j] = index.index(i, j);
                d2_out[i * k + j] = index.dist2(i, j);
                std::memcpy(err_out + (i * k + j) * 2, index.edgeError(i, j), 2 * sizeof(double));
                std::memcpy(jac_out + (i * k + j) * 6, index.edgeJacobian(i, j), 6 * sizeof(double));
            }
        }
    }
    pp_destroy(h);
    return ok;
}
