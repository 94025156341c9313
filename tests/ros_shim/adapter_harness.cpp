// TEST INFRASTRUCTURE ONLY. Drives path_planner_b200::B200GlobalPlanner through the stand-in ROS types the
// way move_base drives a global planner plugin: initialize(name, costmap_ros), then makePlan(start, goal, plan).
#include "b200_global_planner.h"

#include <tf/transform_datatypes.h>

#include <cstring>

extern "C" int adapter_make_plan(const unsigned char *cost, unsigned W, unsigned H, double res, double origin_x, double origin_y,
                                 const float *boundary, int nb, int hole, const char *motion_type, int max_iterations,
                                 const double *start_xyt, const double *goal_xyt, double *plan_xyyaw_out, int max_poses, int *n_out,
                                 int *iterations_out, char *frame_out, int frame_cap)
{
    auto &p = ros_shim::params();
    p.clear();
    // src/path_planner/config/params.yaml:28-45
    p["~/planner/collision/hole"] = static_cast<bool>(hole);
    p["~/planner/collision/footprint"] = std::vector<float>(boundary, boundary + nb);
    p["~/planner/path_planner/minimum_turning_radius"] = 0.84;
    p["~/planner/path_planner/cost_penalty"] = 2.0;
    p["~/planner/path_planner/change_penalty"] = 0.01;
    p["~/planner/path_planner/non_straight_penalty"] = 1.2;
    p["~/planner/path_planner/reverse_penalty"] = 3.0;
    p["~/planner/path_planner/analytic_expansion_ratio"] = 1.0;
    p["~/planner/path_planner/analytic_expansion_tolerance"] = 3.0;
    p["~/planner/path_planner/tolerance"] = 0.6;
    p["~/planner/path_planner/max_iterations"] = max_iterations;
    p["~/planner/path_planner/motion_type"] = std::string(motion_type);

    costmap_2d::Costmap2D grid(W, H, res, origin_x, origin_y);
    std::memcpy(grid.getCharMap(), cost, static_cast<size_t>(W) * H);
    costmap_2d::Costmap2DROS costmap_ros(&grid, "map");
    path_planner_b200::B200GlobalPlanner planner;
    nav_core::BaseGlobalPlanner *plugin = &planner;  // through the interface, as move_base holds it
    plugin->initialize("planner", &costmap_ros);

    geometry_msgs::PoseStamped start, goal;
    start.pose.position.x = start_xyt[0], start.pose.position.y = start_xyt[1];
    start.pose.orientation = tf::createQuaternionMsgFromYaw(start_xyt[2]);
    goal.pose.position.x = goal_xyt[0], goal.pose.position.y = goal_xyt[1];
    goal.pose.orientation = tf::createQuaternionMsgFromYaw(goal_xyt[2]);
    std::vector<geometry_msgs::PoseStamped> plan;
    const bool ok = plugin->makePlan(start, goal, plan);
    *n_out = static_cast<int>(plan.size());
    *iterations_out = planner.lastIterations();
    for (int i = 0; i < *n_out && i < max_poses; ++i)
    {
        plan_xyyaw_out[3 * i] = plan[i].pose.position.x;
        plan_xyyaw_out[3 * i + 1] = plan[i].pose.position.y;
        plan_xyyaw_out[3 * i + 2] = tf::getYaw(plan[i].pose.orientation);
    }
    if (!plan.empty())
        std::snprintf(frame_out, frame_cap, "%s", plan.front().header.frame_id.c_str());
    return ok;
}
