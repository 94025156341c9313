// TEST INFRASTRUCTURE ONLY. Stand-in for <nav_core/base_global_planner.h> (ROS navigation): the
// interface as nav_core declares it.
#pragma once
#include <costmap_2d/costmap_2d_ros.h>
#include <geometry_msgs/PoseStamped.h>
#include <string>
#include <vector>

namespace nav_core
{
class BaseGlobalPlanner
{
public:
    virtual bool makePlan(const geometry_msgs::PoseStamped &start, const geometry_msgs::PoseStamped &goal,
                          std::vector<geometry_msgs::PoseStamped> &plan) = 0;
    virtual void initialize(std::string name, costmap_2d::Costmap2DROS *costmap_ros) = 0;
    virtual ~BaseGlobalPlanner() {}

protected:
    BaseGlobalPlanner() {}
};
}  // namespace nav_core
