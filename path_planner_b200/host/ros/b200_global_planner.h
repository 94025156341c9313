// N4: the ROS1 surface the north star names. A nav_core::BaseGlobalPlanner plugin
// that does what PathPlannerROS::goalSetCallback does with AstarHybrid::createPath
// (src/path_planner/src/path_planner_ros.cpp:110-168), on the GPU tables: costmap_2d grid -> pp_costmap_upload,
// footprint -> pp_footprint_from_boundary, plan -> pp_plan (SE2 volume + heuristic field on the device,
// search on the host). Parameters have the reference's names and defaults (path_planner_ros.cpp:19-52,
// config/params.yaml:28-45) under "~/<name>/collision" and "~/<name>/path_planner".
//
// ROS is not installed where this was written: the file is compiled and exercised against the
// stand-in headers of tests/ros_shim/ (tests/test_ros_adapter.py), never against a real ROS. The
// reference's Smoother (path_planner_ros.cpp:143) is outside the hot path and is not applied.
#pragma once
#include <costmap_2d/costmap_2d_ros.h>
#include <geometry_msgs/PoseStamped.h>
#include <nav_core/base_global_planner.h>
#include <ros/ros.h>

#include <string>
#include <vector>

#include "pp_b200.h"

namespace path_planner_b200
{
class B200GlobalPlanner : public nav_core::BaseGlobalPlanner
{
public:
    B200GlobalPlanner() = default;
    B200GlobalPlanner(std::string name, costmap_2d::Costmap2DROS *costmap_ros) { initialize(name, costmap_ros); }
    ~B200GlobalPlanner() override;

    void initialize(std::string name, costmap_2d::Costmap2DROS *costmap_ros) override;
    bool makePlan(const geometry_msgs::PoseStamped &start, const geometry_msgs::PoseStamped &goal,
                  std::vector<geometry_msgs::PoseStamped> &plan) override;

    // of the last makePlan: expanded nodes, and the two halves of its time
    int lastIterations() const { return _iterations; }
    double lastTablesMs() const { return _tables_ms; }
    double lastSearchMs() const { return _search_ms; }

private:
    bool _initialized = false;
    costmap_2d::Costmap2DROS *_costmap_ros = nullptr;
    pp_handle _pp = nullptr;
    pp_plan_params _params{};
    std::vector<float> _footprint;
    bool _hole = false;
    double _footprint_resolution = 0.0;  // resolution the device footprint was rasterised for
    std::vector<float> _path;
    int _iterations = 0;
    double _tables_ms = 0.0, _search_ms = 0.0;
};
}  // namespace path_planner_b200
