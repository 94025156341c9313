// See b200_global_planner.h.
#include "b200_global_planner.h"

#include <pluginlib/class_list_macros.h>
#include <tf/transform_datatypes.h>

#include <mutex>

PLUGINLIB_EXPORT_CLASS(path_planner_b200::B200GlobalPlanner, nav_core::BaseGlobalPlanner)

namespace path_planner_b200
{
B200GlobalPlanner::~B200GlobalPlanner()
{
    if (_pp)
        pp_destroy(_pp);
}

void B200GlobalPlanner::initialize(std::string name, costmap_2d::Costmap2DROS *costmap_ros)
{
    if (_initialized)
    {
        ROS_WARN("[B200GlobalPlanner] already initialized");
        return;
    }
    _costmap_ros = costmap_ros;
    ros::NodeHandle nh("~/" + name);
    // the reference's parameters, names and defaults: path_planner_ros.cpp:19-52
    ros::NodeHandle collision(nh, "collision");
    collision.param<bool>("hole", _hole, false);
    collision.param<std::vector<float>>("footprint", _footprint, {1.0f});
    ros::NodeHandle pp(nh, "path_planner");
    pp.param<float>("minimum_turning_radius", _params.minimum_turning_radius, 0.5f);
    pp.param<float>("cost_penalty", _params.cost_penalty, 2.0f);
    pp.param<float>("change_penalty", _params.change_penalty, 0.0f);
    pp.param<float>("non_straight_penalty", _params.non_straight_penalty, 1.2f);
    pp.param<float>("reverse_penalty", _params.reverse_penalty, 2.1f);
    pp.param<float>("analytic_expansion_ratio", _params.analytic_expansion_ratio, 1.0f);
    pp.param<float>("analytic_expansion_tolerance", _params.analytic_expansion_tolerance, 3.0f);
    pp.param<float>("tolerance", _params.tolerance, 0.5f);
    int bins, max_iterations, max_on_approach_iterations, device;
    pp.param<int>("angle_quantization_bins", bins, 72);
    pp.param<int>("max_iterations", max_iterations, 100000);
    pp.param<int>("max_on_approach_iterations", max_on_approach_iterations, 1000);
    std::string motion_type;
    pp.param<std::string>("motion_type", motion_type, "ReedsShepp");
    nh.param<int>("device", device, 0);
    bool exact_analytic;
    nh.param<bool>("exact_analytic_expansion", exact_analytic, true);  // continuous-pose checks of analytic expansions, like the reference
    _params.flags = exact_analytic ? PP_PLAN_EXACT_ANALYTIC : 0u;
    _params.n_theta = static_cast<uint32_t>(bins);
    _params.max_iterations = static_cast<uint32_t>(max_iterations);
    _params.max_on_approach_iterations = static_cast<uint32_t>(max_on_approach_iterations);
    if (motion_type == "Dubins")
        _params.motion_type = 1;
    else if (motion_type == "ReedsShepp")
        _params.motion_type = 2;
    else
    {
        ROS_ERROR("[B200GlobalPlanner] Invalid motion type: %s", motion_type.c_str());
        return;
    }
    if (_footprint.size() != 1 && _footprint.size() != 2 && _footprint.size() != 4)
    {
        ROS_ERROR("[B200GlobalPlanner] collision/footprint needs 1, 2 or 4 numbers (collision_checker.cpp:3-81)");
        return;
    }
    if (pp_create(&_pp, device) != PP_OK)
    {
        ROS_ERROR("[B200GlobalPlanner] no CUDA device %d: this planner has no CPU path", device);
        _pp = nullptr;
        return;
    }
    _initialized = true;
}

bool B200GlobalPlanner::makePlan(const geometry_msgs::PoseStamped &start, const geometry_msgs::PoseStamped &goal,
                                 std::vector<geometry_msgs::PoseStamped> &plan)
{
    plan.clear();
    if (!_initialized)
    {
        ROS_ERROR("[B200GlobalPlanner] makePlan before initialize");
        return false;
    }
    costmap_2d::Costmap2D *costmap = _costmap_ros->getCostmap();
    double resolution, origin_x, origin_y;
    {
        // Costmap2DROS updates the grid on its own thread: copy it out under the costmap's lock
        // (pp_costmap_upload copies into its pinned staging buffer before it returns)
        std::lock_guard<costmap_2d::Costmap2D::mutex_t> lock(*(costmap->getMutex()));
        resolution = costmap->getResolution();
        origin_x = costmap->getOriginX();
        origin_y = costmap->getOriginY();
        if (pp_costmap_upload(_pp, costmap->getCharMap(), costmap->getSizeInCellsX(), costmap->getSizeInCellsY(), 0) != PP_OK)
        {
            ROS_ERROR("[B200GlobalPlanner] %s", pp_last_error(_pp));
            return false;
        }
    }
    if (resolution != _footprint_resolution)
    {
        if (pp_footprint_from_boundary(_pp, resolution, _footprint.data(), static_cast<int>(_footprint.size()), _hole) != PP_OK)
        {
            ROS_ERROR("[B200GlobalPlanner] %s", pp_last_error(_pp));
            return false;
        }
        _footprint_resolution = resolution;
    }
    // poses truncated to centimetres, as PathPlannerROS::goalSetCallback does (path_planner_ros.cpp:121-127)
    auto cm = [](double v) { return static_cast<int>(v * 100) / 100.0f; };
    const float s[3] = {cm(start.pose.position.x), cm(start.pose.position.y), cm(tf::getYaw(start.pose.orientation))};
    const float g[3] = {cm(goal.pose.position.x), cm(goal.pose.position.y), cm(tf::getYaw(goal.pose.orientation))};

    const uint32_t max_points = 65536;
    _path.resize(static_cast<size_t>(3) * max_points);
    uint32_t n = 0;
    int32_t found = 0, iterations = 0;
    if (pp_plan(_pp, &_params, resolution, origin_x, origin_y, s, g, _path.data(), max_points, &n, &iterations, &found, &_tables_ms,
                &_search_ms) != PP_OK)
    {
        ROS_ERROR("[B200GlobalPlanner] %s", pp_last_error(_pp));
        return false;
    }
    _iterations = iterations;
    if (!found)
    {
        ROS_WARN("[B200GlobalPlanner] No path found");
        return false;
    }
    if (n > max_points)  // pp_plan reports the full length and writes at most max_points poses
    {
        ROS_ERROR("[B200GlobalPlanner] path of %u poses does not fit the %u-pose buffer", n, max_points);
        return false;
    }
    // the search returns goal -> start; the reference reverses it (path_planner_ros.cpp:140)
    const ros::Time stamp = ros::Time::now();
    plan.reserve(n);
    for (uint32_t i = n; i-- > 0;)
    {
        geometry_msgs::PoseStamped pose;
        pose.header.frame_id = _costmap_ros->getGlobalFrameID();
        pose.header.stamp = stamp;
        pose.pose.position.x = _path[3 * i];
        pose.pose.position.y = _path[3 * i + 1];
        pose.pose.position.z = 0;
        pose.pose.orientation = tf::createQuaternionMsgFromYaw(_path[3 * i + 2]);
        plan.push_back(pose);
    }
    return true;
}
}  // namespace path_planner_b200
