// TEST INFRASTRUCTURE ONLY. Stand-in for <geometry_msgs/PoseStamped.h>.
#pragma once
#include <ros/ros.h>
#include <string>

namespace std_msgs
{
struct Header
{
    unsigned int seq = 0;
    ros::Time stamp;
    std::string frame_id;
};
}  // namespace std_msgs

namespace geometry_msgs
{
struct Point
{
    double x = 0, y = 0, z = 0;
};
struct Quaternion
{
    double x = 0, y = 0, z = 0, w = 1;
};
struct Pose
{
    Point position;
    Quaternion orientation;
};
struct PoseStamped
{
    std_msgs::Header header;
    Pose pose;
};
}  // namespace geometry_msgs
