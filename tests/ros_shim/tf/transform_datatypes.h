// TEST INFRASTRUCTURE ONLY. Stand-in for <tf/transform_datatypes.h>: the two yaw helpers, with the
// formulas of tf (yaw of a quaternion; quaternion of a rotation about z).
#pragma once
#include <cmath>
#include <geometry_msgs/PoseStamped.h>

namespace tf
{
inline double getYaw(const geometry_msgs::Quaternion &q)
{
    return std::atan2(2.0 * (q.w * q.z + q.x * q.y), 1.0 - 2.0 * (q.y * q.y + q.z * q.z));
}
inline geometry_msgs::Quaternion createQuaternionMsgFromYaw(double yaw)
{
    geometry_msgs::Quaternion q;
    q.x = 0.0;
    q.y = 0.0;
    q.z = std::sin(0.5 * yaw);
    q.w = std::cos(0.5 * yaw);
    return q;
}
}  // namespace tf
