// TEST INFRASTRUCTURE ONLY (oracle). Stand-in for ROS costmap_2d/cost_values.h, which is an
// external dependency of the reference (README.md:15) and not present under /root/reference.
// Values are the ones the reference relies on at collision_checker.cpp:138,163,165 and
// teb_follower.cpp:152.
#pragma once
namespace costmap_2d
{
static const unsigned char NO_INFORMATION = 255;
static const unsigned char LETHAL_OBSTACLE = 254;
static const unsigned char INSCRIBED_INFLATED_OBSTACLE = 253;
static const unsigned char FREE_SPACE = 0;
}
