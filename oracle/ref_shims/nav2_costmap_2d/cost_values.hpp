// Stand-in for nav2_costmap_2d/cost_values.hpp (ROS 2 Humble): the published cost constants.
#pragma once
namespace nav2_costmap_2d {
static constexpr unsigned char NO_INFORMATION = 255;
static constexpr unsigned char LETHAL_OBSTACLE = 254;
static constexpr unsigned char INSCRIBED_INFLATED_OBSTACLE = 253;
static constexpr unsigned char MAX_NON_OBSTACLE = 252;
static constexpr unsigned char FREE_SPACE = 0;
}  // namespace nav2_costmap_2d
