// Stand-in for nav_msgs/msg/occupancy_grid.hpp: the fields SensorSimulator::markRay reads and writes.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>
#include "geometry_msgs/msg/pose.hpp"
namespace nav_msgs { namespace msg {
struct OccupancyGrid {
  using SharedPtr = std::shared_ptr<OccupancyGrid>;
  struct { std::string frame_id; } header;
  struct {
    float resolution = 0;
    uint32_t width = 0, height = 0;
    geometry_msgs::msg::Pose origin;
  } info;
  std::vector<int8_t> data;
};
} }
