// Stand-in for sensor_msgs/msg/point_cloud2.hpp.
#pragma once
namespace sensor_msgs { namespace msg { struct PointCloud2 {}; } }
