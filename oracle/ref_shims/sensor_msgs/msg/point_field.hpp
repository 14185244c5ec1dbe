// Stand-in for sensor_msgs/msg/point_field.hpp.
#pragma once
