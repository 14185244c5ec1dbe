// Stand-in for nav2_costmap_2d/costmap_2d_ros.hpp (the real header brings <cfloat> in; theta_star.hpp relies on that).
#pragma once
#include <cfloat>
#include <nav2_costmap_2d/costmap_2d.hpp>
namespace nav2_costmap_2d { class Costmap2DROS {}; }
