// Stand-in for pcl/point_types.h.
#pragma once
namespace pcl { struct PointXYZI { float x, y, z, intensity; }; }
