// Stand-in for pcl/point_cloud.h.
#pragma once
namespace pcl { template <typename T> class PointCloud {}; }
