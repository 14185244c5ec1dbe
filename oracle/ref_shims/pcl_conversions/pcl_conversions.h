// Stand-in for pcl_conversions/pcl_conversions.h.
#pragma once
