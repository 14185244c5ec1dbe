// Stand-in for nav2_costmap_2d::Costmap2D (ROS 2 Humble, package ros-humble-nav2-costmap-2d; not in /root/reference, see
// SURVEY.md 8c).  Only the members the reference's frontier search and ray helpers call, with the published semantics of
// costmap_2d.cpp: cell centres at origin + (m + 0.5) * resolution, worldToMap truncates and rejects coordinates below the
// origin or at / beyond the size, row-major indices.  Test infrastructure for oracle/_ref, never shipped.
#pragma once
#include <algorithm>
#include <cstring>
#include <queue>
#include <vector>
#include "nav2_costmap_2d/cost_values.hpp"
namespace nav2_costmap_2d {
struct MapLocation {
  unsigned int x;
  unsigned int y;
};
class Costmap2D {
public:
  Costmap2D(unsigned int cells_size_x, unsigned int cells_size_y, double resolution, double origin_x, double origin_y,
            unsigned char default_value = 0)
      : size_x_(cells_size_x), size_y_(cells_size_y), resolution_(resolution), origin_x_(origin_x), origin_y_(origin_y),
        costmap_((size_t)cells_size_x * cells_size_y, default_value) {}
  unsigned char getCost(unsigned int mx, unsigned int my) const { return costmap_[getIndex(mx, my)]; }
  unsigned char getCost(unsigned int index) const { return costmap_[index]; }
  void setCost(unsigned int mx, unsigned int my, unsigned char cost) { costmap_[getIndex(mx, my)] = cost; }
  void mapToWorld(unsigned int mx, unsigned int my, double &wx, double &wy) const {
    wx = origin_x_ + (mx + 0.5) * resolution_;
    wy = origin_y_ + (my + 0.5) * resolution_;
  }
  bool worldToMap(double wx, double wy, unsigned int &mx, unsigned int &my) const {
    if (wx < origin_x_ || wy < origin_y_) return false;
    mx = static_cast<unsigned int>((wx - origin_x_) / resolution_);
    my = static_cast<unsigned int>((wy - origin_y_) / resolution_);
    if (mx < size_x_ && my < size_y_) return true;
    return false;
  }
  inline unsigned int getIndex(unsigned int mx, unsigned int my) const { return my * size_x_ + mx; }
  inline void indexToCells(unsigned int index, unsigned int &mx, unsigned int &my) const {
    my = index / size_x_;
    mx = index - (my * size_x_);
  }
  unsigned char *getCharMap() const { return const_cast<unsigned char *>(costmap_.data()); }
  unsigned int getSizeInCellsX() const { return size_x_; }
  unsigned int getSizeInCellsY() const { return size_y_; }
  double getSizeInMetersX() const { return (size_x_ - 1 + 0.5) * resolution_; }
  double getSizeInMetersY() const { return (size_y_ - 1 + 0.5) * resolution_; }
  double getOriginX() const { return origin_x_; }
  double getOriginY() const { return origin_y_; }
  double getResolution() const { return resolution_; }

protected:
  unsigned int size_x_, size_y_;
  double resolution_, origin_x_, origin_y_;
  std::vector<unsigned char> costmap_;
};
}  // namespace nav2_costmap_2d
