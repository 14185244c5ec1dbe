// Stand-in for tf2's Vector3 (double coordinates): construction and the three getters the Fisher manager uses.
#pragma once
namespace tf2 {
class Vector3 {
public:
  Vector3() : v_{0, 0, 0} {}
  Vector3(double x, double y, double z) : v_{x, y, z} {}
  const double &x() const { return v_[0]; }
  const double &y() const { return v_[1]; }
  const double &z() const { return v_[2]; }
private:
  double v_[3];
};
}  // namespace tf2
