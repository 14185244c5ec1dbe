// Shadows roadmap_explorer/Parameters.hpp (a singleton fed by a lifecycle node) for the oracle/_ref build: the same
// `parameterInstance.getValue<T>(key)` call, backed by a map the harness fills.
#pragma once
#include <map>
#include <stdexcept>
#include <string>
class RefParametersRE {
public:
  static RefParametersRE &instance() { static RefParametersRE p; return p; }
  void set(const std::string &key, double v) { values_[key] = v; }
  template <typename T> T getValue(const std::string &key) const {
    auto it = values_.find(key);
    if (it == values_.end()) throw std::runtime_error("Parameter " + key + " is not found in the map");
    return static_cast<T>(it->second);
  }
private:
  std::map<std::string, double> values_;
};
#define parameterInstance (RefParametersRE::instance())
