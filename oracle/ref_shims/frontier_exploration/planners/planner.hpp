// Shadows the reference's planners/planner.hpp (A* over the costmap: out of scope) for the oracle/_ref build.
#pragma once
