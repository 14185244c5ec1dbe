// Shadows the reference's planners/theta_star.hpp (Theta* over the costmap: out of scope) for the oracle/_ref build.
#pragma once
