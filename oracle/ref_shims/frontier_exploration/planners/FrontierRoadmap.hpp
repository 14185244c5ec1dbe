// Stand-in: nothing of the roadmap planner is used by the sources oracle/_ref compiles; the real header brings in the helpers and
// the parameter singleton, which FisherInfoManager.cpp relies on.
#pragma once
#include <frontier_exploration/Frontier.hpp>
#include <frontier_exploration/Helpers.hpp>
#include <frontier_exploration/Parameters.hpp>
