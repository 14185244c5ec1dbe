// Stand-in for nav2_navfn_planner/navfn.hpp: named by roadmap_explorer/Helpers.hpp, not used by what oracle/_ref compiles.
#pragma once
namespace nav2_navfn_planner { class NavFn {}; }
