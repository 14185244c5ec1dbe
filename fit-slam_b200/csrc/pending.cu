// pending.cu -- entry points whose stage is not built yet return FITSLAM_E_INVALID ("not implemented").
// Each block is deleted when its stage file lands.
#include "ctx.h"
using namespace fitslam;
#define NOT_IMPL(c) return fail((c), FITSLAM_E_INVALID, std::string(__func__) + ": not implemented in this build")
namespace fitslam {
#ifndef HAVE_FRONTIER
void destroy_frontier_state(fitslam_ctx *) {}
#endif
#ifndef HAVE_FISHER
void destroy_fisher_state(fitslam_ctx *) {}
#endif
#ifndef HAVE_SELECT
void destroy_select_state(fitslam_ctx *) {}
#endif
#ifndef HAVE_CYCLE
void destroy_cycle_state(fitslam_ctx *) {}
#endif
}
extern "C" {
#ifndef HAVE_FRONTIER
fitslam_status fitslam_search_frontiers(fitslam_ctx *c, double, double, fitslam_frontier *, uint32_t *, uint32_t *, uint32_t *) { NOT_IMPL(c); }
fitslam_status fitslam_set_search_distance(fitslam_ctx *c, double) { NOT_IMPL(c); }
#endif
#ifndef HAVE_FISHER
fitslam_status fitslam_set_map_points(fitslam_ctx *c, const float *, uint64_t) { NOT_IMPL(c); }
fitslam_status fitslam_fisher_generate_lookup(fitslam_ctx *c, const char *) { NOT_IMPL(c); }
fitslam_status fitslam_fisher_load_lookup(fitslam_ctx *c, const char *) { NOT_IMPL(c); }
fitslam_status fitslam_fisher_lookup_dims(const fitslam_ctx *, int32_t *, int32_t *) { return FITSLAM_E_INVALID; }
fitslam_status fitslam_fisher_lookup_copy(fitslam_ctx *c, float *, uint64_t) { NOT_IMPL(c); }
fitslam_status fitslam_score_fisher(fitslam_ctx *c, const fitslam_pose *, uint32_t, float *, uint8_t *, uint32_t *) { NOT_IMPL(c); }
#endif
#ifndef HAVE_SELECT
fitslam_status fitslam_normalise_select(fitslam_ctx *c, const double *, const double *, const uint8_t *, uint32_t, double *, double *, double *, uint8_t *, uint32_t *, fitslam_selection *) { NOT_IMPL(c); }
#endif
#ifndef HAVE_CYCLE
fitslam_status fitslam_cycle(fitslam_ctx *c, double, double, double, const double *, int, fitslam_cycle_result *) { NOT_IMPL(c); }
fitslam_status fitslam_cycle_begin(fitslam_ctx *c, double, double, double, const double *, int, uint32_t *) { NOT_IMPL(c); }
fitslam_status fitslam_cycle_end(fitslam_ctx *c, fitslam_cycle_result *) { NOT_IMPL(c); }
fitslam_status fitslam_cycle_fetch(fitslam_ctx *c, fitslam_frontier *, fitslam_arrival *, double *, double *, double *, float *, uint32_t *, uint32_t) { NOT_IMPL(c); }
#endif
}
