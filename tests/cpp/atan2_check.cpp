// atan2_check.cpp -- the host build of fit-slam_b200/csrc/atan2_cr.h against this machine's libm (tests/test_atan2.py).
// Prints: calls, disagreements, disagreements where `risky` was clear (must be 0), risky calls.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <random>

#include "atan2_cr.h"

static uint64_t bits(double d) { uint64_t u; memcpy(&u, &d, 8); return u; }

int main(int argc, char **argv) {
  const long n_each = argc > 1 ? atol(argv[1]) : 1000000;
  std::mt19937_64 rng(20261104);
  std::uniform_int_distribution<int> di(-40, 40);
  std::uniform_real_distribution<double> ur(-1, 1), ue(-60, 60);
  long n = 0, bad = 0, bad_unflagged = 0, n_risky = 0;
  auto check = [&](double y, double x) {
    bool risky = false;
    const double a = atan2(y, x), b = atan2cr::atan2_cr(y, x, &risky);
    ++n;
    n_risky += risky;
    if (bits(a) != bits(b) && !(a != a && b != b)) {
      ++bad;
      if (!risky) {
        if (bad_unflagged < 5) printf("unflagged disagreement: y=%a x=%a libm=%a cr=%a\n", y, x, a, b);
        ++bad_unflagged;
      }
    }
  };
  for (long i = 0; i < n_each; ++i) {  // cell centres of a 0.05 m grid seen from a centroid
    const double cx = ur(rng) * 0.5, cy = ur(rng) * 0.5;
    check((di(rng) + 0.5) * 0.05 - cy, (di(rng) + 0.5) * 0.05 - cx);
  }
  for (long i = 0; i < n_each; ++i) {  // collinear cells: one coordinate is rounding noise
    const double noise = ur(rng) * 1e-14, along = di(rng) * 0.05 + ur(rng) * 1e-15;
    check(along, noise);
    check(noise, along);
  }
  for (long i = 0; i < n_each; ++i) check(ldexp(ur(rng), (int)ue(rng)), ldexp(ur(rng), (int)ue(rng)));  // any magnitude
  const double sp[] = {0.0, -0.0, 1.0, -1.0, INFINITY, -INFINITY, 1e-310, -1e-310, 1e308, -1e308, 5e-324, 0.05, -0.05};
  for (double y : sp)
    for (double x : sp) check(y, x);
  printf("calls %ld disagreements %ld unflagged %ld risky %ld\n", n, bad, bad_unflagged, n_risky);
  return bad_unflagged ? 1 : 0;
}
