// atan2_cr.h -- atan2 in double, correctly rounded for all practical purposes (double-double evaluation, error < 2^-98
// of the result), the same code on the device and on the host.
//
// Why: the reference sorts the cells of a frontier chunk by atan2 of their offsets from the chunk's centroid
// (SortByMedianFunctor, DEP/include/frontier_exploration/FrontierSearch.hpp:156-184) and takes the median.  On a straight
// axis-aligned frontier the offsets are collinear up to rounding noise of the centroid (dx ~ 1e-15 at 0.05 m), and the
// order of the angles is decided in their last bit.  CUDA's atan2 is within 2 ulp of the exact value, glibc's (the
// reference's) is the correctly rounded value except when the exact value lies within ~2^-90 of a rounding boundary, so
// the two can order such cells differently.  This one returns the correctly rounded value under the same condition (2^-98),
// so it equals glibc's wherever both are correctly rounded: tests/test_atan2.py compares the two over millions of inputs of
// the kind the search produces, on the host (this header compiles for both) and on the device.
//
// Method: atan2(y, x) = +-(0 | pi/2 | pi) +- atan(q), q = min(|x|, |y|) / max(|x|, |y|) in [0, 1] as a double-double;
// atan(q) = atan(k / 64) + atan(s), s = (q - k / 64) / (1 + q k / 64), |s| <= 1 / 127, k = round(64 q): a table of 65
// double-double values and an odd series in s up to s^13 (s^15 / 15 < 2^-108), leading terms in double-double.
#pragma once
#include <math.h>

#ifdef __CUDACC__
#define ATAN2_CR_FN __host__ __device__ inline
#else
#define ATAN2_CR_FN inline
#endif

namespace atan2cr {

struct dd { double hi, lo; };

ATAN2_CR_FN dd two_sum(double a, double b) {
  const double s = a + b, bb = s - a;
  return dd{s, (a - (s - bb)) + (b - bb)};
}
ATAN2_CR_FN dd fast_two_sum(double a, double b) {  // |a| >= |b|
  const double s = a + b;
  return dd{s, b - (s - a)};
}
ATAN2_CR_FN dd two_prod(double a, double b) {
  const double p = a * b;
  return dd{p, fma(a, b, -p)};
}
ATAN2_CR_FN dd add(dd a, dd b) {
  dd s = two_sum(a.hi, b.hi);
  const dd t = two_sum(a.lo, b.lo);
  s.lo += t.hi;
  s = fast_two_sum(s.hi, s.lo);
  s.lo += t.lo;
  return fast_two_sum(s.hi, s.lo);
}
ATAN2_CR_FN dd neg(dd a) { return dd{-a.hi, -a.lo}; }
ATAN2_CR_FN dd mul(dd a, dd b) {
  dd p = two_prod(a.hi, b.hi);
  p.lo += a.hi * b.lo + a.lo * b.hi;
  return fast_two_sum(p.hi, p.lo);
}
ATAN2_CR_FN dd mul_d(dd a, double b) {
  dd p = two_prod(a.hi, b);
  p.lo += a.lo * b;
  return fast_two_sum(p.hi, p.lo);
}
ATAN2_CR_FN dd div(dd a, dd b) {
  const double q1 = a.hi / b.hi;
  dd r = add(a, neg(mul_d(b, q1)));
  const double q2 = r.hi / b.hi;
  r = add(r, neg(mul_d(b, q2)));
  const double q3 = r.hi / b.hi;
  dd q = fast_two_sum(q1, q2);
  q.lo += q3;
  return fast_two_sum(q.hi, q.lo);
}

// atan(k / 64), k = 0 .. 64, as double-double (generated with 80-digit decimal arithmetic)
#ifdef __CUDA_ARCH__
__device__
#endif
static const double ATAN_TAB[65][2] = {
    {0x0.0p+0, 0x0.0p+0},  // atan(0 / 64)
    {0x1.fff555bbb729bp-7, -0x1.220c39d4dff50p-61},  // atan(1 / 64)
    {0x1.ffd55bba97625p-6, -0x1.5ec431444912cp-60},  // atan(2 / 64)
    {0x1.7fb818430da2ap-5, -0x1.86ef8f794f105p-63},  // atan(3 / 64)
    {0x1.ff55bb72cfdeap-5, -0x1.c934d86d23f1dp-60},  // atan(4 / 64)
    {0x1.3f59f0e7c559dp-4, 0x1.ac4ce285df847p-58},  // atan(5 / 64)
    {0x1.7ee182602f10fp-4, -0x1.cfb654c0c3d98p-58},  // atan(6 / 64)
    {0x1.be39ebe6f07c3p-4, 0x1.f7b8f29a05987p-58},  // atan(7 / 64)
    {0x1.fd5ba9aac2f6ep-4, -0x1.cd37686760c17p-59},  // atan(8 / 64)
    {0x1.1e1fafb043727p-3, -0x1.b485914dacf8cp-59},  // atan(9 / 64)
    {0x1.3d6eee8c6626cp-3, 0x1.61a3b0ce9281bp-57},  // atan(10 / 64)
    {0x1.5c9811e3ec26ap-3, -0x1.054ab2c010f3dp-58},  // atan(11 / 64)
    {0x1.7b97b4bce5b02p-3, 0x1.347b0b4f881cap-58},  // atan(12 / 64)
    {0x1.9a6a8e96c8626p-3, 0x1.cf601e7b4348ep-59},  // atan(13 / 64)
    {0x1.b90d7529260a2p-3, 0x1.17b10d2e0e5abp-61},  // atan(14 / 64)
    {0x1.d77d5df205736p-3, 0x1.c648d1534597ep-57},  // atan(15 / 64)
    {0x1.f5b75f92c80ddp-3, 0x1.8ab6e3cf7afbdp-57},  // atan(16 / 64)
    {0x1.09dc597d86362p-2, 0x1.62e47390cb865p-56},  // atan(17 / 64)
    {0x1.18bf5a30bf178p-2, 0x1.30ca4748b1bf9p-57},  // atan(18 / 64)
    {0x1.278372057ef46p-2, -0x1.077cdd36dfc81p-56},  // atan(19 / 64)
    {0x1.362773707ebccp-2, -0x1.963a544b672d8p-57},  // atan(20 / 64)
    {0x1.44aa436c2af0ap-2, -0x1.5d5e43c55b3bap-56},  // atan(21 / 64)
    {0x1.530ad9951cd4ap-2, -0x1.2566480884082p-57},  // atan(22 / 64)
    {0x1.614840309cfe2p-2, -0x1.a725715711f00p-56},  // atan(23 / 64)
    {0x1.6f61941e4def1p-2, -0x1.c63aae6f6e918p-56},  // atan(24 / 64)
    {0x1.7d5604b63b3f7p-2, 0x1.69c885c2b249ap-56},  // atan(25 / 64)
    {0x1.8b24d394a1b25p-2, 0x1.b6d0ba3748fa8p-56},  // atan(26 / 64)
    {0x1.98cd5454d6b18p-2, 0x1.9e6c988fd0a77p-56},  // atan(27 / 64)
    {0x1.a64eec3cc23fdp-2, -0x1.24dec1b50b7ffp-56},  // atan(28 / 64)
    {0x1.b3a911da65c6cp-2, 0x1.ae187b1ca5040p-56},  // atan(29 / 64)
    {0x1.c0db4c94ec9f0p-2, -0x1.cc1ce70934c34p-56},  // atan(30 / 64)
    {0x1.cde53432c1351p-2, -0x1.a2cfa4418f1adp-56},  // atan(31 / 64)
    {0x1.dac670561bb4fp-2, 0x1.a2b7f222f65e2p-56},  // atan(32 / 64)
    {0x1.e77eb7f175a34p-2, 0x1.0e53dc1bf3435p-56},  // atan(33 / 64)
    {0x1.f40dd0b541418p-2, -0x1.a3992dc382a23p-57},  // atan(34 / 64)
    {0x1.0039c73c1a40cp-1, -0x1.b32c949c9d593p-55},  // atan(35 / 64)
    {0x1.0657e94db30d0p-1, -0x1.d5b495f6349e6p-56},  // atan(36 / 64)
    {0x1.0c6145b5b43dap-1, 0x1.974fa13b5404fp-58},  // atan(37 / 64)
    {0x1.1255d9bfbd2a9p-1, -0x1.2bdaee1c0ee35p-58},  // atan(38 / 64)
    {0x1.1835a88be7c13p-1, 0x1.c621cec00c301p-55},  // atan(39 / 64)
    {0x1.1e00babdefeb4p-1, -0x1.928df287a668fp-58},  // atan(40 / 64)
    {0x1.23b71e2cc9e6ap-1, 0x1.c421c9f38224ep-57},  // atan(41 / 64)
    {0x1.2958e59308e31p-1, -0x1.09e73b0c6c087p-56},  // atan(42 / 64)
    {0x1.2ee628406cbcap-1, 0x1.c5d5e9ff0cf8dp-55},  // atan(43 / 64)
    {0x1.345f01cce37bbp-1, 0x1.1021137c71102p-55},  // atan(44 / 64)
    {0x1.39c391cd4171ap-1, -0x1.2304331d8bf46p-55},  // atan(45 / 64)
    {0x1.3f13fb89e96f4p-1, 0x1.ecf8b492644f0p-56},  // atan(46 / 64)
    {0x1.445065b795b56p-1, -0x1.f76d0163f79c8p-56},  // atan(47 / 64)
    {0x1.4978fa3269ee1p-1, 0x1.2419a87f2a458p-56},  // atan(48 / 64)
    {0x1.4e8de5bb6ec04p-1, 0x1.4a33dbeb3796cp-55},  // atan(49 / 64)
    {0x1.538f57b89061fp-1, -0x1.1bb74abda520cp-55},  // atan(50 / 64)
    {0x1.587d81f732fbbp-1, -0x1.5e5c9d8c5a950p-56},  // atan(51 / 64)
    {0x1.5d58987169b18p-1, 0x1.0028e4bc5e7cap-57},  // atan(52 / 64)
    {0x1.6220d115d7b8ep-1, -0x1.2b785350ee8c1p-57},  // atan(53 / 64)
    {0x1.66d663923e087p-1, -0x1.6ea6febe8bbbap-56},  // atan(54 / 64)
    {0x1.6b798920b3d99p-1, -0x1.a80386188c50ep-55},  // atan(55 / 64)
    {0x1.700a7c5784634p-1, -0x1.8c34d25aadef6p-56},  // atan(56 / 64)
    {0x1.748978fba8e0fp-1, 0x1.7b2a6165884a1p-59},  // atan(57 / 64)
    {0x1.78f6bbd5d315ep-1, 0x1.406a089803740p-55},  // atan(58 / 64)
    {0x1.7d528289fa093p-1, 0x1.560821e2f3aa9p-55},  // atan(59 / 64)
    {0x1.819d0b7158a4dp-1, -0x1.bf76229d3b917p-56},  // atan(60 / 64)
    {0x1.85d69576cc2c5p-1, 0x1.6b66e7fc8b8c3p-57},  // atan(61 / 64)
    {0x1.89ff5ff57f1f8p-1, -0x1.55b9a5e177a1bp-55},  // atan(62 / 64)
    {0x1.8e17aa99cc05ep-1, -0x1.ec182ab042f61p-56},  // atan(63 / 64)
    {0x1.921fb54442d18p-1, 0x1.1a62633145c07p-55},  // atan(64 / 64)
};

// `risky` (optional): the exact value lies within 0.03 ulp of a rounding boundary.  glibc's atan2 (the reference's) has an
// error of up to 0.523 ulp (measured: the farthest of 20 000 disagreements in 6e7 calls sat 0.0225 ulp from the boundary),
// so it returns the correctly rounded value whenever this flag is clear, and may return its neighbour when it is set.
ATAN2_CR_FN double atan2_cr(double y, double x, bool *risky = nullptr) {
  if (risky) *risky = false;
  if (x != x || y != y) return x + y;
  const dd PI = {0x1.921fb54442d18p+1, 0x1.1a62633145c07p-53}, PI_2 = {0x1.921fb54442d18p+0, 0x1.1a62633145c07p-54};
  const double ax = fabs(x), ay = fabs(y);
  // zeros and infinities as C99 / glibc: the sign of y carries over, x decides between 0, pi/2 and pi
  if (ay == 0.0) return signbit(x) ? copysign(PI.hi, y) : y;
  if (ax == 0.0) return copysign(PI_2.hi, y);
  if (isinf(ax) || isinf(ay)) {
    double r;
    if (isinf(ax) && isinf(ay)) r = signbit(x) ? 0x1.2d97c7f3321d2p+1 : 0x1.921fb54442d18p-1;  // 3 pi / 4, pi / 4
    else if (isinf(ay)) r = PI_2.hi;
    else r = signbit(x) ? PI.hi : 0.0;
    return copysign(r, y);
  }
  const bool swap = ay > ax;
  double num = swap ? ax : ay, den = swap ? ay : ax;
  // keep the quotient's remainder exact: scale very large / very small operands (powers of two: exact)
  if (den > 0x1p+1000) { num *= 0x1p-100; den *= 0x1p-100; }
  if (den < 0x1p-900) { num *= 0x1p+200; den *= 0x1p+200; }
  dd q;
  q.hi = num / den;
  q.lo = fma(-q.hi, den, num) / den;
  dd r;
  if (q.hi < 0x1p-60) {
    r = q;  // atan(q) = q (1 - q^2 / 3 + ...): the correction is below 2^-120 of q
  } else {
    const double kf = rint(q.hi * 64.0);
    const int k = (int)kf;
    dd s = q;
    if (k > 0) {
      const double c = kf * 0.015625;
      dd n = two_sum(q.hi, -c);
      n.lo += q.lo;
      n = fast_two_sum(n.hi, n.lo);
      dd d = two_prod(q.hi, c);
      d.lo += q.lo * c;
      d = fast_two_sum(d.hi, d.lo);
      d = add(dd{1.0, 0.0}, d);
      s = div(n, d);
    }
    // atan(s) = s - s^3 / 3 + s^5 / 5 - ... + s^13 / 13
    const dd s2 = mul(s, s);
    const double z = s2.hi;
    const double tail = z * (-0x1.2492492492492p-3 + z * (0x1.c71c71c71c71cp-4 + z * (-0x1.745d1745d1746p-4 + z * 0x1.3b13b13b13b14p-4)));  // -1/7, 1/9, -1/11, 1/13
    dd p = add(dd{0x1.999999999999ap-3, -0x1.999999999999ap-57}, dd{tail, 0.0});  // 1/5 + ...
    p = mul(p, s2);
    p = add(dd{-0x1.5555555555555p-2, -0x1.5555555555555p-56}, p);                  // -1/3 + ...
    p = mul(p, s2);
    p = mul(p, s);
    r = add(s, p);
    if (k > 0) r = add(dd{ATAN_TAB[k][0], ATAN_TAB[k][1]}, r);
  }
  if (swap) r = add(PI_2, neg(r));
  if (signbit(x)) r = add(PI, neg(r));
  if (risky) {
    int e;
    frexp(r.hi, &e);  // r.hi = m * 2^e, m in [0.5, 1): ulp(r.hi) = 2^(e - 53)
    *risky = fabs(r.lo) > 0.47 * ldexp(1.0, e - 53);
  }
  return copysign(r.hi + r.lo, y);
}

}  // namespace atan2cr
