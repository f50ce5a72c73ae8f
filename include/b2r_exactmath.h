/*
 * b2r_exactmath.h -- acos, cos and atan2 in fp64, evaluated in double-double arithmetic from +, -, *, /, sqrt and fma only.
 *
 * WHY.  The reference's Open3D calls (estimate_normals' analytic eigen-solver, compute_fpfh_feature's pair features;
 * HelpersRealtime.py:215-224) evaluate acos / cos / atan2 with the host's libm.  CUDA's libm differs from glibc's in the last
 * ulp, which is enough to move a pair feature across a histogram bin and a normal by 1e-7 through the eigen-solver
 * (VERDICT round 1, weak item 3).  This header is included VERBATIM by both poseestimator_b200/csrc/presteps.cu (device) and
 * oracle/presteps_oracle.cpp (host): every operation is an IEEE-754 basic operation or an fma, none is contracted, so the two
 * sides return identical bits by construction.  The results are the correctly rounded values except when the exact value lies
 * within ~2^-100 (relative) of a rounding boundary; tests/test_oracle_exactmath.py checks them against mpmath (correct
 * rounding) and against glibc (<= 1 ulp, the accuracy glibc documents).
 *
 * Test and product both include it; it is arithmetic, not the oracle: nothing under oracle/ is reachable from the product.
 * Constants: tools/gen_exactmath_tables.py (mpmath, 300 bits).
 */
#ifndef B2R_EXACTMATH_H_
#define B2R_EXACTMATH_H_

#include <math.h>

#if defined(__CUDACC__)
#define B2R_XM_FN __device__ __forceinline__
#define B2R_XM_FN_NOINLINE __device__ __noinline__
#define B2R_XM_CONST static __constant__ const
#define XM_ADD(a, b) __dadd_rn((a), (b))
#define XM_SUB(a, b) __dsub_rn((a), (b))
#define XM_MUL(a, b) __dmul_rn((a), (b))
#define XM_DIV(a, b) __ddiv_rn((a), (b))
#define XM_SQRT(a) __dsqrt_rn((a))
#define XM_FMA(a, b, c) __fma_rn((a), (b), (c))
#else
/* host: compile with -ffp-contract=off (oracle/Makefile does); the pragma and attribute below repeat it for this header */
#if defined(__GNUC__) && !defined(__clang__)
#pragma GCC push_options
#pragma GCC optimize("fp-contract=off")
#endif
#define B2R_XM_FN static inline
#define B2R_XM_FN_NOINLINE static
#define B2R_XM_CONST static const
#define XM_ADD(a, b) ((a) + (b))
#define XM_SUB(a, b) ((a) - (b))
#define XM_MUL(a, b) ((a) * (b))
#define XM_DIV(a, b) ((a) / (b))
#define XM_SQRT(a) sqrt((a))
#define XM_FMA(a, b, c) __builtin_fma((a), (b), (c))
#endif

typedef struct b2r_dd { double hi, lo; } b2r_dd;

B2R_XM_CONST double kAtanTab[17][2] = {
    {0.0, 0.0},
    {0.06241880999595735, -1.5490756308295046e-18},
    {0.12435499454676144, -3.1253241424539383e-18},
    {0.18534794999569476, 4.180692268843079e-18},
    {0.24497866312686414, 1.0698755618734451e-17},
    {0.3028848683749714, -1.1010827903001369e-17},
    {0.35877067027057225, -2.4623815582638635e-17},
    {0.4124104415973873, -1.587652227770689e-17},
    {0.4636476090008061, 2.2698777452961687e-17},
    {0.5123894603107377, -2.5462781472855804e-17},
    {0.5585993153435624, -5.4556305485916264e-18},
    {0.6022873461349642, 2.950430737228402e-17},
    {0.6435011087932844, 1.5834785051444286e-17},
    {0.6823165548747481, 6.943223671560008e-18},
    {0.7188299996216245, -2.1478388444456983e-17},
    {0.7531512809621944, -2.4256934659182068e-17},
    {0.7853981633974483, 3.061616997868383e-17},
};
B2R_XM_CONST double kOddInv[14][2] = {
    {1.0, 0.0},
    {0.3333333333333333, 1.850371707708594e-17},
    {0.2, -1.1102230246251566e-17},
    {0.14285714285714285, 7.93016446160826e-18},
    {0.1111111111111111, 6.1679056923619804e-18},
    {0.09090909090909091, -2.523234146875356e-18},
    {0.07692307692307693, -4.270088556250602e-18},
    {0.06666666666666667, 9.251858538542971e-19},
    {0.058823529411764705, 8.163404592832033e-19},
    {0.05263157894736842, 2.921639538487254e-18},
    {0.047619047619047616, 2.64338815386942e-18},
    {0.043478260869565216, 1.206764157201257e-18},
    {0.04, -8.326672684688674e-19},
    {0.037037037037037035, 2.05596856412066e-18},
};
B2R_XM_CONST double kInvFact[30][2] = {
    {1.0, 0.0},
    {1.0, 0.0},
    {0.5, 0.0},
    {0.16666666666666666, 9.25185853854297e-18},
    {0.041666666666666664, 2.3129646346357427e-18},
    {0.008333333333333333, 1.1564823173178714e-19},
    {0.001388888888888889, -5.300543954373577e-20},
    {0.0001984126984126984, 1.7209558293420705e-22},
    {2.48015873015873e-05, 2.1511947866775882e-23},
    {2.7557319223985893e-06, -1.858393274046472e-22},
    {2.755731922398589e-07, 2.3767714622250297e-23},
    {2.505210838544172e-08, -1.448814070935912e-24},
    {2.08767569878681e-09, -1.20734505911326e-25},
    {1.6059043836821613e-10, 1.2585294588752098e-26},
    {1.1470745597729725e-11, 2.0655512752830745e-28},
    {7.647163731819816e-13, 7.03872877733453e-30},
    {4.779477332387385e-14, 4.399205485834081e-31},
    {2.8114572543455206e-15, 1.6508842730861433e-31},
    {1.5619206968586225e-16, 1.1910679660273754e-32},
    {8.22063524662433e-18, 2.2141894119604265e-34},
    {4.110317623312165e-19, 1.4412973378659527e-36},
    {1.9572941063391263e-20, -1.3643503830087908e-36},
    {8.896791392450574e-22, -7.911402614872376e-38},
    {3.868170170630684e-23, -8.843177655482344e-40},
    {1.6117375710961184e-24, -3.6846573564509766e-41},
    {6.446950284384474e-26, -1.9330404233703465e-42},
    {2.4795962632247976e-27, -1.2953730964765229e-43},
    {9.183689863795546e-29, 1.4303150396787322e-45},
    {3.279889237069838e-30, 1.5117542744029879e-46},
    {1.1309962886447716e-31, 1.0498015412959506e-47},
};
B2R_XM_CONST double kPi[2] = {3.141592653589793, 1.2246467991473532e-16};
B2R_XM_CONST double kPiHalf[2] = {1.5707963267948966, 6.123233995736766e-17};
B2R_XM_CONST double kPiQuarter[2] = {0.7853981633974483, 3.061616997868383e-17};
B2R_XM_CONST double kPiHalf3[3] = {1.5707963267948966, 6.123233995736766e-17, -1.4973849048591698e-33};
// 2/pi = 0.6366197723675814

B2R_XM_FN b2r_dd xm_make(double hi, double lo) { b2r_dd r; r.hi = hi; r.lo = lo; return r; }
B2R_XM_FN b2r_dd xm_two_sum(double a, double b) {
  const double s = XM_ADD(a, b), bb = XM_SUB(s, a);
  return xm_make(s, XM_ADD(XM_SUB(a, XM_SUB(s, bb)), XM_SUB(b, bb)));
}
B2R_XM_FN b2r_dd xm_quick_two_sum(double a, double b) {   /* |a| >= |b| */
  const double s = XM_ADD(a, b);
  return xm_make(s, XM_SUB(b, XM_SUB(s, a)));
}
B2R_XM_FN b2r_dd xm_two_prod(double a, double b) {
  const double p = XM_MUL(a, b);
  return xm_make(p, XM_FMA(a, b, -p));
}
B2R_XM_FN b2r_dd xm_add(b2r_dd x, b2r_dd y) {
  b2r_dd s = xm_two_sum(x.hi, y.hi);
  const b2r_dd t = xm_two_sum(x.lo, y.lo);
  s.lo = XM_ADD(s.lo, t.hi);
  s = xm_quick_two_sum(s.hi, s.lo);
  s.lo = XM_ADD(s.lo, t.lo);
  return xm_quick_two_sum(s.hi, s.lo);
}
B2R_XM_FN b2r_dd xm_neg(b2r_dd x) { return xm_make(-x.hi, -x.lo); }
B2R_XM_FN b2r_dd xm_sub(b2r_dd x, b2r_dd y) { return xm_add(x, xm_neg(y)); }
B2R_XM_FN b2r_dd xm_add_d(b2r_dd x, double d) {
  b2r_dd s = xm_two_sum(x.hi, d);
  s.lo = XM_ADD(s.lo, x.lo);
  return xm_quick_two_sum(s.hi, s.lo);
}
B2R_XM_FN b2r_dd xm_mul(b2r_dd x, b2r_dd y) {
  b2r_dd p = xm_two_prod(x.hi, y.hi);
  p.lo = XM_ADD(p.lo, XM_ADD(XM_MUL(x.hi, y.lo), XM_MUL(x.lo, y.hi)));
  return xm_quick_two_sum(p.hi, p.lo);
}
B2R_XM_FN b2r_dd xm_mul_d(b2r_dd x, double d) {
  b2r_dd p = xm_two_prod(x.hi, d);
  p.lo = XM_ADD(p.lo, XM_MUL(x.lo, d));
  return xm_quick_two_sum(p.hi, p.lo);
}
B2R_XM_FN b2r_dd xm_div(b2r_dd x, b2r_dd y) {   /* three quotient digits */
  const double q1 = XM_DIV(x.hi, y.hi);
  b2r_dd r = xm_sub(x, xm_mul_d(y, q1));
  const double q2 = XM_DIV(r.hi, y.hi);
  r = xm_sub(r, xm_mul_d(y, q2));
  const double q3 = XM_DIV(r.hi, y.hi);
  return xm_add_d(xm_quick_two_sum(q1, q2), q3);
}
B2R_XM_FN b2r_dd xm_sqrt(b2r_dd x) {            /* x >= 0 */
  if (!(x.hi > 0.0)) return xm_make(0.0, 0.0);
  const double s = XM_SQRT(x.hi);
  const b2r_dd s2 = xm_two_prod(s, s);
  const b2r_dd r = xm_sub(x, s2);
  return xm_two_sum(s, XM_DIV(r.hi, XM_MUL(2.0, s)));
}

/* atan(z), 0 <= z <= 1: z = c + (z - c) with c = k / 16, atan z = atan c + atan t, t = (z - c) / (1 + z c), |t| <= 1 / 32;
 * Taylor series of atan t to t^27 (truncation < 2^-130). */
B2R_XM_FN_NOINLINE b2r_dd xm_atan01(b2r_dd z) {
  int k = (int)XM_ADD(XM_MUL(z.hi, 16.0), 0.5);
  if (k > 16) k = 16;
  if (k < 0) k = 0;
  const double c = XM_MUL((double)k, 0.0625);
  const b2r_dd num = xm_add_d(z, -c);
  const b2r_dd den = xm_add_d(xm_mul_d(z, c), 1.0);
  const b2r_dd t = xm_div(num, den);
  const b2r_dd t2 = xm_mul(t, t);
  b2r_dd p = xm_make(kOddInv[13][0], kOddInv[13][1]);
  for (int m = 12; m >= 0; --m) p = xm_sub(xm_make(kOddInv[m][0], kOddInv[m][1]), xm_mul(t2, p));
  return xm_add(xm_make(kAtanTab[k][0], kAtanTab[k][1]), xm_mul(t, p));
}

/* angle of (x, y) in [0, pi] for y >= 0 (double-double operands, not both zero) */
B2R_XM_FN b2r_dd xm_atan2_pos(b2r_dd y, b2r_dd x) {
  const int xneg = x.hi < 0.0 || (x.hi == 0.0 && x.lo < 0.0);
  const b2r_dd ax = xneg ? xm_neg(x) : x;
  const int swap = y.hi > ax.hi || (y.hi == ax.hi && y.lo > ax.lo);
  b2r_dd a = xm_atan01(swap ? xm_div(ax, y) : xm_div(y, ax));
  if (swap) a = xm_sub(xm_make(kPiHalf[0], kPiHalf[1]), a);
  if (xneg) a = xm_sub(xm_make(kPi[0], kPi[1]), a);
  return a;
}

B2R_XM_FN double b2r_atan2(double y, double x) {
  if (x != x || y != y) return XM_ADD(x, y);
  const double ay = fabs(y), ax = fabs(x);
  const int xneg = signbit(x) ? 1 : 0;
  if (ay == 0.0) return copysign(xneg ? kPi[0] : 0.0, y);
  if (ax == 0.0) return copysign(kPiHalf[0], y);
  if (isinf(ax) || isinf(ay)) {
    if (isinf(ax) && isinf(ay)) return copysign(xneg ? XM_MUL(3.0, kPiQuarter[0]) : kPiQuarter[0], y);
    if (isinf(ax)) return copysign(xneg ? kPi[0] : 0.0, y);
    return copysign(kPiHalf[0], y);
  }
  const b2r_dd a = xm_atan2_pos(xm_make(ay, 0.0), xm_make(x, 0.0));
  return copysign(a.hi, y);
}

/* acos x = angle of (x, sqrt((1 - x)(1 + x))) */
B2R_XM_FN double b2r_acos(double x) {
  if (x != x) return x;
  if (fabs(x) > 1.0) return XM_DIV(XM_SUB(x, x), XM_SUB(x, x));   /* NaN */
  if (x == 1.0) return 0.0;
  if (x == -1.0) return kPi[0];
  if (x == 0.0) return kPiHalf[0];
  const b2r_dd s = xm_sqrt(xm_mul(xm_two_sum(1.0, -x), xm_two_sum(1.0, x)));
  return xm_atan2_pos(s, xm_make(x, 0.0)).hi;
}

/* cos x for |x| <= 2^20: x = k pi/2 + r with a three-piece pi/2, Taylor series of sin / cos on |r| <= pi/4 (to r^29 / r^28) */
B2R_XM_FN_NOINLINE double b2r_cos(double x) {
  const double ax = fabs(x);
  if (!(ax <= 1048576.0)) return XM_DIV(XM_SUB(x, x), XM_SUB(x, x));   /* NaN: outside the supported range, inf or NaN */
  if (ax < 7.450580596923828125e-9) return 1.0;                         /* 2^-27: 1 - x^2/2 rounds to 1 */
  const double kd = floor(XM_ADD(XM_MUL(ax, 0.6366197723675814), 0.5));
  const int k = (int)kd;
  b2r_dd r = xm_sub(xm_make(ax, 0.0), xm_two_prod(kd, kPiHalf3[0]));
  r = xm_sub(r, xm_two_prod(kd, kPiHalf3[1]));
  r = xm_add_d(r, -XM_MUL(kd, kPiHalf3[2]));
  const b2r_dd r2 = xm_mul(r, r);
  b2r_dd p;
  if (k & 1) {   /* +- sin r */
    p = xm_make(kInvFact[29][0], kInvFact[29][1]);
    for (int n = 27; n >= 1; n -= 2) p = xm_sub(xm_make(kInvFact[n][0], kInvFact[n][1]), xm_mul(r2, p));
    p = xm_mul(r, p);
    return (k & 2) ? p.hi : -p.hi;
  }
  p = xm_make(kInvFact[28][0], kInvFact[28][1]);
  for (int n = 26; n >= 0; n -= 2) p = xm_sub(xm_make(kInvFact[n][0], kInvFact[n][1]), xm_mul(r2, p));
  return (k & 2) ? -p.hi : p.hi;
}

#if !defined(__CUDACC__) && defined(__GNUC__) && !defined(__clang__)
#pragma GCC pop_options
#endif
#endif /* B2R_EXACTMATH_H_ */
