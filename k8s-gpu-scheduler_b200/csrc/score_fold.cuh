// score_fold.cuh — what the two restructured scoring kernels share: the count table, the guarded
// forms of the weighted combine (gpu_plugins.go:675-688, :731-743) and the float32 filter's constants.
#pragma once
#include "score_common.cuh"

namespace gs {

constexpr int kLutSize = 64;  // counts the {1/c, c} table covers

// m / c for a count c: the table form where its preconditions hold, IEEE division otherwise
// (both correctly rounded, hence identical).
__device__ __forceinline__ double div_count(double m, int c, const double2* lut) {
  const double am = fabs(m);
  if (c < kLutSize && (am == 0.0 || (am > 1e-250 && am < 1e300))) {
    const double2 e = lut[c];
    return div_small_int(m, e.x, e.y);
  }
  return __ddiv_rn(m, (double)c);
}

// combine() of score_math.cuh with the divisions by counts done through the table
__device__ __forceinline__ double combine_lut(double neg_sum, double pos_sum, int cn, int cp, const double2* lut) {
  double tmp = 0.0;
  if (cn > 0 && cp > 0) {
    const double k = div_count((double)cn, cn + cp, lut);
    const double a = __dmul_rn(100.0, div_count(__dmul_rn(__dsub_rn(1.0, k), pos_sum), cp, lut));
    const double b = __dmul_rn(100.0, div_count(__dmul_rn(k, neg_sum), cn, lut));
    tmp = __dadd_rn(a, b);
  } else if (cn > 0) {
    tmp = __dmul_rn(100.0, div_count(neg_sum, cn, lut));
  } else if (cp > 0) {
    tmp = __dmul_rn(100.0, div_count(pos_sum, cp, lut));
  }
  return tmp;
}

// {kk, cst} of one branch: everything of the weighted combine that does not depend on the
// candidate's own term, with the operations and their order as in combine() (x*1.0 and 0.0+x are
// exact, so an absent side contributes kk = 1, cst = 0).
__device__ __forceinline__ double2 branch_const(bool neg, double neg_sum, double pos_sum, int cn0, int cp0,
                                                const double2* lut) {
  double kk = 1.0, cst = 0.0;
  if (neg) {
    if (cp0 > 0) {  // candidate lands on the negative side, positives exist
      const double kq = div_count((double)(cn0 + 1), cn0 + 1 + cp0, lut);
      kk = kq;
      cst = __dmul_rn(100.0, div_count(__dmul_rn(__dsub_rn(1.0, kq), pos_sum), cp0, lut));
    }
  } else if (cn0 > 0) {  // candidate lands on the positive side, negatives exist
    const double kq = div_count((double)cn0, cn0 + cp0 + 1, lut);
    kk = __dsub_rn(1.0, kq);
    cst = __dmul_rn(100.0, div_count(__dmul_rn(kq, neg_sum), cn0, lut));
  }
  return make_double2(kk, cst);
}

// Float32 filter constants.  With A, B >= 0 and A + B <= kFilterMaxSum the float32 estimate
//   est = fma(B32, rcp_approx(z32), A32),  z32 = 1 + |t|  or  fma(1 + |t|, 1 + |t|, 1)
// differs from the reference's float64 tmp by less than
//   2u*A (rounding A, the final fma) + 7u*B*v (rounding B, z32: at most 3u relative, the
//   approximate reciprocal: 1 ulp = 2u, the final fma)  <=  7u * kFilterMaxSum = 4.2e-5,  u = 2^-24,
// plus the float64 roundings of tmp itself and of A, B (below 1e-11).  Mathematically A + B <= 100
// (it is tmp at v = 1, a weighted mean of values in [0, 1] times 100); constants outside
// [0, kFilterMaxSum] switch the filter off for their GPU.  kFilterDelta = 2^-14 = 6.1e-5 leaves a
// factor of 1.45 (still 1.1 if the reciprocal were 2 ulp off).  -DGS_FILTER_DELTA_LOG2=13
// -DGS_FILTER_MAX_SUM=128 is the wider margin the first version ran with (5 % slower).
#ifndef GS_FILTER_DELTA_LOG2
#define GS_FILTER_DELTA_LOG2 14
#endif
#ifndef GS_FILTER_MAX_SUM
#define GS_FILTER_MAX_SUM 100.5
#endif
constexpr float kFilterDelta = 1.0f / (float)(1 << GS_FILTER_DELTA_LOG2);
constexpr double kFilterMaxSum = GS_FILTER_MAX_SUM;  // A + B may not exceed this


// ---- one GPU's sums folded into what the pod loops need ------------------------------------
struct FoldOut {
  float An, Bn, Ap, Bp;  // float32 filter: tmp ~ A + B*v per branch; (0, 0) = always take the exact path
  int z_neg, z_pos;      // int64(tmp) when the candidate's term is exactly 0 (pods without an SLO)
  int si;                // int64(tmp) of the residents alone (no candidate term, :667-696)
};

// The streamlined domain of one GPU: non-negative finite sums of ordinary size (they are sums of
// fewer than 64 values in (0, 1]), a finite candidate interference, counts inside the table.
__device__ __forceinline__ bool fold_in_domain(const GpuAcc& acc, float ci) {
  return acc.neg_sum >= 0.0 && acc.neg_sum <= 1.0e6 && acc.pos_sum >= 0.0 && acc.pos_sum <= 1.0e6 &&
         fabsf(ci) <= 3.0e38f && acc.cnt_neg + acc.cnt_pos + 1 < kLutSize;  // false for NaN
}

// Inside the domain: the same operations in the same order as combine() / branch_const(), every
// division through the table and none of them guarded.  An empty side needs no special case: its k
// comes out as exactly 1 or 0 (c/c and 0/c are exact) and its product as exactly 0 (lut[0] = {0, 0}
// turns m/0 into 0), and x*1.0, 0.0+x are exact — so the values equal the guarded form's literals.
__device__ __forceinline__ FoldOut fold_fast_fn(const GpuAcc& acc, const double2* lut) {
  const int n = acc.cnt_neg, p = acc.cnt_pos;
  const double ns = acc.neg_sum, ps = acc.pos_sum;
  const double2 Ln = lut[n], Lp = lut[p], Lnp = lut[n + p], Lnp1 = lut[n + p + 1], Ln1 = lut[n + 1], Lp1 = lut[p + 1];
  auto dv = [](double m, const double2& L) { return div_small_int(m, L.x, L.y); };
  FoldOut o;
  const double k0 = dv((double)n, Lnp);
  const double tmp0 = __dadd_rn(__dmul_rn(100.0, dv(__dmul_rn(__dsub_rn(1.0, k0), ps), Lp)),
                                __dmul_rn(100.0, dv(__dmul_rn(k0, ns), Ln)));
  o.si = score_int(tmp0);
  const double kk_n = dv((double)(n + 1), Lnp1);  // candidate on the negative side: k = (n+1)/(n+1+p)
  const double cst_n = __dmul_rn(100.0, dv(__dmul_rn(__dsub_rn(1.0, kk_n), ps), Lp));
  const double kq_p = dv((double)n, Lnp1);        // candidate on the positive side: k = n/(n+p+1)
  const double kk_p = __dsub_rn(1.0, kq_p);
  const double cst_p = __dmul_rn(100.0, dv(__dmul_rn(kq_p, ns), Ln));
  const double tn = __dadd_rn(cst_n, __dmul_rn(100.0, dv(__dmul_rn(kk_n, __dadd_rn(ns, 0.0)), Ln1)));
  const double tp = __dadd_rn(cst_p, __dmul_rn(100.0, dv(__dmul_rn(kk_p, __dadd_rn(ps, 0.0)), Lp1)));
  o.z_neg = score_int(tn);
  o.z_pos = score_int(tp);
  const double Bn = 100.0 * kk_n * Ln1.x, Bp = 100.0 * kk_p * Lp1.x;
  const bool okn = tn >= 0.0 && Bn >= 0.0 && tn + Bn <= kFilterMaxSum, okp = tp >= 0.0 && Bp >= 0.0 && tp + Bp <= kFilterMaxSum;
  o.An = okn ? (float)tn : 0.f;
  o.Bn = okn ? (float)Bn : 0.f;
  o.Ap = okp ? (float)tp : 0.f;
  o.Bp = okp ? (float)Bp : 0.f;
  return o;
}

// Outside the domain (guarded divisions, IEEE division where the table does not reach): the
// residents-only score and the zero-term scores stay exact; the filter is off (candidates of this
// node go through logic_eval).
static __device__ __noinline__ FoldOut fold_general_fn(const GpuAcc& acc, const double2* lut) {
  const int cn0 = acc.cnt_neg, cp0 = acc.cnt_pos;
  FoldOut o;
  o.si = score_int(combine_lut(acc.neg_sum, acc.pos_sum, cn0, cp0, lut));
  const double2 kn = branch_const(true, acc.neg_sum, acc.pos_sum, cn0, cp0, lut);  // {kk, cst}
  const double2 kp = branch_const(false, acc.neg_sum, acc.pos_sum, cn0, cp0, lut);
  const double tn = __dadd_rn(kn.y, __dmul_rn(100.0, div_count(__dmul_rn(kn.x, __dadd_rn(acc.neg_sum, 0.0)), cn0 + 1, lut)));
  const double tp = __dadd_rn(kp.y, __dmul_rn(100.0, div_count(__dmul_rn(kp.x, __dadd_rn(acc.pos_sum, 0.0)), cp0 + 1, lut)));
  o.z_neg = score_int(tn);
  o.z_pos = score_int(tp);
  o.An = o.Bn = o.Ap = o.Bp = 0.f;
  return o;
}

}  // namespace gs
