/*
 * fold_check.c — test infrastructure (CPU).  The scoring kernel's set-up (score_fast.cu: fold_fast)
 * evaluates the weighted combine of gpu_plugins.go:731-743 and its per-branch constants WITHOUT the
 * reference's case distinction on empty sides, relying on c/c = 1, 0/c = 0 and "m / 0 -> 0" (the
 * {1/c, c} table's entry 0) to reproduce the literals 1.0 and 0.0 of the guarded form.  This program
 * compares the two forms bit for bit (memcmp, so signs of zero count) over every count pair with
 * n + p + 1 < 64 and pseudo-random sums.  IEEE division stands in for the kernel's table division,
 * which is correctly rounded (gs_selftest checks that on the GPU).  Build with -ffp-contract=off.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <stdint.h>
static double dvq(double m, int c) { return c == 0 ? 0.0 : m / (double)c; }  /* table form: lut[0] -> 0 */
typedef struct { double tmp0, kk_n, cst_n, kk_p, cst_p, tn, tp; } R;
static R old_fold(double ns, double ps, int n, int p) {
  R r; double tmp = 0.0;
  if (n > 0 && p > 0) { double k = (double)n / (double)(n + p); double a = 100.0 * (((1.0 - k) * ps) / p); double b = 100.0 * ((k * ns) / n); tmp = a + b; }
  else if (n > 0) tmp = 100.0 * (ns / n);
  else if (p > 0) tmp = 100.0 * (ps / p);
  r.tmp0 = tmp;
  r.kk_n = 1.0; r.cst_n = 0.0; r.kk_p = 1.0; r.cst_p = 0.0;
  if (p > 0) { double kq = (double)(n + 1) / (double)(n + 1 + p); r.kk_n = kq; r.cst_n = 100.0 * (((1.0 - kq) * ps) / p); }
  if (n > 0) { double kq = (double)n / (double)(n + p + 1); r.kk_p = 1.0 - kq; r.cst_p = 100.0 * ((kq * ns) / n); }
  r.tn = r.cst_n + 100.0 * ((r.kk_n * (ns + 0.0)) / (n + 1));
  r.tp = r.cst_p + 100.0 * ((r.kk_p * (ps + 0.0)) / (p + 1));
  return r;
}
static R new_fold(double ns, double ps, int n, int p) {
  R r;
  double k0 = dvq((double)n, n + p);
  r.tmp0 = 100.0 * dvq((1.0 - k0) * ps, p) + 100.0 * dvq(k0 * ns, n);
  r.kk_n = dvq((double)(n + 1), n + p + 1);
  r.cst_n = 100.0 * dvq((1.0 - r.kk_n) * ps, p);
  double kq_p = dvq((double)n, n + p + 1);
  r.kk_p = 1.0 - kq_p;
  r.cst_p = 100.0 * dvq(kq_p * ns, n);
  r.tn = r.cst_n + 100.0 * dvq(r.kk_n * (ns + 0.0), n + 1);
  r.tp = r.cst_p + 100.0 * dvq(r.kk_p * (ps + 0.0), p + 1);
  return r;
}
int main(void) {
  uint64_t x = 88172645463325252ull; long bad = 0, tot = 0;
  for (int n = 0; n < 63; ++n) for (int p = 0; n + p + 1 < 64; ++p) for (int it = 0; it < 300; ++it) {
    x ^= x << 13; x ^= x >> 7; x ^= x << 17;
    double u1 = (double)(x >> 11) / 9007199254740992.0; x ^= x << 13; x ^= x >> 7; x ^= x << 17;
    double u2 = (double)(x >> 11) / 9007199254740992.0;
    double ns = n ? n * u1 : 0.0, ps = p ? p * u2 : 0.0;   /* empty side -> sum 0 */
    if (it % 7 == 0) ns *= 1e-30; if (it % 11 == 0) ps *= 1e-60;
    R a = old_fold(ns, ps, n, p), b = new_fold(ns, ps, n, p);
    ++tot; if (memcmp(&a, &b, sizeof a)) { if (bad < 5) printf("diff n=%d p=%d ns=%g ps=%g: %.17g %.17g | %.17g %.17g | %.17g %.17g\n", n, p, ns, ps, a.tmp0, b.tmp0, a.kk_n, b.kk_n, a.cst_p, b.cst_p); ++bad; }
  }
  printf("%ld cases, %ld differ\n", tot, bad); return bad != 0;
}
