// tridiag.cu — BayesianRidge for the wide tables, in a tridiagonal basis (sm_100a).
//
// BayesianRidge.fit takes an SVD of the centred design matrix per feature
// (pkg/recommender/recommender/recommender.py:15-17 -> sklearn linear_model/_bayes.py:262-306) and then
// iterates alpha, lambda with, per round, coef = V diag(1 / (S^2 + lambda/alpha)) V^T X^T y,
// gamma = sum(alpha S^2 / (lambda + alpha S^2)) and the residual (_bayes.py:309-339).  Every one of these is a
// function of the Gram matrix G = X^T X (without feature j) through (G + mu I)^-1, mu = lambda/alpha:
//   coef  = (G + mu I)^-1 b,              b = X^T y,
//   gamma = m - mu tr((G + mu I)^-1),
//   sse   = y^T y - coef.b - mu |coef|^2.
// None of it needs eigenvectors.  With G = Q T Q^T, T tridiagonal (Householder), all three cost O(m) per round
// in the basis of T: a solve with T + mu I, and the diagonal of its inverse from the pivots of its LDL^T and
// UDU^T factorisations.  The reduction is one pass of m - 2 reflections instead of the 6 - 10 Jacobi sweeps an
// eigendecomposition takes (eig_jacobi.cu), about a tenth of the arithmetic and no host synchronisation.
//
//   tri_reduce_kernel  one cooperative launch, the matrix rows dealt out cyclically to the CTAs and kept in
//                      SHARED MEMORY for the whole reduction (1 000 columns: 7 rows of 8 KB per CTA on 148 SMs);
//                      per column ONE exchange between the CTAs and no barrier: every CTA forms the reflector
//                      itself from the pivot row, multiplies its own rows by it, publishes those entries of
//                      p = tau A v as self-validating values, collects everybody's, and applies the rank-2
//                      update A -= v w^T + w v^T to its rows.  b~ = Q^T b rides along.
//   tri_bayes_kernel   one CTA: the evidence iteration with T (two threads walk the pivot recurrences from
//                      both ends, the rest is parallel), the stopping rule and the final coefficients in the
//                      original basis by applying the stored reflections in reverse.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#include "tables.h"
#include "tridiag.h"

namespace gs {

constexpr int kTriThreads = 512;
constexpr int kTriRowsPerCta = 4;   // target when fewer CTAs than SMs suffice
constexpr int kBrMaxIterT = 300;    // BayesianRidge defaults (as complete.cu)
constexpr double kBrTolT = 1e-3;
constexpr double kBrPriorT = 1e-6;

struct TriBuf {
  double* Vh;      // [m][ld] reflector k in row k, entries k+1 .. m-1 (entry k+1 is 1)
  double* tau;     // [ld]
  double* dT;      // [ld] diagonal of T
  double* eT;      // [ld] off-diagonal of T (eT[i] couples i and i+1)
  double* bt;      // [ld] Q^T b
  double* xbuf;    // [2][ld][2] exchange slots, tagged (tri_publish)
  long long* prof; // [16] (GS_TRI_PROFILE builds)
};

static inline int tri_ld(int m) { return (m + 7) & ~7; }

size_t tri_work_doubles(int m) {
  const size_t ld = (size_t)tri_ld(m);
  return (size_t)m * ld + 8 * ld + 16 + 8;
}

static TriBuf tri_carve(double* work, int m) {
  const size_t ld = (size_t)tri_ld(m);
  TriBuf b;
  b.Vh = work;
  b.tau = work + (size_t)m * ld;
  b.dT = b.tau + ld;
  b.eT = b.dT + ld;
  b.bt = b.eT + ld;
  b.xbuf = b.bt + ld;
  b.prof = reinterpret_cast<long long*>(b.xbuf + 4 * ld);
  return b;
}

static inline int tri_grid(int m, int n_sm) {
  static const int rows = [] {
    const char* v = getenv("GS_TRI_ROWS");  // measurement
    const int r = v ? atoi(v) : kTriRowsPerCta;
    return r < 1 ? 1 : r;
  }();
  int g = (m + rows - 1) / rows;
  if (g > n_sm) g = n_sm;
  return g < 1 ? 1 : g;
}

static inline size_t tri_reduce_smem(int m, int grid) {
  const size_t ld = (size_t)tri_ld(m);
  const size_t rows = (size_t)((m + grid - 1) / grid);
  return sizeof(double) * ((rows + 4) * ld + 128);
}

int tri_max_m(int n_sm) {
  // rows + 4 vectors of ld doubles within 200 KB of shared memory per CTA
  int m = 32;
  while (m < 4096 && tri_reduce_smem(m + 32, tri_grid(m + 32, n_sm)) <= 200 * 1024) m += 32;
  return m;
}

// sums of a and b over the CTA, returned to every thread; the same order of additions in every CTA.
// Callers alternate between two scratch areas, or otherwise make sure s_red is no longer being read.
__device__ __forceinline__ void tri_sum2(double& a, double& b, double* s_red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if (lane == 0) {
    s_red[warp] = a;
    s_red[32 + warp] = b;
  }
  __syncthreads();
  a = lane < nw ? s_red[lane] : 0.0;
  b = lane < nw ? s_red[32 + lane] : 0.0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
  }
}

__device__ __forceinline__ void tri_sum1(double& a, double* s_red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if (lane == 0) s_red[warp] = a;
  __syncthreads();
  a = lane < nw ? s_red[lane] : 0.0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
}

// ---- the exchange between CTAs: values that carry their own "ready" mark ----
// A published double gives up its lowest mantissa bit to a tag that flips every time its slot is reused; a reader
// spins on the slot until the tag is the one it expects.  The value IS the signal: no flag, no fence, no barrier
// — about one trip through L2 per column.  Every CTA, the writer included, uses the value as published.
// Measured on B200 (profiles/ubench/ub_exchange.cu, 148 CTAs x 999 values): 1.2 us per exchange this way,
// 2.2 us with cooperative-groups grid.sync(), 6.6 us with per-CTA release/acquire flags.  Every CTA reads the
// whole vector, 148 x 16 KB per column, so a round of polls is about half a microsecond of L2 bandwidth: polls
// that start before the data can have arrived delay the very stores they wait for (the reduction ran 1.5x
// slower per column with them), hence the sleep before the first poll.
constexpr unsigned kTriFirstPollNs = 600;  // GS_TRI_SLEEP overrides (measurement)

__device__ __forceinline__ void tri_publish(double* slot, double p, double a, unsigned tag) {
  const unsigned long long pb = ((unsigned long long)__double_as_longlong(p) & ~1ull) | tag;
  const unsigned long long ab = ((unsigned long long)__double_as_longlong(a) & ~1ull) | tag;
  asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(slot), "l"(pb), "l"(ab) : "memory");
}
__device__ __forceinline__ void tri_poll(const double* slot, unsigned long long* pb, unsigned long long* ab) {
  asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(*pb), "=l"(*ab) : "l"(slot) : "memory");
}
__device__ __forceinline__ double tri_untag(unsigned long long bits) { return __longlong_as_double((long long)(bits & ~1ull)); }

__device__ __forceinline__ double tri_rcp(double x);

// Householder reduction of A = (G without row/column j) to tridiagonal form, LAPACK dsytd2's arithmetic
// (H = I - tau v v^T, v[0] = 1; p = tau A v; w = p - (tau/2)(p.v) v; A -= v w^T + w v^T), full rows instead of
// one triangle.  Row i lives in CTA i % gridDim.x.  Loop state at step k: s_r = row k of the current matrix in
// every CTA.  Per step a CTA publishes, for each of its rows i, p_i and the entry A[i][k+1] as it is before the
// update (by symmetry the gathered entries are row k+1), collects everybody's, and forms w and the next pivot row
// itself.  Two sets of slots suffice: a CTA can be two steps ahead of another only after that one has published
// the step in between, which it does after it has finished reading.  Thread t owns the vector elements
// k+1+t and k+1+t+blockDim.x (m <= 2 blockDim.x).
__global__ void __launch_bounds__(kTriThreads, 1) tri_reduce_kernel(const double* __restrict__ G, int d, int j, TriBuf tb, unsigned first_poll_ns) {
  extern __shared__ double sm[];
  const int m = d - 1, ld = (m + 7) & ~7;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nt = blockDim.x, nw = nt >> 5;
  const int nc = gridDim.x, c = blockIdx.x;
  const int R = (m + nc - 1) / nc;  // local row lr <-> global row c + nc * lr
  double* s_rows = sm;
  double* s_r = s_rows + (size_t)R * ld;
  double* s_v = s_r + ld;
  double* s_p = s_v + ld;
  double* s_b = s_p + ld;
  double* s_red = s_b + ld;  // [2][64]
  auto gi = [&](int x) { return x < j ? x : x + 1; };
  for (int lr = 0; lr < R; ++lr) {
    const int i = c + nc * lr;
    if (i >= m) break;
    for (int jj = tid; jj < m; jj += nt) s_rows[(size_t)lr * ld + jj] = G[(size_t)gi(i) * d + gi(jj)];
  }
  for (int jj = tid; jj < m; jj += nt) {
    s_r[jj] = G[(size_t)gi(0) * d + gi(jj)];
    s_b[jj] = G[(size_t)gi(jj) * d + j];  // X^T y
  }
  __syncthreads();
  unsigned red_flip = 0;
#ifdef GS_TRI_PROFILE
  long long tacc[6] = {0, 0, 0, 0, 0, 0}, t0 = clock64(), t1;
#define TRI_T(i) { t1 = clock64(); tacc[i] += t1 - t0; t0 = t1; }
#else
#define TRI_T(i)
#endif
  for (int k = 0; k + 2 < m; ++k) {
    double* slots = tb.xbuf + (size_t)(k & 1) * 2 * ld;  // [ld] pairs (p_i, A[i][k+1])
    const unsigned tag = (unsigned)(k >> 1) & 1u;
    const int j0 = k + 1 + tid, j1 = j0 + nt;
    const bool h0 = j0 < m, h1 = j1 < m;
    // the reflector that zeroes row k beyond k+1 (dlarfg), and v.b for b~
    const double x0 = s_r[k + 1], b0 = s_b[k + 1];
    const double r0 = h0 ? s_r[j0] : 0.0, r1 = h1 ? s_r[j1] : 0.0;
    const double bb0 = h0 ? s_b[j0] : 0.0, bb1 = h1 ? s_b[j1] : 0.0;
    double ss = tid ? r0 * r0 : 0.0, xb = tid ? r0 * bb0 : 0.0;  // from k+2 on
    ss = fma(r1, r1, ss);
    xb = fma(r1, bb1, xb);
    tri_sum2(ss, xb, s_red + 64 * (red_flip ^= 1));
    double tau = 0.0, beta = x0, scale = 0.0;
    if (ss > 0.0) {
      // 1 / |x| by Newton from the hardware estimate, then beta = -sign(x0) |x|, tau = (beta - x0) / beta
      const double n2 = fma(x0, x0, ss);
      double rs;
      asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(rs) : "d"(n2));
      rs = rs * fma(-0.5 * n2, rs * rs, 1.5);
      rs = rs * fma(-0.5 * n2, rs * rs, 1.5);
      double nrm = n2 * rs;
      nrm = fma(fma(-nrm, nrm, n2), 0.5 * rs, nrm);
      beta = -copysign(nrm, x0);
      tau = (beta - x0) * -copysign(rs, x0);
      scale = tri_rcp(x0 - beta);
    }
    const double tvb = tau * fma(scale, xb, b0);  // tau (v . b)
    const bool keeper = c == k % nc;               // this CTA records step k
    const double v0 = tid ? r0 * scale : 1.0, v1 = r1 * scale;
    if (h0) {
      s_v[j0] = v0;
      s_b[j0] = fma(-tvb, v0, bb0);
      if (keeper) tb.Vh[(size_t)k * ld + j0] = v0;
    }
    if (h1) {
      s_v[j1] = v1;
      s_b[j1] = fma(-tvb, v1, bb1);
      if (keeper) tb.Vh[(size_t)k * ld + j1] = v1;
    }
    if (keeper && tid == 0) {
      tb.tau[k] = tau;
      tb.dT[k] = s_r[k];
      tb.eT[k] = beta;
    }
    __syncthreads();
    TRI_T(0)
    // p = tau A v for the rows this CTA holds: a warp per row
    for (int lr = warp; lr < R; lr += nw) {
      const int i = c + nc * lr;
      if (i <= k || i >= m) continue;
      const double* row = s_rows + (size_t)lr * ld;
      double a0 = 0.0, a1 = 0.0;
      int jj = k + 1 + lane;
      for (; jj + 32 < m; jj += 64) {
        a0 = fma(row[jj], s_v[jj], a0);
        a1 = fma(row[jj + 32], s_v[jj + 32], a1);
      }
      if (jj < m) a0 = fma(row[jj], s_v[jj], a0);
      a0 += a1;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) a0 += __shfl_xor_sync(0xffffffffu, a0, o);
      if (lane == 0) tri_publish(slots + 2 * (size_t)i, tau * a0, row[k + 1], tag);
    }
    __syncthreads();  // nobody polls before this CTA has published
    TRI_T(1)
    __nanosleep(first_poll_ns);
    // collect: both of this thread's slots in flight together
    unsigned long long p0 = 0, a0 = 0, p1 = 0, a1 = 0;
    {
      bool d0 = !h0, d1 = !h1;
      unsigned spins = 0;
      while (!(d0 && d1)) {
        if (!d0) tri_poll(slots + 2 * (size_t)j0, &p0, &a0);
        if (!d1) tri_poll(slots + 2 * (size_t)j1, &p1, &a1);
        d0 = d0 || ((p0 & 1ull) == tag && (a0 & 1ull) == tag);
        d1 = d1 || ((p1 & 1ull) == tag && (a1 & 1ull) == tag);
        if (++spins > (1u << 25)) __trap();  // seconds, not microseconds: a peer died; fail instead of hanging
      }
    }
    const double pp0 = h0 ? tri_untag(p0) : 0.0, pp1 = h1 ? tri_untag(p1) : 0.0;
    if (h0) s_p[j0] = pp0;
    if (h1) s_p[j1] = pp1;
    double pv = fma(pp0, v0, pp1 * v1);
    TRI_T(2)
    tri_sum1(pv, s_red + 64 * (red_flip ^= 1));  // its barrier also completes s_p
    const double cw = -0.5 * tau * pv;                  // w = p + cw v
    TRI_T(3)
    const double w1 = s_p[k + 1] + cw;  // v[k+1] = 1
    const double wj0 = fma(cw, v0, pp0), wj1 = fma(cw, v1, pp1);
    // products first, then one sum: A[i][jj] and A[jj][i] stay equal bit for bit, in whichever CTA they live.
    // Row k+1 after the update is the next pivot row.
    if (h0) s_r[j0] = tri_untag(a0) - (wj0 + w1 * v0);
    if (h1) s_r[j1] = tri_untag(a1) - (wj1 + w1 * v1);
    for (int lr = 0; lr < R; ++lr) {
      const int i = c + nc * lr;
      if (i <= k || i >= m) continue;
      const double vi = s_v[i], wi = fma(cw, vi, s_p[i]);
      double* row = s_rows + (size_t)lr * ld;
      if (h0) row[j0] = row[j0] - (vi * wj0 + wi * v0);
      if (h1) row[j1] = row[j1] - (vi * wj1 + wi * v1);
    }
    __syncthreads();
    TRI_T(4)
  }
#ifdef GS_TRI_PROFILE
  if (tid == 0 && (c == 0 || c == nc - 1))
    for (int i = 0; i < 5; ++i) tb.prof[(c ? 8 : 0) + i] = tacc[i];
#endif
  // the last 2 x 2 block, and b~
  if (c == 0) {
    if (tid == 0) {
      if (m >= 2) {
        tb.dT[m - 2] = s_r[m - 2];
        tb.eT[m - 2] = s_r[m - 1];
      } else {
        tb.dT[0] = s_r[0];
      }
    }
    for (int jj = tid; jj < m; jj += nt) tb.bt[jj] = s_b[jj];
  }
  if (m >= 2 && (m - 1) % nc == c && tid == 0) tb.dT[m - 1] = s_rows[(size_t)((m - 1) / nc) * ld + m - 1];
}

__device__ __forceinline__ double tri_rcp(double x) {  // 1 / x for normal x > 0, to rounding
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  y = fma(y, fma(-x, y, 1.0), y);
  y = fma(y, fma(-x, y, 1.0), y);
  return y;
}

constexpr int kTriChunk = 8;  // elements of the recurrences fetched ahead of the dependent chain

// The evidence iteration (_bayes.py:262-306) with T in place of the eigenvalues.  Per round, mu = lambda/alpha:
//   (T + mu I) x = b~ by LDL^T: pivots dl[i] = t_ii + mu - e[i-1]^2 / dl[i-1]             (thread 0, forwards)
//   the UDU^T pivots rho[i] = t_ii + mu - e[i]^2 / rho[i+1]                              (thread 32, backwards)
//   ((T + mu I)^-1)_ii = 1 / (dl[i] + rho[i] - t_ii - mu)   ->   gamma = m - mu * trace
//   |coef|^2 = |x|^2,   sse = y.y - x.b~ - mu |x|^2   (x^T T x = x.(b~ - mu x)).
// The stopping rule sum |coef_old - coef| < tol lives in the original basis: ||.||_2 <= ||.||_1 <= sqrt(m) ||.||_2
// and Q is orthogonal, so Q (x_old - x) is formed (m - 2 reflections in reverse) only while ||x_old - x||_2 lies in
// [tol / sqrt(m), tol).  Thread t owns elements t and t + blockDim.x of every vector (m <= 2 blockDim.x).
__global__ void __launch_bounds__(kTriThreads, 1) tri_bayes_kernel(const double* __restrict__ G, const double* __restrict__ mu,
                                                                  const int* __restrict__ n_obs_p, int d, int j, TriBuf tb,
                                                                  double* __restrict__ coef_out,
                                                                  double* __restrict__ intercept_out,
                                                                  int* __restrict__ iters_out) {
  extern __shared__ double sm[];
  const int m = d - 1, ld = ((m + 7) & ~7) + 2 * kTriChunk;
  const int tid = threadIdx.x;
  double* s_d = sm;             // t_ii
  double* s_e = s_d + ld;       // e[i] couples i and i+1; 0 from m-1 on
  double* s_bt = s_e + ld;
  double* s_inv = s_bt + ld;    // 1 / dl[i]
  double* s_dl = s_inv + ld;
  double* s_rho = s_dl + ld;
  double* s_z = s_rho + ld;
  double* s_x = s_z + ld;
  double* s_tau = s_x + ld;
  double* s_red = s_tau + ld;   // [2][64]
  unsigned red_flip = 0;
  const size_t vld = (size_t)((m + 7) & ~7);
  for (int i = tid; i < ld; i += kTriThreads) {
    s_d[i] = i < m ? tb.dT[i] : 0.0;
    s_e[i] = i < m - 1 ? tb.eT[i] : 0.0;
    s_bt[i] = i < m ? tb.bt[i] : 0.0;
    s_tau[i] = i < m - 2 ? tb.tau[i] : 0.0;
  }
  __syncthreads();
  const int e0 = tid, e1 = tid + kTriThreads;
  const bool has0 = e0 < m, has1 = e1 < m;
  const double n = (double)*n_obs_p;
  const double yy = G[(size_t)j * d + j];
  const double eps = 2.220446049250313e-16;

  auto solve = [&](double mu_) {
    if (tid == 0) {
      double dl = s_d[0] + mu_, inv = tri_rcp(dl), z = s_bt[0];
      s_dl[0] = dl;
      s_inv[0] = inv;
      s_z[0] = z;
      for (int i0 = 1; i0 < m; i0 += kTriChunk) {
        double e[kTriChunk], a[kTriChunk], bb[kTriChunk];
#pragma unroll
        for (int t = 0; t < kTriChunk; ++t) {  // padded: reading past m - 1 is harmless
          e[t] = s_e[i0 + t - 1];
          a[t] = s_d[i0 + t] + mu_;
          bb[t] = s_bt[i0 + t];
        }
#pragma unroll
        for (int t = 0; t < kTriChunk; ++t) {
          const double l = e[t] * inv;
          dl = fma(-l, e[t], a[t]);
          z = fma(-l, z, bb[t]);
          inv = tri_rcp(dl);
          if (i0 + t < m) {
            s_dl[i0 + t] = dl;
            s_inv[i0 + t] = inv;
            s_z[i0 + t] = z;
          }
        }
      }
    } else if (tid == 32) {
      double rho = s_d[m - 1] + mu_;
      s_rho[m - 1] = rho;
      for (int i0 = m - 2; i0 >= 0; i0 -= kTriChunk) {
        double e2[kTriChunk], a[kTriChunk];
#pragma unroll
        for (int t = 0; t < kTriChunk; ++t) {
          const int i = i0 - t < 0 ? 0 : i0 - t;
          e2[t] = s_e[i] * s_e[i];
          a[t] = s_d[i] + mu_;
        }
#pragma unroll
        for (int t = 0; t < kTriChunk; ++t) {
          if (i0 - t >= 0) {
            rho = fma(-e2[t], tri_rcp(rho), a[t]);
            s_rho[i0 - t] = rho;
          }
        }
      }
    }
    __syncthreads();
    if (tid == 0) {
      double x = s_z[m - 1] * s_inv[m - 1];
      s_x[m - 1] = x;
      for (int i0 = m - 2; i0 >= 0; i0 -= kTriChunk) {
        double e[kTriChunk], z[kTriChunk], iv[kTriChunk];
#pragma unroll
        for (int t = 0; t < kTriChunk; ++t) {
          const int i = i0 - t < 0 ? 0 : i0 - t;
          e[t] = s_e[i];
          z[t] = s_z[i];
          iv[t] = s_inv[i];
        }
#pragma unroll
        for (int t = 0; t < kTriChunk; ++t) {
          if (i0 - t >= 0) {
            x = fma(-e[t], x, z[t]) * iv[t];
            s_x[i0 - t] = x;
          }
        }
      }
    }
    __syncthreads();
  };

  // vec <- Q vec = H_0 H_1 ... H_{m-3} vec; thread-owned elements in registers.  One reduction per reflection
  // is the critical path, so the reflectors' entries are fetched four reflections ahead of it.
  auto load_v = [&](int k, double* a, double* b) {
    *a = (k >= 0 && has0 && e0 >= k + 1) ? tb.Vh[(size_t)k * vld + e0] : 0.0;
    *b = (k >= 0 && has1 && e1 >= k + 1) ? tb.Vh[(size_t)k * vld + e1] : 0.0;
  };
  auto apply_q = [&](double& r0, double& r1) {
    constexpr int kAhead = 4;
    double nv0[kAhead], nv1[kAhead];
#pragma unroll
    for (int t = 0; t < kAhead; ++t) load_v(m - 3 - t, &nv0[t], &nv1[t]);
    for (int k = m - 3; k >= 0; k -= kAhead) {
      double v0[kAhead], v1[kAhead];
#pragma unroll
      for (int t = 0; t < kAhead; ++t) {
        v0[t] = nv0[t];
        v1[t] = nv1[t];
        load_v(k - kAhead - t, &nv0[t], &nv1[t]);
      }
#pragma unroll
      for (int t = 0; t < kAhead; ++t) {
        if (k - t < 0) break;
        double dot = fma(v0[t], r0, v1[t] * r1);
        tri_sum1(dot, s_red + 64 * (red_flip ^= 1));
        const double f = s_tau[k - t] * dot;
        r0 = fma(-f, v0[t], r0);
        r1 = fma(-f, v1[t], r1);
      }
    }
  };

  double alpha = 1.0 / (yy / n + eps);  // 1 / (var(y) + eps), population variance
  double lambda = 1.0;
  double old0 = 0.0, old1 = 0.0;  // the previous round's x
  const double sqrt_m = sqrt((double)m);
  int it = 0;
  for (; it < kBrMaxIterT; ++it) {
    const double mu_ = lambda / alpha;
    solve(mu_);
    const double x0 = has0 ? s_x[e0] : 0.0, x1 = has1 ? s_x[e1] : 0.0;
    double tr = 0.0, xb = 0.0, sq = 0.0, d2 = 0.0;
    if (has0) {
      tr += 1.0 / (s_dl[e0] + s_rho[e0] - (s_d[e0] + mu_));
      xb = fma(x0, s_bt[e0], xb);
      sq = fma(x0, x0, sq);
      d2 = fma(old0 - x0, old0 - x0, d2);
    }
    if (has1) {
      tr += 1.0 / (s_dl[e1] + s_rho[e1] - (s_d[e1] + mu_));
      xb = fma(x1, s_bt[e1], xb);
      sq = fma(x1, x1, sq);
      d2 = fma(old1 - x1, old1 - x1, d2);
    }
    tri_sum2(tr, xb, s_red + 64 * (red_flip ^= 1));
    tri_sum2(sq, d2, s_red + 64 * (red_flip ^= 1));
    const double gamma = (double)m - mu_ * tr;
    const double sse = fmax(yy - xb - mu_ * sq, 0.0);
    const double dnorm = sqrt(d2);
    lambda = (gamma + 2.0 * kBrPriorT) / (sq + 2.0 * kBrPriorT);
    alpha = (n - gamma + 2.0 * kBrPriorT) / (sse + 2.0 * kBrPriorT);
    if (it != 0) {
      bool converged = sqrt_m * dnorm < kBrTolT;
      if (!converged && dnorm < kBrTolT) {  // undecided: evaluate the rule itself
        double q0 = old0 - x0, q1 = old1 - x1;
        apply_q(q0, q1);
        double l1 = fabs(q0) + fabs(q1);
        tri_sum1(l1, s_red + 64 * (red_flip ^= 1));
        converged = l1 < kBrTolT;
      }
      if (converged) break;
    }
    old0 = x0;
    old1 = x1;
  }
  solve(lambda / alpha);  // the coefficients with the final alpha, lambda
  double c0 = has0 ? s_x[e0] : 0.0, c1 = has1 ? s_x[e1] : 0.0;
  apply_q(c0, c1);
  double dot = 0.0;
  if (has0) {
    coef_out[e0] = c0;
    dot = fma(mu[e0 < j ? e0 : e0 + 1], c0, dot);
  }
  if (has1) {
    coef_out[e1] = c1;
    dot = fma(mu[e1 < j ? e1 : e1 + 1], c1, dot);
  }
  tri_sum1(dot, s_red + 64 * (red_flip ^= 1));
  if (tid == 0) {
    *intercept_out = mu[j] - dot;
    *iters_out = it < kBrMaxIterT ? it + 1 : kBrMaxIterT;
  }
}

static inline size_t tri_bayes_smem(int m) { return sizeof(double) * (9 * (size_t)(tri_ld(m) + 2 * kTriChunk) + 128); }

cudaError_t tri_bayes_ridge(const double* G, const double* mu, const int* n_obs, int d, int j, double* work, int n_sm,
                            double* coef_out, double* intercept_out, int* iters_out, cudaStream_t s) {
  const int m = d - 1;
  if (m < 3 || m > 2 * kTriThreads || m > tri_max_m(n_sm)) return cudaErrorInvalidValue;
  TriBuf tb = tri_carve(work, m);
  const int grid = tri_grid(m, n_sm);
  const size_t smem = tri_reduce_smem(m, grid);
  const size_t bsmem = tri_bayes_smem(m);
  cudaError_t e;
  // per call: the attribute belongs to the current device, and a thread may drive several
  if ((e = cudaFuncSetAttribute(tri_reduce_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
  if ((e = cudaFuncSetAttribute(tri_bayes_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bsmem)) != cudaSuccess) return e;
  // every slot starts with the tag the first two steps do NOT expect (all ones: lowest bit 1, they wait for 0)
  if ((e = cudaMemsetAsync(tb.xbuf, 0xFF, sizeof(double) * 4 * (size_t)tri_ld(m), s)) != cudaSuccess) return e;
  int dd = d, jj = j;
  static const unsigned first_ns = getenv("GS_TRI_SLEEP") ? (unsigned)atoi(getenv("GS_TRI_SLEEP")) : kTriFirstPollNs;  // measurement
  unsigned fns = first_ns;
  void* args[] = {(void*)&G, (void*)&dd, (void*)&jj, (void*)&tb, (void*)&fns};
  note_launch();
  if ((e = cudaLaunchCooperativeKernel((const void*)tri_reduce_kernel, dim3((unsigned)grid), dim3(m <= 256 ? 256 : kTriThreads), args, smem, s)) !=
      cudaSuccess)
    return e;
#ifdef GS_TRI_PROFILE
  {
    static int shown = 0;
    if (shown < 4 || (m > 500 && shown < 8)) {
      ++shown;
      long long h[16];
      cudaStreamSynchronize(s);
      cudaMemcpy(h, tb.prof, sizeof h, cudaMemcpyDeviceToHost);
      fprintf(stderr, "tri profile m=%d grid=%d cycles/col  cta0: a %.0f b %.0f wait %.0f c %.0f d %.0f | last: a %.0f b %.0f wait %.0f c %.0f d %.0f\n",
              m, grid, h[0] / (double)m, h[1] / (double)m, h[2] / (double)m, h[3] / (double)m, h[4] / (double)m,
              h[8] / (double)m, h[9] / (double)m, h[10] / (double)m, h[11] / (double)m, h[12] / (double)m);
    }
  }
#endif
  note_launch();
  tri_bayes_kernel<<<1, kTriThreads, bsmem, s>>>(G, mu, n_obs, d, j, tb, coef_out, intercept_out, iters_out);
  return cudaGetLastError();
}

}  // namespace gs
