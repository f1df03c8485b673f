// complete.cu — recommender completion on the GPU: fit / transform of the reference's
// Imputer = sklearn IterativeImputer(max_iter=10, random_state=0) with BayesianRidge
// (pkg/recommender/recommender/recommender.py:15-27).  The arithmetic restated here is
// scikit-learn's (impute/_iterative.py:723-927, linear_model/_bayes.py:217-428 in 1.9.0),
// reformulated for the GPU:
//   * per feature j the estimator needs only second moments of the rows where j is observed:
//     the centred Gram matrix G of ALL d columns over those rows (gram_* kernels, float64).
//     X^T X = G[others][others], X^T y = G[others][j], y^T y = G[j][j], var(y) = G[j][j]/n;
//   * up to 24 features (the reference's own tables) the SVD of X in BayesianRidge.fit becomes the symmetric
//     eigendecomposition of X^T X (parallel cyclic Jacobi, one CTA): eigen_vals = S^2, Vh = V^T, and the
//     alpha/lambda fixed-point iteration (<= 300 rounds) costs O(d) per round in the eigenbasis:
//     coef~ = (V^T X^T y)/(s + lambda/alpha), coef = V coef~, sse = y^T y - 2 coef~.b~ + sum(coef~^2 s);
//   * wider tables need no eigendecomposition: tridiag.cu reduces X^T X to tridiagonal form and runs the same
//     iteration there (eig_jacobi.cu, a block Jacobi eigendecomposition across CTAs, is the cross-check);
//   * rows where j is missing get x.coef + intercept (predict_kernel).
// Everything is float64; parity bar: 1e-5 relative against Imputer.predict (tests/golden).

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <mutex>
#include <new>
#include <vector>

#include "eig_jacobi.h"
#include "tridiag.h"

#include "../../include/gpusched.h"
#include "context.h"

namespace gs {

constexpr int kMaxFeatures = 1024; // d - 1 <= 1023 (shared arrays of the evidence iteration)
constexpr int kJacobiMax = 24;     // d - 1 <= 24: one-CTA Jacobi (covers the reference tables); beyond: block Jacobi (eig_jacobi.cu)
constexpr int kGramTile = 32;
constexpr int kRounds = 10;        // IterativeImputer(max_iter=10), recommender.py:16
constexpr double kRoundTol = 1e-3; // IterativeImputer tol
constexpr int kBrMaxIter = 300;    // BayesianRidge defaults
constexpr double kBrTol = 1e-3;
constexpr double kBrPrior = 1e-6;  // alpha_1 = alpha_2 = lambda_1 = lambda_2

// ---- initial imputation: column means over the observed entries (SimpleImputer "mean") ----
// stats = [sum(d) | count(d) | n_rows(1) | max_abs(d)]: the first 2d+1 are summed over ranks,
// the last d maximised, before finish_means_kernel turns them into means.
__global__ void __launch_bounds__(256) column_stats_kernel(const double* __restrict__ X, int n, int d,
                                                          double* __restrict__ stats) {
  __shared__ double s_sum[256];
  __shared__ double s_max[256];
  __shared__ int s_cnt[256];
  const int c = blockIdx.x;
  double sum = 0.0, mx = 0.0;
  int cnt = 0;
  for (int r = threadIdx.x; r < n; r += blockDim.x) {
    double v = X[(size_t)r * d + c];
    if (v == v) {
      sum += v;
      mx = fmax(mx, fabs(v));
      ++cnt;
    }
  }
  s_sum[threadIdx.x] = sum;
  s_max[threadIdx.x] = mx;
  s_cnt[threadIdx.x] = cnt;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      s_sum[threadIdx.x] += s_sum[threadIdx.x + s];
      s_max[threadIdx.x] = fmax(s_max[threadIdx.x], s_max[threadIdx.x + s]);
      s_cnt[threadIdx.x] += s_cnt[threadIdx.x + s];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    stats[c] = s_sum[0];
    stats[d + c] = (double)s_cnt[0];
    stats[2 * d + 1 + c] = s_max[0];
    if (c == 0) stats[2 * d] = (double)n;
  }
}

__global__ void finish_means_kernel(const double* __restrict__ stats, int d, double* __restrict__ means) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < d) means[c] = stats[d + c] > 0.0 ? stats[c] / stats[d + c] : 0.0;
}

__global__ void fill_kernel(const double* __restrict__ X, const double* __restrict__ means, size_t total, int d,
                            double* __restrict__ Xt, uint8_t* __restrict__ mask) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  double v = X[i];
  bool miss = !(v == v);
  mask[i] = miss ? 1 : 0;
  Xt[i] = miss ? means[i % d] : v;
}

// ---- moments over the rows where feature j is observed ----
// Column sums (all d columns) + row count as sums[d].  Rows are read once, coalesced: a CTA takes
// a block of rows with thread t on columns t, t+256, ...; per-block partials are then added in
// block order (deterministic).  Summed over ranks before finish_mu_kernel divides.
constexpr int kSumRows = 256;  // rows per CTA
constexpr int kSumColsPerThread = (kMaxFeatures + 255) / 256;

__global__ void __launch_bounds__(256) obs_sums_partial_kernel(const double* __restrict__ Xt,
                                                              const uint8_t* __restrict__ mask, int n, int d, int j,
                                                              double* __restrict__ partial /*[blocks][d+1]*/) {
  __shared__ uint8_t s_live[kSumRows];
  __shared__ int s_cnt;
  const int r0 = blockIdx.x * kSumRows;
  const int rows = min(kSumRows, n - r0);
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  int live = 0;
  if ((int)threadIdx.x < rows) live = !mask[(size_t)(r0 + threadIdx.x) * d + j];
  s_live[threadIdx.x] = (uint8_t)live;
  if (live) atomicAdd(&s_cnt, 1);
  __syncthreads();
  double acc[kSumColsPerThread];
#pragma unroll
  for (int q = 0; q < kSumColsPerThread; ++q) acc[q] = 0.0;
  for (int r = 0; r < rows; ++r) {
    if (!s_live[r]) continue;
    const double* row = Xt + (size_t)(r0 + r) * d;
#pragma unroll
    for (int q = 0; q < kSumColsPerThread; ++q) {
      const int c = threadIdx.x + q * 256;
      if (c < d) acc[q] += row[c];
    }
  }
  double* out = partial + (size_t)blockIdx.x * (d + 1);
#pragma unroll
  for (int q = 0; q < kSumColsPerThread; ++q) {
    const int c = threadIdx.x + q * 256;
    if (c < d) out[c] = acc[q];
  }
  if (threadIdx.x == 0) out[d] = (double)s_cnt;
}

__global__ void obs_sums_reduce_kernel(const double* __restrict__ partial, int blocks, int d,
                                       double* __restrict__ sums) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c > d) return;
  double s = 0.0;
  for (int b = 0; b < blocks; ++b) s += partial[(size_t)b * (d + 1) + c];
  sums[c] = s;
}

__global__ void finish_mu_kernel(const double* __restrict__ sums, int d, double* __restrict__ mu,
                                 int* __restrict__ n_obs) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  const double cnt = sums[d];
  if (c < d) mu[c] = cnt > 0.0 ? sums[c] / cnt : 0.0;
  if (c == 0) *n_obs = (int)cnt;
}

// out[0] = max(in[0..n))
__global__ void __launch_bounds__(256) max_small_kernel(double* __restrict__ v, int n) {
  __shared__ double s_red[256];
  double mx = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) mx = fmax(mx, v[i]);
  s_red[threadIdx.x] = mx;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) s_red[threadIdx.x] = fmax(s_red[threadIdx.x], s_red[threadIdx.x + s]);
    __syncthreads();
  }
  if (threadIdx.x == 0) v[0] = s_red[0];
}

// Partial centred Gram: grid = (tiles_b, tiles_a, row_parts); CTA 16x16 threads, each a 2x2
// block of the 32x32 output tile; rows of the part staged 32 at a time in shared memory,
// centred, rows where j is missing zeroed.  Only tiles with tile_b >= tile_a are computed.
__global__ void __launch_bounds__(256) gram_partial_kernel(const double* __restrict__ Xt,
                                                          const uint8_t* __restrict__ mask, int n, int d, int j,
                                                          const double* __restrict__ mu, int rows_per_part,
                                                          double* __restrict__ partial) {
  if (blockIdx.x < blockIdx.y) return;
  __shared__ double sa[kGramTile][kGramTile + 1];  // [row][col a]
  __shared__ double sb[kGramTile][kGramTile + 1];
  const int a0 = blockIdx.y * kGramTile, b0 = blockIdx.x * kGramTile;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int r_begin = blockIdx.z * rows_per_part;
  const int r_end = min(n, r_begin + rows_per_part);
  double acc[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
  for (int r0 = r_begin; r0 < r_end; r0 += kGramTile) {
    // 256 threads load 32 rows x 32 cols for each tile: thread -> (row = t / 8 .. , 4 cols)
    for (int e = threadIdx.x; e < kGramTile * kGramTile; e += 256) {
      const int rr = e / kGramTile, cc = e % kGramTile;
      const int r = r0 + rr;
      const bool live = r < r_end && !mask[(size_t)r * d + j];
      const int ca = a0 + cc, cb = b0 + cc;
      sa[rr][cc] = (live && ca < d) ? Xt[(size_t)r * d + ca] - mu[ca] : 0.0;
      sb[rr][cc] = (live && cb < d) ? Xt[(size_t)r * d + cb] - mu[cb] : 0.0;
    }
    __syncthreads();
#pragma unroll 8
    for (int rr = 0; rr < kGramTile; ++rr) {
      const double x0 = sa[rr][ty * 2], x1 = sa[rr][ty * 2 + 1];
      const double y0 = sb[rr][tx * 2], y1 = sb[rr][tx * 2 + 1];
      acc[0][0] = fma(x0, y0, acc[0][0]);
      acc[0][1] = fma(x0, y1, acc[0][1]);
      acc[1][0] = fma(x1, y0, acc[1][0]);
      acc[1][1] = fma(x1, y1, acc[1][1]);
    }
    __syncthreads();
  }
  double* out = partial + (size_t)blockIdx.z * d * d;
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int a = a0 + ty * 2 + i, b = b0 + tx * 2 + k;
      if (a < d && b < d) out[(size_t)a * d + b] = acc[i][k];
    }
}

// G[a][b] = sum over parts (fixed order: deterministic), mirrored to the lower triangle.  With
// `whole` (the Gram over all rows) the parts hold the rows to leave out: G = whole - sum.
__global__ void gram_reduce_kernel(const double* __restrict__ partial, int d, int parts, int tile,
                                   const double* __restrict__ whole, double* __restrict__ G) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= d * d) return;
  int a = idx / d, b = idx % d;
  if (b / tile < a / tile) return;  // tile not computed; filled by its mirror
  double s = 0.0;
  for (int p = 0; p < parts; ++p) s += partial[(size_t)p * d * d + idx];
  if (whole) s = whole[idx] - s;
  G[idx] = s;
  if (a / tile != b / tile) G[(size_t)b * d + a] = s;
}

// Wide tables keep every Gram centred at one fixed vector m (the initial column means), so that
// sums over row sets can be added and subtracted; the centring the estimator wants follows from
//   sum_S (x-mu)(x-mu)^T = sum_S (x-m)(x-m)^T - |S| (mu-m)(mu-m)^T,  mu = mean over S.
__global__ void recentre_kernel(double* __restrict__ G, int d, const double* __restrict__ mu,
                                const double* __restrict__ m, const int* __restrict__ n_obs) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= d * d) return;
  int a = idx / d, b = idx % d;
  G[idx] -= (double)*n_obs * ((mu[a] - m[a]) * (mu[b] - m[b]));
}

// Row j of the all-rows Gram after column j changed: partial[blk][b] = sum over the block's rows of
// (x[r][j]-m[j]) (x[r][b]-m[b]); h_row_reduce_kernel sums the blocks in order into H[j][.] and H[.][j].
constexpr int kHRows = 128;
__global__ void __launch_bounds__(256) h_row_partial_kernel(const double* __restrict__ Xt, int n, int d, int j,
                                                           const double* __restrict__ m,
                                                           double* __restrict__ partial /*[blocks][d]*/) {
  const int r0 = blockIdx.x * kHRows;
  const int rows = min(kHRows, n - r0);
  double acc[kSumColsPerThread], mc[kSumColsPerThread];
#pragma unroll
  for (int q = 0; q < kSumColsPerThread; ++q) {
    const int c = threadIdx.x + q * 256;
    acc[q] = 0.0;
    mc[q] = c < d ? m[c] : 0.0;
  }
  const double mj = m[j];
#pragma unroll 4
  for (int r = 0; r < rows; ++r) {
    const double* row = Xt + (size_t)(r0 + r) * d;
    const double xj = row[j] - mj;
#pragma unroll
    for (int q = 0; q < kSumColsPerThread; ++q) {
      const int c = threadIdx.x + q * 256;
      if (c < d) acc[q] = fma(xj, row[c] - mc[q], acc[q]);
    }
  }
#pragma unroll
  for (int q = 0; q < kSumColsPerThread; ++q) {
    const int c = threadIdx.x + q * 256;
    if (c < d) partial[(size_t)blockIdx.x * d + c] = acc[q];
  }
}

__global__ void h_row_reduce_kernel(const double* __restrict__ partial, int blocks, int d, int j,
                                    double* __restrict__ H) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= d) return;
  double s = 0.0;
  for (int k = 0; k < blocks; ++k) s += partial[(size_t)k * d + b];
  H[(size_t)j * d + b] = s;
  H[(size_t)b * d + j] = s;
}

// The same partial Gram for wide tables (d >= 128): 128x128 output tile per CTA, 8x8 outputs per
// thread (64 independent DFMA chains), rows streamed 16 at a time through shared memory with the
// next slab prefetched into registers.  Bound: the FP64 pipe (2*128*128*16 flop per 32 KB staged).
// Thread (ty,tx) owns rows {2ty,2ty+1}+32i and columns {2tx,2tx+1}+32i, i=0..3, so that every
// 128-bit shared load of a warp is contiguous (no bank conflicts) or a broadcast.
constexpr int kBT = 128, kBK = 16;

// kList = false: rows 0..n-1, those where feature j is observed (j < 0: every row).
// kList = true: the rows rows[0..n-1] (a list of row numbers), all of them.
// the 64 accumulators seen as [4 row blocks][8 column blocks][2 fragment elements]
#define GS_ACCM(i, jb, c) acc[((i) * 16 + 2 * (jb) + (c)) >> 3][((i) * 16 + 2 * (jb) + (c)) & 7]
template <bool kList, bool kMma>
__global__ void __launch_bounds__(256, 1) gram128_partial_kernel(const double* __restrict__ Xt,
                                                                const uint8_t* __restrict__ mask,
                                                                const int* __restrict__ rows, int n, int d,
                                                                int j, const double* __restrict__ mu,
                                                                int rows_per_part, double* __restrict__ partial) {
  if (blockIdx.x < blockIdx.y) return;
  // rows padded by 4 doubles: the tensor-core fragments read 4 rows x 8 columns per load (a half-warp's
  // sixteen 8-byte words then fall into sixteen different bank pairs)
  __shared__ __align__(16) double sa[kBK][kBT + 4];
  __shared__ __align__(16) double sb[kBK][kBT + 4];
  const int a0 = blockIdx.y * kBT, b0 = blockIdx.x * kBT;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int r_begin = blockIdx.z * rows_per_part;
  const int r_end = min(n, r_begin + rows_per_part);
  // staging role: 2048 elements per matrix and slab, 8 per thread as 4 rows x 2 adjacent columns:
  // consecutive threads take consecutive 16-byte chunks of a row, so that the global loads are
  // coalesced and the 128-bit shared stores conflict-free
  const int lcol = (tid & 63) * 2;   // 2 adjacent columns
  const int lrow0 = tid >> 6;        // rows lrow0 + 4q, q = 0..3
  double acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[i][k] = 0.0;
  double pa[8], pb[8];
  const int ca = a0 + lcol, cb = b0 + lcol;
  const double mua0 = ca < d ? mu[ca] : 0.0, mua1 = ca + 1 < d ? mu[ca + 1] : 0.0;
  const double mub0 = cb < d ? mu[cb] : 0.0, mub1 = cb + 1 < d ? mu[cb + 1] : 0.0;
  // list mode: the row numbers of the slab after the next one are loaded with each fetch, so
  // that a fetch never waits for its own indices
  int nidx[4];
  if (kList) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int r = r_begin + lrow0 + 4 * q;
      nidx[q] = r < r_end ? rows[r] : 0;
    }
  }
  auto fetch = [&](int r0) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int r = r0 + lrow0 + 4 * q;
      const bool in = r < r_end;
      const size_t rr = kList ? (size_t)nidx[q] : (size_t)(in ? r : r_begin);
      const double* row = Xt + rr * d;
      // the row's values are loaded whether or not feature j is observed in it (one level of load
      // latency instead of mask -> value), and zeroed afterwards
      const bool live = kList ? in : (in && (j < 0 || !mask[rr * d + j]));
      const double xa0 = ca < d ? row[ca] : 0.0, xa1 = ca + 1 < d ? row[ca + 1] : 0.0;
      const double xb0 = cb < d ? row[cb] : 0.0, xb1 = cb + 1 < d ? row[cb + 1] : 0.0;
      pa[2 * q] = (live && ca < d) ? xa0 - mua0 : 0.0;
      pa[2 * q + 1] = (live && ca + 1 < d) ? xa1 - mua1 : 0.0;
      pb[2 * q] = (live && cb < d) ? xb0 - mub0 : 0.0;
      pb[2 * q + 1] = (live && cb + 1 < d) ? xb1 - mub1 : 0.0;
      if (kList) nidx[q] = r + kBK < r_end ? rows[r + kBK] : 0;
    }
  };
  auto stage = [&]() {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      *reinterpret_cast<double2*>(&sa[lrow0 + 4 * q][lcol]) = make_double2(pa[2 * q], pa[2 * q + 1]);
      *reinterpret_cast<double2*>(&sb[lrow0 + 4 * q][lcol]) = make_double2(pb[2 * q], pb[2 * q + 1]);
    }
  };
  if (r_begin < r_end) {
    fetch(r_begin);
    stage();
  }
  __syncthreads();
  for (int r0 = r_begin; r0 < r_end; r0 += kBK) {
    const bool more = r0 + kBK < r_end;
    if (more) fetch(r0 + kBK);
    if constexpr (kMma) {
      // FP64 tensor cores: mma.sync m8n8k4.  Warp w owns rows 32*(w>>1).., columns 64*(w&1).. of the tile:
      // 4 x 8 blocks of 8 x 8; GS_ACCM(i, jb, c) = block (i, jb), fragment element c (the same 64 registers).
      const int lane = tid & 31, wrp = tid >> 5;
      const int fa = 32 * (wrp >> 1) + (lane >> 2), fb = 64 * (wrp & 1) + (lane >> 2), fk = lane & 3;
#pragma unroll
      for (int k0 = 0; k0 < kBK; k0 += 4) {
        double af[4], bf[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) af[i] = sa[k0 + fk][fa + 8 * i];
#pragma unroll
        for (int jb = 0; jb < 8; ++jb) bf[jb] = sb[k0 + fk][fb + 8 * jb];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int jb = 0; jb < 8; ++jb)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(GS_ACCM(i, jb, 0)), "+d"(GS_ACCM(i, jb, 1))
                         : "d"(af[i]), "d"(bf[jb]));
      }
    } else {
#pragma unroll
    for (int kk = 0; kk < kBK; ++kk) {
      double av[8], bv[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const double2 a2 = *reinterpret_cast<const double2*>(&sa[kk][2 * ty + 32 * i]);
        const double2 b2 = *reinterpret_cast<const double2*>(&sb[kk][2 * tx + 32 * i]);
        av[2 * i] = a2.x;
        av[2 * i + 1] = a2.y;
        bv[2 * i] = b2.x;
        bv[2 * i + 1] = b2.y;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[i][k] = fma(av[i], bv[k], acc[i][k]);
    }
    }
    __syncthreads();
    if (more) stage();
    __syncthreads();
  }
  double* out = partial + (size_t)blockIdx.z * d * d;
  if constexpr (kMma) {
    const int lane = tid & 31, wrp = tid >> 5;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int a = a0 + 32 * (wrp >> 1) + 8 * i + (lane >> 2);
#pragma unroll
      for (int jb = 0; jb < 8; ++jb)
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const int b = b0 + 64 * (wrp & 1) + 8 * jb + 2 * (lane & 3) + c;
          if (a < d && b < d) out[(size_t)a * d + b] = GS_ACCM(i, jb, c);
        }
    }
    return;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int a = a0 + 2 * ty + 32 * (i >> 1) + (i & 1);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int b = b0 + 2 * tx + 32 * (k >> 1) + (k & 1);
      if (a < d && b < d) out[(size_t)a * d + b] = acc[i][k];
    }
  }
}

// ---- symmetric eigendecomposition: parallel cyclic Jacobi, one CTA ----
// A (m x m, row-major, destroyed: its diagonal ends as the eigenvalues), V (m x m) the
// eigenvectors in columns.  Round-robin ordering gives m/2 disjoint rotations per step.
__global__ void __launch_bounds__(1024) jacobi_eig_kernel(const double* __restrict__ G, int d, int j,
                                                         double* __restrict__ A, double* __restrict__ V,
                                                         double* __restrict__ evals, int* __restrict__ sweeps_out) {
  __shared__ double s_c[kJacobiMax / 2 + 1], s_s[kJacobiMax / 2 + 1];
  __shared__ int s_p[kJacobiMax / 2 + 1], s_q[kJacobiMax / 2 + 1];
  __shared__ double s_red[1024];
  __shared__ double s_off, s_norm;
  const int m = d - 1;
  const int tid = threadIdx.x, nt = blockDim.x;
  // A = G without row/column j; V = I
  for (int e = tid; e < m * m; e += nt) {
    int a = e / m, b = e % m;
    int ga = a < j ? a : a + 1, gb = b < j ? b : b + 1;
    A[e] = G[(size_t)ga * d + gb];
    V[e] = a == b ? 1.0 : 0.0;
  }
  __syncthreads();
  const int M = (m + 1) & ~1;  // players, even
  const int half = M / 2;
  int sweep = 0;
  double prev_off = -1.0;
  for (; sweep < 60; ++sweep) {
    // off-diagonal and total Frobenius norms
    double off = 0.0, tot = 0.0;
    for (int e = tid; e < m * m; e += nt) {
      double v = A[e];
      tot += v * v;
      if (e / m != e % m) off += v * v;
    }
    s_red[tid] = off;
    __syncthreads();
    for (int s = nt >> 1; s > 0; s >>= 1) {
      if (tid < s) s_red[tid] += s_red[tid + s];
      __syncthreads();
    }
    if (tid == 0) s_off = s_red[0];
    __syncthreads();
    s_red[tid] = tot;
    __syncthreads();
    for (int s = nt >> 1; s > 0; s >>= 1) {
      if (tid < s) s_red[tid] += s_red[tid + s];
      __syncthreads();
    }
    if (tid == 0) s_norm = s_red[0];
    __syncthreads();
    // converged at off/||A|| <= 1e-15, or stagnating on the rounding floor (below 1e-12 and no
    // longer halving per sweep)
    if (s_off <= 1e-30 * s_norm || s_norm == 0.0) break;
    if (s_off <= 1e-24 * s_norm && prev_off >= 0.0 && s_off > 0.25 * prev_off) break;
    prev_off = s_off;
    for (int step = 0; step < M - 1; ++step) {
      // pairs of this step
      if (tid < half) {
        int p, q;
        if (tid == 0) {
          p = M - 1;
          q = step;
        } else {
          p = (step + tid) % (M - 1);
          q = (step - tid + (M - 1)) % (M - 1);
        }
        if (p > q) {
          int t = p;
          p = q;
          q = t;
        }
        double c = 1.0, s = 0.0;
        if (q < m) {
          const double apq = A[(size_t)p * m + q];
          const double app = A[(size_t)p * m + p], aqq = A[(size_t)q * m + q];
          if (fabs(apq) > 1e-300 && fabs(apq) > 1e-18 * sqrt(fabs(app * aqq))) {
            const double theta = (aqq - app) / (2.0 * apq);
            const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
            c = 1.0 / sqrt(t * t + 1.0);
            s = t * c;
          }
        } else {
          q = -1;  // dummy player
        }
        s_p[tid] = p;
        s_q[tid] = q;
        s_c[tid] = c;
        s_s[tid] = s;
      }
      __syncthreads();
      // columns: A <- A J, V <- V J
      for (int e = tid; e < half * m; e += nt) {
        const int pr = e / m, i = e % m;
        const int q = s_q[pr];
        if (q < 0 || s_s[pr] == 0.0) continue;
        const int p = s_p[pr];
        const double c = s_c[pr], s = s_s[pr];
        const double aip = A[(size_t)i * m + p], aiq = A[(size_t)i * m + q];
        A[(size_t)i * m + p] = c * aip - s * aiq;
        A[(size_t)i * m + q] = s * aip + c * aiq;
        const double vip = V[(size_t)i * m + p], viq = V[(size_t)i * m + q];
        V[(size_t)i * m + p] = c * vip - s * viq;
        V[(size_t)i * m + q] = s * vip + c * viq;
      }
      __syncthreads();
      // rows: A <- J^T A
      for (int e = tid; e < half * m; e += nt) {
        const int pr = e / m, i = e % m;
        const int q = s_q[pr];
        if (q < 0 || s_s[pr] == 0.0) continue;
        const int p = s_p[pr];
        const double c = s_c[pr], s = s_s[pr];
        const double api = A[(size_t)p * m + i], aqi = A[(size_t)q * m + i];
        A[(size_t)p * m + i] = c * api - s * aqi;
        A[(size_t)q * m + i] = s * api + c * aqi;
      }
      __syncthreads();
    }
  }
  for (int k = tid; k < m; k += nt) evals[k] = fmax(A[(size_t)k * m + k], 0.0);  // S^2 >= 0
  if (tid == 0) *sweeps_out = sweep;
}

// ---- BayesianRidge's evidence iteration in the eigenbasis, one CTA ----
__device__ __forceinline__ double block_sum(double v, double* s_red) {
  const int tid = threadIdx.x;
  __syncthreads();
  s_red[tid] = v;
  __syncthreads();
  for (int s = blockDim.x >> 1; s > 0; s >>= 1) {
    if (tid < s) s_red[tid] += s_red[tid + s];
    __syncthreads();
  }
  const double r = s_red[0];
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(256) bayes_ridge_kernel(const double* __restrict__ G, const double* __restrict__ mu,
                                                         const int* __restrict__ n_obs_p, int d, int j,
                                                         const double* __restrict__ V, int v_si, int v_sk,
                                                         const double* __restrict__ evals,
                                                         double* __restrict__ coef_out, double* __restrict__ intercept_out,
                                                         int* __restrict__ iters_out) {
  __shared__ double s_red[256];
  extern __shared__ double s_dyn[];  // 6 vectors of m doubles
  const int m = d - 1, tid = threadIdx.x, nt = blockDim.x;
  double *s_b = s_dyn, *s_bt = s_dyn + m, *s_ct = s_dyn + 2 * m, *s_coef = s_dyn + 3 * m, *s_old = s_dyn + 4 * m,
         *s_ev = s_dyn + 5 * m;
  const double n = (double)*n_obs_p;
  for (int i = tid; i < m; i += nt) {
    const int gi = i < j ? i : i + 1;
    s_b[i] = G[(size_t)gi * d + j];  // X^T y
    s_ev[i] = fmax(evals[i], 0.0);  // S^2 >= 0
    s_old[i] = 0.0;
  }
  __syncthreads();
  const double yy = G[(size_t)j * d + j];
  for (int k = tid; k < m; k += nt) {  // b~ = V^T b
    double acc = 0.0;
    for (int i = 0; i < m; ++i) acc += V[(size_t)i * v_si + (size_t)k * v_sk] * s_b[i];
    s_bt[k] = acc;
  }
  __syncthreads();
  const double eps = 2.220446049250313e-16;
  double alpha = 1.0 / (yy / n + eps);  // 1/(var(y)+eps), population variance
  double lambda = 1.0;
  int it = 0;
  // One round in the eigenbasis costs O(m): coef~ = b~/(s + lambda/alpha), sse, gamma, and
  // sum(coef^2) = sum(coef~^2) (V is orthonormal).  Only the stopping rule sum|coef_old - coef| < tol
  // lives in the original basis; since ||x||_2 <= ||x||_1 <= sqrt(m)||x||_2 and ||V x||_2 = ||x||_2,
  // the O(m^2) product V (coef~ - coef~_old) is needed only when ||coef~ - coef~_old||_2 lies in
  // [tol/sqrt(m), tol) — the last round or two.
  auto round_eigen = [&](double al, double la, double* sse) {
    const double ratio = la / al;
    for (int k = tid; k < m; k += nt) s_ct[k] = s_bt[k] / (s_ev[k] + ratio);
    double t1 = 0.0, t2 = 0.0;
    for (int k = tid; k < m; k += nt) {
      t1 += s_ct[k] * s_bt[k];
      t2 += s_ct[k] * s_ct[k] * s_ev[k];
    }
    const double a = block_sum(t1, s_red), b = block_sum(t2, s_red);
    *sse = fmax(yy - 2.0 * a + b, 0.0);
  };
  const double sqrt_m = sqrt((double)m);
  for (it = 0; it < kBrMaxIter; ++it) {
    double sse;
    round_eigen(alpha, lambda, &sse);
    double g = 0.0, sq = 0.0, d2 = 0.0;
    for (int k = tid; k < m; k += nt) {
      g += (alpha * s_ev[k]) / (lambda + alpha * s_ev[k]);
      sq += s_ct[k] * s_ct[k];
      const double df = s_old[k] - s_ct[k];  // s_old holds the previous round's coef~
      d2 += df * df;
    }
    const double gamma = block_sum(g, s_red);
    const double sumsq = block_sum(sq, s_red);
    const double dnorm = sqrt(block_sum(d2, s_red));
    lambda = (gamma + 2.0 * kBrPrior) / (sumsq + 2.0 * kBrPrior);
    alpha = (n - gamma + 2.0 * kBrPrior) / (sse + 2.0 * kBrPrior);
    if (it != 0) {
      bool converged = sqrt_m * dnorm < kBrTol;
      if (!converged && dnorm < kBrTol) {  // undecided: evaluate the rule itself
        for (int k = tid; k < m; k += nt) s_coef[k] = s_old[k] - s_ct[k];  // difference in the eigenbasis
        __syncthreads();
        double dl = 0.0;
        for (int i = tid; i < m; i += nt) {
          double acc = 0.0;
          for (int k = 0; k < m; ++k) acc += V[(size_t)i * v_si + (size_t)k * v_sk] * s_coef[k];
          dl += fabs(acc);
        }
        converged = block_sum(dl, s_red) < kBrTol;
      }
      if (converged) break;
    }
    __syncthreads();
    for (int k = tid; k < m; k += nt) s_old[k] = s_ct[k];
    __syncthreads();
  }
  double sse;
  round_eigen(alpha, lambda, &sse);
  __syncthreads();
  for (int i = tid; i < m; i += nt) {  // coef = V coef~ with the final alpha, lambda
    double acc = 0.0;
    for (int k = 0; k < m; ++k) acc += V[(size_t)i * v_si + (size_t)k * v_sk] * s_ct[k];
    s_coef[i] = acc;
  }
  __syncthreads();
  double dot = 0.0;
  for (int i = tid; i < m; i += nt) {
    const int gi = i < j ? i : i + 1;
    coef_out[i] = s_coef[i];
    dot += mu[gi] * s_coef[i];
  }
  const double xo = block_sum(dot, s_red);
  if (tid == 0) {
    *intercept_out = mu[j] - xo;
    *iters_out = it < kBrMaxIter ? it + 1 : kBrMaxIter;
  }
}

// rows where feature j is missing: Xt[r][j] = Xt[r][others] . coef + intercept; one warp per row
__global__ void __launch_bounds__(256) predict_kernel(double* __restrict__ Xt, const uint8_t* __restrict__ mask, int n,
                                                     int d, int j, const double* __restrict__ coef,
                                                     const double* __restrict__ intercept) {
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (r >= n || !mask[(size_t)r * d + j]) return;
  double acc = 0.0;
  for (int c = lane; c < d; c += 32) {
    if (c == j) continue;
    acc += Xt[(size_t)r * d + c] * coef[c < j ? c : c - 1];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
  if (lane == 0) Xt[(size_t)r * d + j] = acc + *intercept;
}

// np.linalg.norm(Xt - Xt_prev, ord=inf, axis=None) of a 2-D array = the largest row sum of
// absolute differences (_iterative.py:853).  One warp per row; out[block] = max over its rows.
__global__ void __launch_bounds__(256) max_row_diff_kernel(const double* __restrict__ a, const double* __restrict__ b,
                                                          int n, int d, double* __restrict__ out) {
  __shared__ double s_red[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double mx = 0.0;
  for (int r = blockIdx.x * 8 + warp; r < n; r += gridDim.x * 8) {
    double sum = 0.0;
    for (int c = lane; c < d; c += 32) sum += fabs(a[(size_t)r * d + c] - b[(size_t)r * d + c]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xFFFFFFFFu, sum, o);
    mx = fmax(mx, sum);
  }
  if (lane == 0) s_red[warp] = mx;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) mx = fmax(mx, s_red[w]);
    out[blockIdx.x] = mx;
  }
}

// transform: one warp per row, the row in shared memory, every stored step replayed in order
// (impute/_iterative.py:907-915); observed entries never change.
__global__ void __launch_bounds__(128) transform_kernel(const double* __restrict__ X, int m_rows, int d,
                                                       const double* __restrict__ means, const int* __restrict__ feat,
                                                       const double* __restrict__ coef,
                                                       const double* __restrict__ intercept, int n_steps,
                                                       double* __restrict__ out) {
  extern __shared__ double s_rows[];  // [4 warps][d] values, then [4 warps][d] mask bytes packed as doubles' tail
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * 4 + warp;
  if (r >= m_rows) return;
  double* row = s_rows + (size_t)warp * d;
  uint8_t* miss = reinterpret_cast<uint8_t*>(s_rows + (size_t)4 * d) + (size_t)warp * d;
  for (int c = lane; c < d; c += 32) {
    const double v = X[(size_t)r * d + c];
    const bool ms = !(v == v);
    miss[c] = ms ? 1 : 0;
    row[c] = ms ? means[c] : v;
  }
  __syncwarp();
  for (int s = 0; s < n_steps; ++s) {
    const int j = feat[s];
    if (!miss[j]) continue;
    const double* cf = coef + (size_t)s * (d - 1);
    double acc = 0.0;
    for (int c = lane; c < d; c += 32) {
      if (c == j) continue;
      acc += row[c] * cf[c < j ? c : c - 1];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
    __syncwarp();
    if (lane == 0) row[j] = acc + intercept[s];
    __syncwarp();
  }
  for (int c = lane; c < d; c += 32) out[(size_t)r * d + c] = row[c];
}

}  // namespace gs

// ------------------------------------------------------------------------------------------

struct gs_model {
  int d = 0, n_rounds = 0, n_steps = 0;
  std::vector<double> means, coef, intercept;
  std::vector<int> feat;
};

namespace {

struct DevMem {
  std::vector<void*> ptrs;
  ~DevMem() {
    for (void* p : ptrs) cudaFree(p);
  }
  template <typename T>
  cudaError_t alloc(T** out, size_t count) {
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, sizeof(T) * (count ? count : 1));
    if (e == cudaSuccess) ptrs.push_back(p);
    *out = static_cast<T*>(p);
    return e;
  }
};

// Device time of a fit's three parts, sampled on at most 64 estimator fits (events on the stream; read after the
// fit's last synchronisation): moments (column sums over the observed rows), Gram (kernels, all-reduce,
// recentring), estimator (decomposition / reduction and the evidence iteration).
struct FitTimer {
  std::vector<cudaEvent_t> ev;  // 4 per sample, created up front
  size_t used = 0;
  int stride = 1;
  explicit FitTimer(int max_steps) : stride(max_steps > 64 ? (max_steps + 63) / 64 : 1) {
    const size_t want = 4 * (size_t)((max_steps + stride - 1) / stride);
    for (size_t k = 0; k < want; ++k) {
      cudaEvent_t e = nullptr;
      if (cudaEventCreate(&e) != cudaSuccess) {  // no timing then; the fit itself is unaffected
        cudaGetLastError();
        clear();
        return;
      }
      ev.push_back(e);
    }
  }
  ~FitTimer() { clear(); }
  void clear() {
    for (cudaEvent_t e : ev) cudaEventDestroy(e);
    ev.clear();
  }
  bool sampled(int step) const { return !ev.empty() && step % stride == 0 && used + 4 <= ev.size(); }
  void mark(cudaStream_t s) {
    if (used < ev.size()) cudaEventRecord(ev[used++], s);
  }
  void summarise(double out[4]) const {
    out[0] = out[1] = out[2] = out[3] = 0.0;
    const size_t n = used / 4;
    for (size_t k = 0; k < n; ++k)
      for (int part = 0; part < 3; ++part) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ev[4 * k + part], ev[4 * k + part + 1]) == cudaSuccess) out[1 + part] += ms;
      }
    out[0] = (double)n;
    if (n)
      for (int part = 1; part < 4; ++part) out[part] /= (double)n;
  }
};

#define CK(call)                                              \
  do {                                                        \
    cudaError_t _e = (call);                                  \
    if (_e != cudaSuccess) return c->cuda_fail(_e, #call);    \
  } while (0)

}  // namespace

extern "C" {

int gs_complete_fit(gs_ctx* c, const double* X, int32_t n, int32_t d, gs_model** out) {
  return gs_complete_fit_transform(c, X, n, d, nullptr, nullptr, nullptr, out);
}

int gs_complete_fit_sharded(gs_ctx* c, const double* X, int32_t n, int32_t d, gs_allreduce_fn reduce, void* user,
                            gs_model** out) {
  return gs_complete_fit_transform(c, X, n, d, reduce, user, nullptr, out);
}

int gs_complete_fit_transform(gs_ctx* c, const double* X, int32_t n, int32_t d, gs_allreduce_fn reduce, void* user,
                              double* out_completed, gs_model** out) {
  using namespace gs;
  if (out) *out = nullptr;
  if (!c) return GS_ERR_INVALID;
  std::lock_guard<std::mutex> lk(c->mu);
  if (!X || !out || n < 1 || d < 1) return c->fail(GS_ERR_INVALID, "gs_complete_fit: null pointer or empty matrix");
  // the path's exchange steps: sums / maxima of small float64 buffers over the row shards
  auto exchange = [&](double* buf, int64_t count, int op, cudaStream_t st) -> int {
    if (!reduce) return GS_OK;
    int rc = reduce(buf, count, op, (void*)st, user);
    return rc == 0 ? GS_OK : c->fail(GS_ERR_CUDA, "gs_complete_fit_sharded: the all-reduce callback failed (%d)", rc);
  };
  if (d - 1 >= kMaxFeatures)
    return c->fail(GS_ERR_UNSUPPORTED, "gs_complete_fit: %d columns; this build covers at most %d", d, kMaxFeatures);
  CK(cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  const size_t total = (size_t)n * d;
  DevMem mem;
  double *dX, *dXt, *dPrev, *dMeans, *dStats, *dSums, *dSumPart, *dMu, *dG, *dPartial, *dA, *dV, *dEv, *dCoef, *dIcpt, *dDiff;
  uint8_t* dMask;
  int *dNobs, *dIters, *dSweeps;
  // Gram kernel: 128x128 tiles for wide tables, 32x32 otherwise; rows split so that the grid fills the GPU
  const bool big_gram = d >= 128;
  const int gtile = big_gram ? kBT : kGramTile;
  const int tiles = (d + gtile - 1) / gtile;
  int parts;
  if (big_gram) {
    const int upper = tiles * (tiles + 1) / 2;
    parts = (2 * 148 + upper - 1) / upper;
    if (parts > (n + 255) / 256) parts = (n + 255) / 256;
  } else {
    parts = (n + 255) / 256;
    if (parts > 128) parts = 128;
  }
  if (parts < 1) parts = 1;
  int rows_per_part = (n + parts - 1) / parts;
  rows_per_part = (rows_per_part + 31) / 32 * 32;
  parts = (n + rows_per_part - 1) / rows_per_part;
  const int max_steps = kRounds * d;
  CK(mem.alloc(&dX, total));
  CK(mem.alloc(&dXt, total));
  CK(mem.alloc(&dPrev, total));
  CK(mem.alloc(&dMask, total));
  CK(mem.alloc(&dMeans, d));
  CK(mem.alloc(&dStats, 3 * (size_t)d + 1));
  CK(mem.alloc(&dSums, (size_t)d + 1));
  const int sum_blocks = (n + kSumRows - 1) / kSumRows;
  CK(mem.alloc(&dSumPart, (size_t)sum_blocks * ((size_t)d + 1)));
  CK(mem.alloc(&dMu, d));
  CK(mem.alloc(&dNobs, 1));
  CK(mem.alloc(&dG, (size_t)d * d));
  CK(mem.alloc(&dPartial, (size_t)parts * d * d));
  CK(mem.alloc(&dA, (size_t)d * d));
  CK(mem.alloc(&dV, (size_t)d * d));
  CK(mem.alloc(&dEv, d));
  CK(mem.alloc(&dCoef, (size_t)max_steps * (d > 1 ? d - 1 : 1)));
  CK(mem.alloc(&dIcpt, max_steps));
  CK(mem.alloc(&dIters, max_steps));
  CK(mem.alloc(&dSweeps, max_steps));
  CK(mem.alloc(&dDiff, 256));
  CK(cudaMemcpyAsync(dX, X, sizeof(double) * total, cudaMemcpyHostToDevice, s));
  // Wide tables: the Gram over the rows where feature j is observed is taken as (Gram over all
  // rows, kept in dH and refreshed one row per step) minus (Gram over the rows where j is missing,
  // listed per feature in dDead), whenever fewer rows are missing than observed.
  double *dH = nullptr, *dHPart = nullptr;
  int* dDead = nullptr;
  std::vector<size_t> dead_off;
  const int h_blocks = (n + kHRows - 1) / kHRows;
  if (big_gram) {
    dead_off.assign((size_t)d + 1, 0);
    for (size_t r = 0; r < (size_t)n; ++r)
      for (int k = 0; k < d; ++k) dead_off[k + 1] += std::isnan(X[r * d + k]) ? 1 : 0;
    for (int k = 0; k < d; ++k) dead_off[k + 1] += dead_off[k];
    std::vector<int> dead(dead_off[d] ? dead_off[d] : 1);
    std::vector<size_t> cur(dead_off.begin(), dead_off.end() - 1);
    for (size_t r = 0; r < (size_t)n; ++r)
      for (int k = 0; k < d; ++k)
        if (std::isnan(X[r * d + k])) dead[cur[k]++] = (int)r;
    CK(mem.alloc(&dDead, dead.size()));
    CK(cudaMemcpyAsync(dDead, dead.data(), sizeof(int) * dead.size(), cudaMemcpyHostToDevice, s));
    CK(cudaStreamSynchronize(s));  // `dead` goes out of scope
    CK(mem.alloc(&dH, (size_t)d * d));
    CK(mem.alloc(&dHPart, (size_t)h_blocks * d));
  }
  // eigen-solver: the one-CTA Jacobi up to kJacobiMax, the multi-CTA block Jacobi (eig_jacobi.cu) beyond
  const int mm = d - 1;
  // ... unless the tridiagonal path (tridiag.cu) takes the wide tables: no eigendecomposition at all
  // (GS_EIG_TRI=0 keeps the block Jacobi)
  int n_sm = 148;
  {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
      n_sm = 148;
  }
  const char* tri_env = getenv("GS_EIG_TRI");
  const bool use_tri = mm > kJacobiMax && !(tri_env && atoi(tri_env) == 0) && mm <= tri_max_m(n_sm) && mm <= 1024;
  const bool use_block = mm > kJacobiMax && !use_tri;
  double* dTriWork = nullptr;
  if (use_tri) {
    CK(mem.alloc(&dTriWork, tri_work_doubles(mm)));
    CK(cudaMemsetAsync(dSweeps, 0, sizeof(int) * max_steps, s));  // no sweeps on this path
  }
  double* dEigWork = nullptr;
  const int eig_mp = use_block ? eig_padded_size(mm) : 0;
  std::vector<int> block_sweeps;
  // warm start: keep every feature's eigenvectors for the next round (the Gram matrix of a feature
  // changes little from one round to the next), if d * mp^2 doubles fit the budget (GS_EIG_WARM_GB, default 24)
  double* dVstore = nullptr;
  std::vector<char> have_v((size_t)d, 0);
  if (use_block) {
    size_t warm_gb = 24;
    if (const char* v = getenv("GS_EIG_WARM_GB")) warm_gb = (size_t)strtoull(v, nullptr, 10);
    const size_t per = (size_t)eig_mp * eig_mp;
    if (per * (size_t)d * sizeof(double) <= (warm_gb << 30) && mem.alloc(&dVstore, per * (size_t)d) != cudaSuccess) {
      cudaGetLastError();
      dVstore = nullptr;  // no room: every decomposition starts cold
    }
  }
  if (use_block) {
    CK(mem.alloc(&dEigWork, eig_work_doubles(mm)));
    CK(cudaMemsetAsync(dSweeps, 0, sizeof(int) * max_steps, s));
  }
  const size_t bayes_smem = sizeof(double) * 6 * (size_t)(mm > 0 ? mm : 1);
  CK(cudaFuncSetAttribute(bayes_ridge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bayes_smem));

  std::unique_ptr<gs_model> owner(new (std::nothrow) gs_model());  // freed on every early return
  gs_model* m = owner.get();
  if (!m) return c->fail(GS_ERR_NOMEM, "gs_complete_fit: out of host memory");
  m->d = d;
  m->means.resize(d);
  std::vector<long long> missing(d);
  std::vector<double> stats(3 * (size_t)d + 1);

  gs::note_launch();
  column_stats_kernel<<<d, 256, 0, s>>>(dX, n, d, dStats);
  int xrc;
  if ((xrc = exchange(dStats, 2 * (int64_t)d + 1, GS_REDUCE_SUM, s)) || (xrc = exchange(dStats + 2 * d + 1, d, GS_REDUCE_MAX, s))) {
    return xrc;
  }
  gs::note_launch();
  finish_means_kernel<<<(d + 255) / 256, 256, 0, s>>>(dStats, d, dMeans);
  gs::note_launch();
  fill_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(dX, dMeans, total, d, dXt, dMask);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpyAsync(m->means.data(), dMeans, sizeof(double) * d, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(stats.data(), dStats, sizeof(double) * stats.size(), cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) {
    return c->cuda_fail(e, "completion: initial imputation");
  }
  const long long n_all = (long long)stats[2 * d];  // rows over all shards
  for (int k = 0; k < d; ++k) missing[k] = n_all - (long long)stats[d + k];
  const double* max_abs = stats.data() + 2 * d + 1;
  bool all_missing = true;
  for (int k = 0; k < d; ++k) {
    if (missing[k] == n_all) {
      return c->fail(GS_ERR_UNSUPPORTED, "gs_complete_fit: column %d has no observed value (scikit-learn drops such "
                     "columns; not served)", k);
    }
    if (missing[k] != n_all) all_missing = false;
  }
  if (d == 1 || all_missing) {  // _iterative.py:783-789: nothing to iterate
    if (out_completed) {
      CK(cudaMemcpyAsync(out_completed, dXt, sizeof(double) * total, cudaMemcpyDeviceToHost, s));
      CK(cudaStreamSynchronize(s));
    }
    *out = owner.release();
    return GS_OK;
  }
  // ascending missing fraction, stable (np.argsort(kind="mergesort")), all columns (skip_complete=False)
  std::vector<int> order(d);
  for (int k = 0; k < d; ++k) order[k] = k;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return missing[a] < missing[b]; });
  double mx = 0.0;
  for (int k = 0; k < d; ++k) mx = std::fmax(mx, max_abs[k]);
  const double norm_tol = kRoundTol * mx;

  // wide tables: the Gram tiles on the FP64 tensor cores (mma.sync m8n8k4) unless GS_GRAM_MMA=0
  bool gram_mma = true;
  if (const char* v = getenv("GS_GRAM_MMA")) gram_mma = atoi(v) != 0;
  if (big_gram) {  // Gram over all rows, centred at the initial means
    gs::note_launch();
    if (gram_mma)
      gram128_partial_kernel<false, true><<<dim3(tiles, tiles, parts), 256, 0, s>>>(dXt, dMask, nullptr, n, d, -1, dMeans,
                                                                                 rows_per_part, dPartial);
    else
      gram128_partial_kernel<false, false><<<dim3(tiles, tiles, parts), 256, 0, s>>>(dXt, dMask, nullptr, n, d, -1, dMeans,
                                                                                  rows_per_part, dPartial);
    gs::note_launch();
    gram_reduce_kernel<<<(d * d + 255) / 256, 256, 0, s>>>(dPartial, d, parts, gtile, nullptr, dH);
  }
  int step = 0;
  FitTimer timer(max_steps);
  std::vector<double> diff(256);
  CK(cudaMemcpyAsync(dPrev, dXt, sizeof(double) * total, cudaMemcpyDeviceToDevice, s));
  int max_rounds = kRounds;
  if (const char* ov = getenv("GS_COMPLETE_MAX_ROUNDS")) max_rounds = std::max(1, std::min(kRounds, atoi(ov)));  // probing only
  for (int rnd = 1; rnd <= max_rounds; ++rnd) {
    m->n_rounds = rnd;
    for (int oi = 0; oi < d; ++oi, ++step) {
      const int j = order[oi];
      const bool timed = timer.sampled(step);
      if (timed) timer.mark(s);
      gs::note_launch();
      obs_sums_partial_kernel<<<sum_blocks, 256, 0, s>>>(dXt, dMask, n, d, j, dSumPart);
      gs::note_launch();
      obs_sums_reduce_kernel<<<(d + 1 + 255) / 256, 256, 0, s>>>(dSumPart, sum_blocks, d, dSums);
      if ((xrc = exchange(dSums, (int64_t)d + 1, GS_REDUCE_SUM, s))) {
        return xrc;
      }
      gs::note_launch();
      finish_mu_kernel<<<(d + 255) / 256, 256, 0, s>>>(dSums, d, dMu, dNobs);
      if (timed) timer.mark(s);
      if (big_gram) {
        const int n_dead = (int)(dead_off[j + 1] - dead_off[j]);
        if (2 * (size_t)n_dead < (size_t)n) {
          int rpp = (n_dead + parts - 1) / parts;
          rpp = (rpp + kBK - 1) / kBK * kBK;
          gs::note_launch();
          if (gram_mma)
            gram128_partial_kernel<true, true><<<dim3(tiles, tiles, parts), 256, 0, s>>>(
                dXt, dMask, dDead + dead_off[j], n_dead, d, j, dMeans, rpp > 0 ? rpp : kBK, dPartial);
          else
            gram128_partial_kernel<true, false><<<dim3(tiles, tiles, parts), 256, 0, s>>>(
                dXt, dMask, dDead + dead_off[j], n_dead, d, j, dMeans, rpp > 0 ? rpp : kBK, dPartial);
          gs::note_launch();
          gram_reduce_kernel<<<(d * d + 255) / 256, 256, 0, s>>>(dPartial, d, parts, gtile, dH, dG);
        } else {
          gs::note_launch();
          if (gram_mma)
            gram128_partial_kernel<false, true><<<dim3(tiles, tiles, parts), 256, 0, s>>>(dXt, dMask, nullptr, n, d, j,
                                                                                       dMeans, rows_per_part, dPartial);
          else
            gram128_partial_kernel<false, false><<<dim3(tiles, tiles, parts), 256, 0, s>>>(dXt, dMask, nullptr, n, d, j,
                                                                                        dMeans, rows_per_part, dPartial);
          gs::note_launch();
          gram_reduce_kernel<<<(d * d + 255) / 256, 256, 0, s>>>(dPartial, d, parts, gtile, nullptr, dG);
        }
      } else {
        gs::note_launch();
        gram_partial_kernel<<<dim3(tiles, tiles, parts), 256, 0, s>>>(dXt, dMask, n, d, j, dMu, rows_per_part, dPartial);
        gs::note_launch();
        gram_reduce_kernel<<<(d * d + 255) / 256, 256, 0, s>>>(dPartial, d, parts, gtile, nullptr, dG);
      }
      if ((xrc = exchange(dG, (int64_t)d * d, GS_REDUCE_SUM, s))) {
        return xrc;
      }
      if (big_gram) {
        gs::note_launch();
        recentre_kernel<<<(d * d + 255) / 256, 256, 0, s>>>(dG, d, dMu, dMeans, dNobs);
      }
      if (timed) timer.mark(s);
      if (use_tri) {
        const cudaError_t ee = tri_bayes_ridge(dG, dMu, dNobs, d, j, dTriWork, n_sm, dCoef + (size_t)step * (d - 1),
                                               dIcpt + step, dIters + step, s);
        if (ee != cudaSuccess) return c->cuda_fail(ee, "completion: tridiagonal BayesianRidge");
      } else if (use_block) {
        int sw = 0;
        double* vj = dVstore ? dVstore + (size_t)j * eig_mp * eig_mp : nullptr;
        const cudaError_t ee =
            eig_block_jacobi(dG, d, j, dEigWork, dEv, &sw, s, (vj && have_v[j]) ? vj : nullptr, vj);
        if (vj) have_v[j] = 1;
        if (ee != cudaSuccess) return c->cuda_fail(ee, "completion: block Jacobi eigensolver");
        block_sweeps.push_back(sw);
        // eigenvectors in the columns of V = work + mp*mp, row stride mp
        gs::note_launch();
        bayes_ridge_kernel<<<1, 256, bayes_smem, s>>>(dG, dMu, dNobs, d, j, dEigWork + (size_t)eig_mp * eig_mp, eig_mp, 1,
                                                      dEv, dCoef + (size_t)step * (d - 1), dIcpt + step, dIters + step);
      } else {
        int jt = 32;
        while (jt < 1024 && jt < mm * mm / 4) jt <<= 1;
        gs::note_launch();
        jacobi_eig_kernel<<<1, jt, 0, s>>>(dG, d, j, dA, dV, dEv, dSweeps + step);
        gs::note_launch();
        bayes_ridge_kernel<<<1, 256, bayes_smem, s>>>(dG, dMu, dNobs, d, j, dV, mm, 1, dEv,
                                                      dCoef + (size_t)step * (d - 1), dIcpt + step, dIters + step);
      }
      if (timed) timer.mark(s);
      if (missing[j] > 0) {
        gs::note_launch();
        predict_kernel<<<(n + 7) / 8, 256, 0, s>>>(dXt, dMask, n, d, j, dCoef + (size_t)step * (d - 1), dIcpt + step);
      }
      if (big_gram && dead_off[j + 1] > dead_off[j]) {  // column j changed on this rank: row/column j of dH
        gs::note_launch();
        h_row_partial_kernel<<<h_blocks, 256, 0, s>>>(dXt, n, d, j, dMeans, dHPart);
        gs::note_launch();
        h_row_reduce_kernel<<<(d + 255) / 256, 256, 0, s>>>(dHPart, h_blocks, d, j, dH);
      }
      m->feat.push_back(j);
    }
    gs::note_launch();
    max_row_diff_kernel<<<256, 256, 0, s>>>(dXt, dPrev, n, d, dDiff);
    gs::note_launch();
    max_small_kernel<<<1, 256, 0, s>>>(dDiff, 256);
    if ((xrc = exchange(dDiff, 1, GS_REDUCE_MAX, s))) {
      return xrc;
    }
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(diff.data(), dDiff, sizeof(double), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) {
      return c->cuda_fail(e, "completion: round");
    }
    const double inf_norm = diff[0];
    if (inf_norm < norm_tol) break;  // _iterative.py:853-862 (matrix infinity norm)
    CK(cudaMemcpyAsync(dPrev, dXt, sizeof(double) * total, cudaMemcpyDeviceToDevice, s));
  }
  m->n_steps = step;
  m->coef.resize((size_t)step * (d - 1));
  m->intercept.resize(step);
  std::vector<int> sweeps(step);
  e = cudaMemcpyAsync(m->coef.data(), dCoef, sizeof(double) * m->coef.size(), cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(m->intercept.data(), dIcpt, sizeof(double) * step, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(sweeps.data(), dSweeps, sizeof(int) * step, cudaMemcpyDeviceToHost, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  if (e != cudaSuccess) {
    return c->cuda_fail(e, "completion: model download");
  }
  timer.summarise(c->fit_timing);
  if (use_block)
    for (int k = 0; k < step && k < (int)block_sweeps.size(); ++k) sweeps[k] = block_sweeps[k] >= 40 ? 60 : block_sweeps[k];
  for (int k = 0; k < step; ++k)
    if (sweeps[k] >= 60) {
      return c->fail(GS_ERR_NOCONV, "gs_complete_fit: the eigen-solver did not converge at step %d", k);
    }
  // the completed training rows: what transform(X) would return (observed entries are never written)
  if (out_completed) {
    CK(cudaMemcpyAsync(out_completed, dXt, sizeof(double) * total, cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
  }
  *out = owner.release();
  return GS_OK;
}

int gs_complete_last_timing(gs_ctx* c, double* out4) {
  if (!c) return GS_ERR_INVALID;
  std::lock_guard<std::mutex> lk(c->mu);
  if (!out4) return c->fail(GS_ERR_INVALID, "gs_complete_last_timing: null output");
  for (int k = 0; k < 4; ++k) out4[k] = c->fit_timing[k];
  return GS_OK;
}

uint64_t gs_eig_stats(uint64_t* out_sweeps) {
  unsigned long long f = 0, s = 0;
  gs::eig_stats(&f, &s);
  if (out_sweeps) *out_sweeps = s;
  return f;
}

int gs_complete_transform(gs_ctx* c, const gs_model* m, const double* X, int32_t rows, double* out) {
  using namespace gs;
  if (!c) return GS_ERR_INVALID;
  std::lock_guard<std::mutex> lk(c->mu);
  if (!m || rows < 0 || (rows && (!X || !out))) return c->fail(GS_ERR_INVALID, "gs_complete_transform: null pointer");
  if (rows == 0) return GS_OK;
  CK(cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  const int d = m->d;
  const size_t total = (size_t)rows * d;
  DevMem mem;
  double *dX, *dOut, *dMeans, *dCoef, *dIcpt;
  int* dFeat;
  CK(mem.alloc(&dX, total));
  CK(mem.alloc(&dOut, total));
  CK(mem.alloc(&dMeans, d));
  CK(mem.alloc(&dCoef, m->coef.size()));
  CK(mem.alloc(&dIcpt, m->intercept.size()));
  CK(mem.alloc(&dFeat, m->feat.size()));
  CK(cudaMemcpyAsync(dX, X, sizeof(double) * total, cudaMemcpyHostToDevice, s));
  CK(cudaMemcpyAsync(dMeans, m->means.data(), sizeof(double) * d, cudaMemcpyHostToDevice, s));
  if (m->n_steps) {
    CK(cudaMemcpyAsync(dCoef, m->coef.data(), sizeof(double) * m->coef.size(), cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(dIcpt, m->intercept.data(), sizeof(double) * m->n_steps, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(dFeat, m->feat.data(), sizeof(int) * m->n_steps, cudaMemcpyHostToDevice, s));
  }
  const size_t smem = (size_t)4 * d * sizeof(double) + (size_t)4 * d + 16;
  CK(cudaFuncSetAttribute(transform_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  gs::note_launch();
  transform_kernel<<<(rows + 3) / 4, 128, smem, s>>>(dX, rows, d, dMeans, dFeat, dCoef, dIcpt, m->n_steps, dOut);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(out, dOut, sizeof(double) * total, cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  return GS_OK;
}

int gs_model_info(const gs_model* m, int32_t* n_features, int32_t* n_rounds, int32_t* n_steps) {
  if (!m) return GS_ERR_INVALID;
  if (n_features) *n_features = m->d;
  if (n_rounds) *n_rounds = m->n_rounds;
  if (n_steps) *n_steps = m->n_steps;
  return GS_OK;
}

int gs_model_step(const gs_model* m, int32_t s, int32_t* feat, double* coef, double* intercept) {
  if (!m || s < 0 || s >= m->n_steps) return GS_ERR_INVALID;
  if (feat) *feat = m->feat[s];
  if (coef)
    for (int k = 0; k < m->d - 1; ++k) coef[k] = m->coef[(size_t)s * (m->d - 1) + k];
  if (intercept) *intercept = m->intercept[s];
  return GS_OK;
}

void gs_model_destroy(gs_model* m) { delete m; }

}  // extern "C"
