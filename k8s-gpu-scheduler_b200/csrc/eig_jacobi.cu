// eig_jacobi.cu — symmetric eigendecomposition for the recommender's completion (sm_100a).
//
// BayesianRidge.fit takes an SVD of the centred design matrix per feature
// (pkg/recommender/recommender/recommender.py:15-17 -> sklearn linear_model/_bayes.py:262-306:
// eigen_vals_ = S^2, Vh); here that is the eigendecomposition of the centred Gram matrix G without
// row/column j.  Up to 24 columns complete.cu's one-CTA Jacobi serves; beyond, this file: a
// two-sided BLOCK Jacobi across CTAs (round 1 called cuSOLVER's syevd there).
//
//   The m x m matrix is padded to mp = nb * kB (nb even) and split into nb block columns of kB = 16.
//   A sweep is nb - 1 steps of a round-robin tournament; in a step the nb / 2 disjoint block pairs
//   (I, J) are processed independently:
//     bj_pivot_kernel  one CTA per pair: the 32 x 32 pivot [A_II A_IJ; A_JI A_JJ] is diagonalised in
//                      shared memory by cyclic Jacobi (parallel ordering, 16 rotations per step) and
//                      the accumulated rotation Q (32 x 32, orthogonal) written out;
//     bj_update_kernel every 32 x 32 block (pair's rows x pair's columns) of A becomes Q_i^T A_blk Q_j and
//                      every block of V becomes V_blk Q_j — one pass, each CTA owns its block.
//   Sweeps repeat until off(A)^2 <= 1e-30 ||A||^2 (or the off-norm stagnates on the rounding floor),
//   the same stopping rule as the one-CTA kernel.  Padding rows/columns are zero and never rotate.
#include <atomic>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

#include "eig_jacobi.h"
#include "tables.h"

namespace gs {

static std::atomic<unsigned long long> g_eig_fits{0}, g_eig_sweeps{0};
void eig_stats(unsigned long long* fits, unsigned long long* sweeps) {
  *fits = g_eig_fits.load();
  *sweeps = g_eig_sweeps.load();
}

constexpr int kB = 16;       // block size
constexpr int kP = 2 * kB;   // pivot size
constexpr int kInnerSweeps = 1;  // default; GS_EIG_INNER overrides (measurement)

__host__ __device__ inline void bj_pair(int nb, int step, int p, int* I, int* J) {
  int a, b;
  if (p == 0) {
    a = nb - 1;
    b = step;
  } else {
    a = (step + p) % (nb - 1);
    b = (step - p + (nb - 1)) % (nb - 1);
  }
  *I = a < b ? a : b;
  *J = a < b ? b : a;
}

// A = G without row/column j, padded with zeros to mp x mp; V = identity
__global__ void bj_init_kernel(const double* __restrict__ G, int d, int j, int mp, double* __restrict__ A,
                               double* __restrict__ V) {
  const int m = d - 1;
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (size_t)mp * mp) return;
  const int a = (int)(e / mp), b = (int)(e % mp);
  double v = 0.0;
  if (a < m && b < m) v = G[(size_t)(a < j ? a : a + 1) * d + (b < j ? b : b + 1)];
  A[e] = v;
  if (V) V[e] = a == b ? 1.0 : 0.0;
}

// norms[0] += sum of squares off the diagonal, norms[1] += sum of all squares
__global__ void __launch_bounds__(256) bj_norm_kernel(const double* __restrict__ A, int mp, double* norms) {
  __shared__ double s_off[256], s_tot[256];
  double off = 0.0, tot = 0.0;
  const size_t n = (size_t)mp * mp;
  for (size_t e = (size_t)blockIdx.x * 256 + threadIdx.x; e < n; e += (size_t)gridDim.x * 256) {
    const double v = A[e];
    tot += v * v;
    if (e / mp != e % mp) off += v * v;
  }
  s_off[threadIdx.x] = off;
  s_tot[threadIdx.x] = tot;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      s_off[threadIdx.x] += s_off[threadIdx.x + s];
      s_tot[threadIdx.x] += s_tot[threadIdx.x + s];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    atomicAdd(&norms[0], s_off[0]);
    atomicAdd(&norms[1], s_tot[0]);
  }
}

// One CTA per block pair: diagonalise the 32 x 32 pivot, write the accumulated rotation Q.
__global__ void __launch_bounds__(256) bj_pivot_kernel(const double* __restrict__ A, int mp, int nb, int step,
                                                      double* __restrict__ Qout, int inner_sweeps) {
  __shared__ double S[kP][kP + 1], Q[kP][kP + 1];
  __shared__ double s_c[kB], s_s[kB];
  __shared__ int s_p[kB], s_q[kB];
  const int tid = threadIdx.x;
  int I, J;
  bj_pair(nb, step, blockIdx.x, &I, &J);
  for (int e = tid; e < kP * kP; e += 256) {
    const int r = e / kP, c = e % kP;
    const int gr = (r < kB ? I * kB + r : J * kB + r - kB), gc = (c < kB ? I * kB + c : J * kB + c - kB);
    S[r][c] = A[(size_t)gr * mp + gc];
    Q[r][c] = r == c ? 1.0 : 0.0;
  }
  __syncthreads();
  // a few cyclic sweeps: the pivot need not be diagonalised exactly for the block iteration to
  // converge, and every barrier here is latency on the critical path of the step.  A sweep in which
  // no rotation passes the threshold ends the loop.
  for (int sweep = 0; sweep < inner_sweeps; ++sweep) {
    bool rotated = false;
    for (int st = 0; st < kP - 1; ++st) {
      if (tid < kB) {
        int p, q;
        if (tid == 0) {
          p = kP - 1;
          q = st;
        } else {
          p = (st + tid) % (kP - 1);
          q = (st - tid + (kP - 1)) % (kP - 1);
        }
        if (p > q) {
          const int t = p;
          p = q;
          q = t;
        }
        double c = 1.0, s = 0.0;
        const double apq = S[p][q], app = S[p][p], aqq = S[q][q];
        if (fabs(apq) > 1e-300 && fabs(apq) > 1e-18 * sqrt(fabs(app * aqq))) {
          const double theta = (aqq - app) / (2.0 * apq);
          const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
          c = 1.0 / sqrt(t * t + 1.0);
          s = t * c;
        }
        s_p[tid] = p;
        s_q[tid] = q;
        s_c[tid] = c;
        s_s[tid] = s;
        rotated = rotated || s != 0.0;
      }
      if (!__syncthreads_or(tid < kB && s_s[tid] != 0.0)) continue;  // nothing to rotate in this step
      // columns: S <- S J, Q <- Q J
      for (int e = tid; e < kB * kP; e += 256) {
        const int pr = e / kP, i = e % kP;
        const double c = s_c[pr], s = s_s[pr];
        if (s == 0.0) continue;
        const int p = s_p[pr], q = s_q[pr];
        const double aip = S[i][p], aiq = S[i][q];
        S[i][p] = c * aip - s * aiq;
        S[i][q] = s * aip + c * aiq;
        const double vip = Q[i][p], viq = Q[i][q];
        Q[i][p] = c * vip - s * viq;
        Q[i][q] = s * vip + c * viq;
      }
      __syncthreads();
      // rows: S <- J^T S
      for (int e = tid; e < kB * kP; e += 256) {
        const int pr = e / kP, i = e % kP;
        const double c = s_c[pr], s = s_s[pr];
        if (s == 0.0) continue;
        const int p = s_p[pr], q = s_q[pr];
        const double api = S[p][i], aqi = S[q][i];
        S[p][i] = c * api - s * aqi;
        S[q][i] = s * api + c * aqi;
      }
      __syncthreads();
    }
    if (!__syncthreads_or(rotated)) break;
  }
  double* out = Qout + (size_t)blockIdx.x * kP * kP;
  for (int e = tid; e < kP * kP; e += 256) out[e] = Q[e / kP][e % kP];
}

// One step's whole update in one pass: every 32 x 32 block (pair pi's rows x pair pj's columns) of A becomes
// Q_pi^T A_blk Q_pj, and every block (row tile, pair pj's columns) of V becomes V_blk Q_pj.  A CTA owns its
// block, so the update is in place without any ordering between CTAs.  grid (pairs, pairs, 2): z = 0 A, 1 V.
__global__ void __launch_bounds__(256) bj_update_kernel(double* __restrict__ A, double* __restrict__ V, int mp, int nb,
                                                       int step, const double* __restrict__ Qin) {
  __shared__ double X[kP][kP + 1], T[kP][kP + 1], Qi[kP][kP + 1], Qj[kP][kP + 1];
  const bool is_v = blockIdx.z == 1;
  double* M = is_v ? V : A;
  if (!M) return;
  int Ii, Ji, Ij, Jj;
  bj_pair(nb, step, blockIdx.y, &Ii, &Ji);
  bj_pair(nb, step, blockIdx.x, &Ij, &Jj);
  const double* qi = Qin + (size_t)blockIdx.y * kP * kP;
  const double* qj = Qin + (size_t)blockIdx.x * kP * kP;
  for (int e = threadIdx.x; e < kP * kP; e += 256) {
    const int r = e / kP, c = e % kP;
    // rows: the pair's rows for A, a plain 32-row tile for V
    const int gr = is_v ? (int)blockIdx.y * kP + r : (r < kB ? Ii * kB + r : Ji * kB + r - kB);
    const int gc = c < kB ? Ij * kB + c : Jj * kB + c - kB;
    X[r][c] = M[(size_t)gr * mp + gc];
    Qj[r][c] = qj[e];
    if (!is_v) Qi[r][c] = qi[e];
  }
  __syncthreads();
  // T = X Qj : thread -> row r, four columns
  const int r = threadIdx.x >> 3, c0 = (threadIdx.x & 7) * 4;
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
  for (int k = 0; k < kP; ++k) {
    const double x = X[r][k];
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i] += x * Qj[k][c0 + i];
  }
  if (!is_v) {
#pragma unroll
    for (int i = 0; i < 4; ++i) T[r][c0 + i] = acc[i];
    __syncthreads();
    // out = Qi^T T
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i] = 0.0;
#pragma unroll
    for (int k = 0; k < kP; ++k) {
      const double q = Qi[k][r];
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[i] += q * T[k][c0 + i];
    }
  }
  const int gr = is_v ? (int)blockIdx.y * kP + r : (r < kB ? Ii * kB + r : Ji * kB + r - kB);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = c0 + i;
    M[(size_t)gr * mp + (c < kB ? Ij * kB + c : Jj * kB + c - kB)] = acc[i];
  }
}

__global__ void bj_diag_kernel(const double* __restrict__ A, int mp, int m, double* __restrict__ evals) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < m) evals[k] = fmax(A[(size_t)k * mp + k], 0.0);  // S^2 >= 0
}

// C[i][j] = sum_k X[k][i] * Y[k][j]  (X^T Y; all n x n row-major).  64 x 64 tile per CTA, 4 x 4 per thread.
// For the warm start: T = G V0 (G symmetric, so G V0 = G^T V0) and A = V0^T T.
__global__ void __launch_bounds__(256) bj_gemm_tn_kernel(const double* __restrict__ X, const double* __restrict__ Y,
                                                        double* __restrict__ C, int n) {
  __shared__ double Xs[16][64 + 4], Ys[16][64 + 4];
  const int i0 = blockIdx.y * 64, j0 = blockIdx.x * 64;
  const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
  double acc[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
  for (int k0 = 0; k0 < n; k0 += 16) {
    for (int e = threadIdx.x; e < 16 * 64; e += 256) {
      const int kk = e >> 6, c = e & 63;
      Xs[kk][c] = (k0 + kk < n && i0 + c < n) ? X[(size_t)(k0 + kk) * n + i0 + c] : 0.0;
      Ys[kk][c] = (k0 + kk < n && j0 + c < n) ? Y[(size_t)(k0 + kk) * n + j0 + c] : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      double xa[4], yb[4];
#pragma unroll
      for (int a = 0; a < 4; ++a) xa[a] = Xs[kk][4 * ty + a];
#pragma unroll
      for (int b = 0; b < 4; ++b) yb[b] = Ys[kk][4 * tx + b];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] += xa[a] * yb[b];
    }
    __syncthreads();
  }
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int i = i0 + 4 * ty + a, j = j0 + 4 * tx + b;
      if (i < n && j < n) C[(size_t)i * n + j] = acc[a][b];
    }
}

// A <- (A + A^T) / 2: V0^T (G V0) is symmetric only up to rounding
__global__ void bj_symmetrise_kernel(double* __restrict__ A, int n) {
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= (size_t)n * n) return;
  const int i = (int)(e / n), j = (int)(e % n);
  if (i < j) {
    const double v = 0.5 * (A[(size_t)i * n + j] + A[(size_t)j * n + i]);
    A[(size_t)i * n + j] = v;
    A[(size_t)j * n + i] = v;
  }
}

int eig_padded_size(int m) {
  int nb = (m + kB - 1) / kB;
  if (nb & 1) ++nb;
  if (nb < 2) nb = 2;
  return nb * kB;
}

size_t eig_work_doubles(int m) {
  const size_t mp = (size_t)eig_padded_size(m);
  return 3 * mp * mp + (mp / kP) * kP * kP + 8;
}

// Eigendecomposition of G (d x d) without row/column j: evals[m], eigenvectors in the columns of
// V = work + mp*mp (row-major, row stride mp).  Synchronises the stream once per sweep (the stopping
// rule is evaluated on the host).  *sweeps_out (host memory) = sweeps run.
cudaError_t eig_block_jacobi(const double* G, int d, int j, double* work, double* evals, int* sweeps_out,
                             cudaStream_t s, const double* V0, double* Vsave) {
  const int m = d - 1, mp = eig_padded_size(m), nb = mp / kB;
  double* A = work;
  double* V = work + (size_t)mp * mp;
  double* Tm = V + (size_t)mp * mp;
  double* Q = Tm + (size_t)mp * mp;
  double* norms = Q + (size_t)(nb / 2) * kP * kP;
  const size_t n2 = (size_t)mp * mp;
  note_launch();
  bj_init_kernel<<<(unsigned)((n2 + 255) / 256), 256, 0, s>>>(G, d, j, mp, A, V0 ? nullptr : V);
  if (V0) {
    // warm start from an earlier decomposition of (nearly) the same matrix: iterate on V0^T G V0, which
    // is close to diagonal, and accumulate the rotations onto V0
    const dim3 gg((mp + 63) / 64, (mp + 63) / 64);
    note_launch();
    bj_gemm_tn_kernel<<<gg, 256, 0, s>>>(A, V0, Tm, mp);   // Tm = G V0
    note_launch();
    bj_gemm_tn_kernel<<<gg, 256, 0, s>>>(V0, Tm, A, mp);   // A = V0^T Tm
    note_launch();
    bj_symmetrise_kernel<<<(unsigned)((n2 + 255) / 256), 256, 0, s>>>(A, mp);
    cudaError_t e = cudaMemcpyAsync(V, V0, sizeof(double) * n2, cudaMemcpyDeviceToDevice, s);
    if (e != cudaSuccess) return e;
  }
  static const int inner = [] {
    const char* v = getenv("GS_EIG_INNER");
    const int k = v ? atoi(v) : kInnerSweeps;
    return k < 1 ? 1 : (k > 16 ? 16 : k);
  }();
  double prev_off = -1.0;
  int sweep = 0;
  for (; sweep < 40; ++sweep) {
    cudaError_t e = cudaMemsetAsync(norms, 0, 2 * sizeof(double), s);
    if (e != cudaSuccess) return e;
    note_launch();
    bj_norm_kernel<<<296, 256, 0, s>>>(A, mp, norms);
    double h[2];
    if ((e = cudaMemcpyAsync(h, norms, sizeof h, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return e;
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return e;
    // converged at off/||A|| <= 1e-15, or stagnating on the rounding floor
    if (h[0] <= 1e-30 * h[1] || h[1] == 0.0) break;
    if (h[0] <= 1e-24 * h[1] && prev_off >= 0.0 && h[0] > 0.25 * prev_off) break;
    prev_off = h[0];
    for (int step = 0; step < nb - 1; ++step) {
      note_launch();
      bj_pivot_kernel<<<nb / 2, 256, 0, s>>>(A, mp, nb, step, Q, inner);
      note_launch();
      bj_update_kernel<<<dim3(nb / 2, nb / 2, 2), 256, 0, s>>>(A, V, mp, nb, step, Q);
    }
  }
  note_launch();
  bj_diag_kernel<<<(m + 255) / 256, 256, 0, s>>>(A, mp, m, evals);
  if (Vsave) {
    cudaError_t e = cudaMemcpyAsync(Vsave, V, sizeof(double) * n2, cudaMemcpyDeviceToDevice, s);
    if (e != cudaSuccess) return e;
  }
  if (sweeps_out) *sweeps_out = sweep;
  g_eig_fits.fetch_add(1, std::memory_order_relaxed);
  g_eig_sweeps.fetch_add((unsigned long long)sweep, std::memory_order_relaxed);
  return cudaGetLastError();
}

}  // namespace gs
