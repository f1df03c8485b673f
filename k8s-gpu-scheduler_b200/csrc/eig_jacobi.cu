// eig_jacobi.cu — symmetric eigendecomposition for the recommender's completion (sm_100a).
//
// BayesianRidge.fit takes an SVD of the centred design matrix per feature
// (pkg/recommender/recommender/recommender.py:15-17 -> sklearn linear_model/_bayes.py:262-306:
// eigen_vals_ = S^2, Vh); here that is the eigendecomposition of the centred Gram matrix G without
// row/column j.  Up to 24 columns complete.cu's one-CTA Jacobi serves; beyond, tridiag.cu needs no
// eigendecomposition at all and is what runs by default.  This file is the other round-2 solver (round 1 called
// cuSOLVER's syevd there), selected with GS_EIG_TRI=0 or when tridiag.cu's matrix would not fit shared memory,
// and kept as a cross-check: a two-sided BLOCK Jacobi across CTAs.
//
//   The m x m matrix is padded to mp = nb * kB (nb even) and split into nb block columns of kB = 16.
//   A sweep is nb - 1 steps of a round-robin tournament; in a step the nb / 2 disjoint block pairs
//   (I, J) are processed independently:
//     bj_pivot_kernel  one CTA per pair: the 32 x 32 pivot [A_II A_IJ; A_JI A_JJ] is rotated in shared memory
//                      (16 Jacobi rotations per parallel step; all index pairs in step 0 of a sweep, only the
//                      pairs across the two blocks afterwards) and the accumulated rotation Q (32 x 32,
//                      orthogonal) written out;
//     bj_update_kernel every 32 x 32 block (pair's rows x pair's columns) of A becomes Q_i^T A_blk Q_j (upper
//                      triangle of blocks computed, mirrored below) and every block of V becomes V_blk Q_j —
//                      one pass, each CTA owns what it writes; the products run on the FP64 tensor cores.
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

__device__ __forceinline__ double bj_rcp_approx(double x) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  return y;
}

// The Jacobi rotation for the pair (p, q): t = sgn(theta) / (|theta| + sqrt(theta^2 + 1)),
// theta = (aqq - app) / (2 apq), c = 1 / sqrt(1 + t^2), s = t c.  This is the critical path of every pivot
// step, so t is computed in float32 from an approximate quotient: with t good to ~1e-7 the rotated apq
// shrinks by that factor instead of vanishing, and the next sweep (which is needed anyway) takes care of
// the rest.  c gets two double-precision Newton steps: c^2 (1 + t^2) = 1 to rounding — whatever t is — is what
// keeps the accumulated Q orthogonal.  No branches: a rotation that would change nothing becomes (1, 0).
__device__ __forceinline__ void bj_angle(double app, double aqq, double apq, double* c, double* s) {
  const bool rotate = fabs(apq) > 1e-300 && fabs(apq) > 1e-18 * fmax(fabs(app), fabs(aqq));
  const double dl = aqq - app, u = 2.0 * apq;
  const bool big = fabs(dl) >= fabs(u);
  const double num = big ? u : dl, den = big ? dl : u;  // |num / den| <= 1; |den| >= 2 |apq| is not subnormal
  const float r = (float)(num * bj_rcp_approx(rotate ? den : 1.0));  // big: 1 / theta, else theta
  const float w1 = fmaf(r, r, 1.0f);  // in [1, 2]
  const float sq = w1 * rsqrtf(w1);
  const float tn = big ? r : (r >= 0.0f ? 1.0f : -1.0f);
  const float td = (big ? 1.0f : fabsf(r)) + sq;
  const double t = rotate ? (double)__fdividef(tn, td) : 0.0;
  const double w = fma(t, t, 1.0);  // in [1, 2]
  double yc = (double)rsqrtf((float)w);
  yc = yc * fma(-0.5 * w, yc * yc, 1.5);
  yc = yc * fma(-0.5 * w, yc * yc, 1.5);
  *c = yc;
  *s = t * yc;
}

// One CTA per block pair: rotate the 32 x 32 pivot [A_II A_IJ; A_JI A_JJ] in shared memory and write the
// accumulated rotation Q.  `full` (step 0 of a sweep, where every block sits in exactly one pair): a cyclic
// sweep over all 496 index pairs, 31 parallel steps of 16 rotations; otherwise only the 256 pairs ACROSS the
// two blocks, 16 parallel steps — so an outer sweep rotates every index pair of the matrix exactly once.
// A parallel step is one barrier: thread (a, b) owns the 2 x 2 block (pair a's rows) x (pair b's columns),
// reads it from one copy of S, applies both rotations and writes it to the other copy; lane b of every warp
// derives pair b's rotation, pair a's arrives by shuffle.
__global__ void __launch_bounds__(256) bj_pivot_kernel(const double* __restrict__ A, int mp, int nb, int step,
                                                      double* __restrict__ Qout, int inner_sweeps, int full) {
  __shared__ double S[2][kP][kP + 1], Q[kP][kP + 1];
  const int tid = threadIdx.x;
  int I, J;
  bj_pair(nb, step, blockIdx.x, &I, &J);
  for (int e = tid; e < kP * kP; e += 256) {
    const int r = e / kP, c = e % kP;
    const int gr = (r < kB ? I * kB + r : J * kB + r - kB), gc = (c < kB ? I * kB + c : J * kB + c - kB);
    S[0][r][c] = A[(size_t)gr * mp + gc];
    Q[r][c] = r == c ? 1.0 : 0.0;
  }
  __syncthreads();
  const int a = tid >> 4, b = tid & 15;
  const int nst = full ? kP - 1 : kB;
  int cur = 0;
  for (int sweep = 0; sweep < inner_sweeps; ++sweep) {
    for (int st = 0; st < nst; ++st) {
      int pa, qa, pb, qb;
      if (full) {
        // round robin over the 32 indices, smaller index first
        int x = a == 0 ? kP - 1 : (st + a) % (kP - 1), y = a == 0 ? st : (st - a + (kP - 1)) % (kP - 1);
        pa = min(x, y);
        qa = max(x, y);
        x = b == 0 ? kP - 1 : (st + b) % (kP - 1);
        y = b == 0 ? st : (st - b + (kP - 1)) % (kP - 1);
        pb = min(x, y);
        qb = max(x, y);
      } else {
        pa = a;
        qa = kB + ((a + st) & (kB - 1));
        pb = b;
        qb = kB + ((b + st) & (kB - 1));
      }
      const double(*Sc)[kP + 1] = S[cur];
      double(*Sn)[kP + 1] = S[cur ^ 1];
      // every load of the step first, so that they overlap the rotation arithmetic
      const double bpp = Sc[pb][pb], bqq = Sc[qb][qb], bpq = Sc[pb][qb];
      const double x00 = Sc[pa][pb], x01 = Sc[pa][qb], x10 = Sc[qa][pb], x11 = Sc[qa][qb];
      const double v0 = Q[a][pb], v1 = Q[a][qb], v2 = Q[a + kB][pb], v3 = Q[a + kB][qb];
      double cb, sb;
      bj_angle(bpp, bqq, bpq, &cb, &sb);
      const double ca = __shfl_sync(0xffffffffu, cb, a), sa = __shfl_sync(0xffffffffu, sb, a);  // lane a holds pair a
      // columns (pair b), then rows (pair a): J^T S J
      const double y00 = cb * x00 - sb * x01, y01 = sb * x00 + cb * x01;
      const double y10 = cb * x10 - sb * x11, y11 = sb * x10 + cb * x11;
      Sn[pa][pb] = ca * y00 - sa * y10;
      Sn[pa][qb] = ca * y01 - sa * y11;
      Sn[qa][pb] = sa * y00 + ca * y10;
      Sn[qa][qb] = sa * y01 + ca * y11;
      // Q <- Q J: rows a and a + 16, pair b's columns (in place: nobody else touches them in this step)
      Q[a][pb] = cb * v0 - sb * v1;
      Q[a][qb] = sb * v0 + cb * v1;
      Q[a + kB][pb] = cb * v2 - sb * v3;
      Q[a + kB][qb] = sb * v2 + cb * v3;
      cur ^= 1;
      __syncthreads();
    }
  }
  double* out = Qout + (size_t)blockIdx.x * kP * kP;
  for (int e = tid; e < kP * kP; e += 256) out[e] = Q[e / kP][e % kP];
}

// C (8 x 8, two elements per lane) += A (8 x 4, row-major fragment) * B (4 x 8, column-major fragment) on the
// FP64 tensor cores.  Lane l holds A[l / 4][l % 4], B[l % 4][l / 4], C[l / 4][2 (l % 4) + {0, 1}].
__device__ __forceinline__ void bj_dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

constexpr int kLd = kP + 4;  // shared row stride: the DMMA fragment loads below touch 16 distinct banks per half warp

// One step's update, A and V in one launch.
//   A: CTAs over the upper triangle of pair blocks; the 32 x 32 block (pair pi's rows) x (pair pj's columns),
//      pi <= pj, becomes Z = Q_pi^T A_blk Q_pj and its mirror image (pj, pi) becomes Z^T — A stays exactly
//      symmetric off the pivots and a third of the step's flops goes away.
//   V: CTAs over (row tile of 32 * vt rows, pair pj): every 32-row slab becomes V_blk Q_pj, Q_pj staged once.
// A CTA owns what it writes, so the update is in place without any ordering between CTAs.  Products on the
// FP64 tensor cores: warp w owns the 8-row band w / 2 and two 8-column tiles of a 32 x 32 result.
__global__ void __launch_bounds__(256) bj_update_kernel(double* __restrict__ A, double* __restrict__ V, int mp, int nb,
                                                       int step, const double* __restrict__ Qin, int vt) {
  __shared__ double X[kP][kLd], T[kP][kLd], Qi[kP][kLd], Qj[kP][kLd];
  const int np = nb / 2, n_a = np * (np + 1) / 2;
  const bool is_v = (int)blockIdx.x >= n_a;
  int bi, bj;  // A: pair indices (rows, columns); V: row tile, pair index
  if (is_v) {
    bi = ((int)blockIdx.x - n_a) / np;
    bj = ((int)blockIdx.x - n_a) % np;
  } else {
    int rest = (int)blockIdx.x;
    bi = 0;
    while (rest >= np - bi) {
      rest -= np - bi;
      ++bi;
    }
    bj = bi + rest;
  }
  double* M = is_v ? V : A;
  int Ii = 0, Ji = 0, Ij, Jj;
  if (!is_v) bj_pair(nb, step, bi, &Ii, &Ji);
  bj_pair(nb, step, bj, &Ij, &Jj);
  const double* qi = Qin + (size_t)bi * kP * kP;
  const double* qj = Qin + (size_t)bj * kP * kP;
  auto col_of = [&](int c) { return c < kB ? Ij * kB + c : Jj * kB + c - kB; };
  const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
  const int fr = 8 * (wrp >> 1) + (lane >> 2), fk = lane & 3;  // fragment row / k offset
  const int fc = 16 * (wrp & 1) + (lane >> 2);                 // first tile's fragment column
  const int cr = fr, cc = 16 * (wrp & 1) + 2 * (lane & 3);     // accumulator row / first column
  for (int e = threadIdx.x; e < kP * kP; e += 256) {
    Qj[e / kP][e % kP] = qj[e];
    if (!is_v) Qi[e / kP][e % kP] = qi[e];
  }
  const int n_sub = is_v ? vt : 1;
  for (int sub = 0; sub < n_sub; ++sub) {
    auto row_of = [&](int r) { return is_v ? (bi * vt + sub) * kP + r : (r < kB ? Ii * kB + r : Ji * kB + r - kB); };
    if (sub) __syncthreads();  // the previous slab's stores out of X are done
    for (int e = threadIdx.x; e < kP * kP; e += 256) {
      const int r = e / kP, c = e % kP;
      X[r][c] = M[(size_t)row_of(r) * mp + col_of(c)];
    }
    __syncthreads();
    double acc[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
    // T = X Qj
#pragma unroll
    for (int k0 = 0; k0 < kP; k0 += 4) {
      const double a = X[fr][k0 + fk];
      bj_dmma(acc[0][0], acc[0][1], a, Qj[k0 + fk][fc]);
      bj_dmma(acc[1][0], acc[1][1], a, Qj[k0 + fk][fc + 8]);
    }
    if (!is_v) {
#pragma unroll
      for (int t = 0; t < 2; ++t) {
        T[cr][cc + 8 * t] = acc[t][0];
        T[cr][cc + 8 * t + 1] = acc[t][1];
        acc[t][0] = acc[t][1] = 0.0;
      }
      __syncthreads();
      // Z = Qi^T T
#pragma unroll
      for (int k0 = 0; k0 < kP; k0 += 4) {
        const double a = Qi[k0 + fk][fr];
        bj_dmma(acc[0][0], acc[0][1], a, T[k0 + fk][fc]);
        bj_dmma(acc[1][0], acc[1][1], a, T[k0 + fk][fc + 8]);
      }
    }
    // through shared memory (X again, once every warp is done reading it), then row segments to global memory
    __syncthreads();
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      X[cr][cc + 8 * t] = acc[t][0];
      X[cr][cc + 8 * t + 1] = acc[t][1];
    }
    __syncthreads();
    const bool mirror = !is_v && bi != bj;
    for (int e = threadIdx.x; e < kP * kP; e += 256) {
      const int r = e / kP, c = e % kP;
      M[(size_t)row_of(r) * mp + col_of(c)] = X[r][c];
      if (mirror) {
        // block (bj, bi): rows are pair bj's, columns pair bi's; element (r, c) there is Z[c][r]
        const int gr = r < kB ? Ij * kB + r : Jj * kB + r - kB, gc = c < kB ? Ii * kB + c : Ji * kB + c - kB;
        M[(size_t)gr * mp + gc] = X[c][r];
      }
    }
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

// work: A, V, Tm (mp * mp each), one step's rotations, norms
static size_t q_set_doubles(int nb) { return (size_t)(nb / 2) * kP * kP; }

size_t eig_work_doubles(int m) {
  const size_t mp = (size_t)eig_padded_size(m);
  const int nb = (int)(mp / kB);
  return 3 * mp * mp + q_set_doubles(nb) + 8;
}

// Eigendecomposition of G (d x d) without row/column j: evals[m], eigenvectors in the columns of
// V = work + mp*mp (row-major, row stride mp).  Synchronises the stream once per sweep (the stopping
// rule is evaluated on the host).  *sweeps_out (host memory) = sweeps run.
cudaError_t eig_block_jacobi(const double* G, int d, int j, double* work, double* evals, int* sweeps_out,
                             cudaStream_t s, const double* V0, double* Vsave) {
  const int m = d - 1, mp = eig_padded_size(m), nb = mp / kB, np = nb / 2;
  double* A = work;
  double* V = work + (size_t)mp * mp;
  double* Tm = V + (size_t)mp * mp;
  double* Q = Tm + (size_t)mp * mp;
  double* norms = Q + q_set_doubles(nb);
  const size_t n2 = (size_t)mp * mp;
  const int n_tri = np * (np + 1) / 2;
  const int vt = (mp >= 512 && np % 2 == 0) ? 2 : 1;  // 64-row slabs of V per CTA for the wide matrices
  const int n_vcta = (mp / (kP * vt)) * np;
  cudaError_t e;
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
    if ((e = cudaMemcpyAsync(V, V0, sizeof(double) * n2, cudaMemcpyDeviceToDevice, s)) != cudaSuccess) return e;
  }
  static const int inner = [] {
    const char* v = getenv("GS_EIG_INNER");
    const int k = v ? atoi(v) : kInnerSweeps;
    return k < 1 ? 1 : (k > 16 ? 16 : k);
  }();
  // GS_EIG_CROSS=0: a full cyclic sweep of every pivot (measurement; rotates pairs inside a block nb - 1 times per sweep)
  static const int cross_only = [] {
    const char* v = getenv("GS_EIG_CROSS");
    return v ? atoi(v) : 1;
  }();
  double prev_off = -1.0;
  int sweep = 0;
  for (; sweep < 40; ++sweep) {
    if ((e = cudaMemsetAsync(norms, 0, 2 * sizeof(double), s)) != cudaSuccess) return e;
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
      bj_pivot_kernel<<<np, 256, 0, s>>>(A, mp, nb, step, Q, inner, step == 0 || cross_only == 0);
      note_launch();
      bj_update_kernel<<<n_tri + n_vcta, 256, 0, s>>>(A, V, mp, nb, step, Q, vt);
    }
  }
  note_launch();
  bj_diag_kernel<<<(m + 255) / 256, 256, 0, s>>>(A, mp, m, evals);
  if (Vsave && (e = cudaMemcpyAsync(Vsave, V, sizeof(double) * n2, cudaMemcpyDeviceToDevice, s)) != cudaSuccess) return e;
  if (sweeps_out) *sweeps_out = sweep;
  g_eig_fits.fetch_add(1, std::memory_order_relaxed);
  g_eig_sweeps.fetch_add((unsigned long long)sweep, std::memory_order_relaxed);
  return cudaGetLastError();
}

}  // namespace gs
