// tridiag.h — BayesianRidge in a tridiagonal basis (tridiag.cu) for complete.cu
#pragma once
#include <cstddef>
#include <cuda_runtime.h>

namespace gs {
size_t tri_work_doubles(int m);  // doubles of work space for an m x m problem
int tri_max_m(int n_sm);         // widest problem the reduction keeps in shared memory on this device
// One feature's estimator: Householder reduction of (G without row/column j) to tridiagonal T, then the
// evidence iteration in that basis; writes the d-1 coefficients, the intercept and the iteration count.
// No host synchronisation.  `n_sm` = SMs of the device `s` runs on.
cudaError_t tri_bayes_ridge(const double* G, const double* mu, const int* n_obs, int d, int j, double* work, int n_sm,
                            double* coef_out, double* intercept_out, int* iters_out, cudaStream_t s);
}  // namespace gs
