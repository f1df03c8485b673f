// eig_jacobi.h — block Jacobi eigensolver (eig_jacobi.cu) for complete.cu
#pragma once
#include <cstddef>
#include <cuda_runtime.h>

namespace gs {
int eig_padded_size(int m);        // mp: row stride of the work matrices for an m x m problem
size_t eig_work_doubles(int m);    // doubles of work space: A (mp*mp), V (mp*mp), rotations, norms
// V0 (optional, mp x mp): eigenvectors of an earlier, nearby matrix to start from; Vsave (optional): a copy of V
cudaError_t eig_block_jacobi(const double* G, int d, int j, double* work, double* evals, int* sweeps_host,
                             cudaStream_t s, const double* V0 = nullptr, double* Vsave = nullptr);
void eig_stats(unsigned long long* fits, unsigned long long* sweeps);  // decompositions and sweeps so far (process-wide)
}  // namespace gs
