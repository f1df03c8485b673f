// ub_alu.cu — per-SM throughput of the arithmetic the scoring kernel is made of (B200).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o ub_alu ub_alu.cu
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
#define ILP 4
template <int OP>
__global__ void k(double* out, double seed) {
  double a[ILP];
  float f[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) { a[i] = seed + threadIdx.x * 1e-3 + i; f[i] = (float)a[i]; }
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      if (OP == 0) a[i] = __fma_rn(a[i], 1.0000001, 1e-9);
      if (OP == 1) a[i] = __dadd_rn(a[i], 1e-9);
      if (OP == 2) a[i] = __dmul_rn(a[i], 1.0000001);
      if (OP == 3) a[i] = __drcp_rn(__dadd_rn(a[i], 1.5));
      if (OP == 4) a[i] = __ddiv_rn(a[i], 3.0);
      if (OP == 5) a[i] = __ddiv_rn(1.0 + a[i], a[i] + 2.0);
      if (OP == 6) f[i] = __fmaf_rn(f[i], 1.0000001f, 1e-9f);
      if (OP == 7) f[i] = __fdiv_rn(1.0f, __fadd_rn(f[i], 1.5f));
      if (OP == 8) f[i] = __frcp_rn(__fadd_rn(f[i], 1.5f));
      if (OP == 9) { float r; asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(f[i] + 1.5f)); f[i] = r; }
      if (OP == 10) a[i] = (double)(float)a[i] + 1e-9;                 // F2F both ways + DADD
      if (OP == 11) a[i] = (double)__double2int_rz(a[i]) + 1.5;        // F2I + I2F
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += a[i] + f[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int OP>
void run(const char* name, int opsPerIter) {
  int dev; cudaGetDevice(&dev); cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
  int blocks = p.multiProcessorCount * 4, threads = 512;
  double* out; cudaMalloc(&out, sizeof(double) * blocks * threads);
  k<OP><<<blocks, threads>>>(out, 1.0); cudaDeviceSynchronize();
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a); k<OP><<<blocks, threads>>>(out, 1.0); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  double ops = (double)blocks * threads * ITERS * ILP * opsPerIter;
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, dev);
  printf("%-28s %8.3f ms  %8.2f Gop/s  %6.2f ops/clk/SM (at %d MHz nominal)\n", name, ms, ops / ms * 1e-6,
         ops / (ms * 1e-3) / p.multiProcessorCount / (clk * 1e3), clk / 1000);
  cudaFree(out);
}
int main() {
  run<0>("DFMA", 1); run<1>("DADD", 1); run<2>("DMUL", 1); run<3>("DADD+__drcp_rn", 1); run<4>("__ddiv_rn(x,3)", 1);
  run<5>("2 DADD+__ddiv_rn(x,y)", 1); run<6>("FFMA", 1); run<7>("FADD+__fdiv_rn(1,x)", 1); run<8>("FADD+__frcp_rn", 1);
  run<9>("FADD+rcp.approx", 1); run<10>("F2F.64.32+F2F.32.64+DADD", 1); run<11>("F2I+I2F+DADD", 1);
  return 0;
}
