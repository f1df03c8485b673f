// ub_exchange.cu — how fast can the CTAs of one grid exchange a vector per step?  (tridiag.cu's inner loop)
// Every step each CTA publishes its share of m values and needs all m before the next step.
//   nvcc -arch=sm_100a -O3 -rdc=true -o ub_exchange ub_exchange.cu && ./ub_exchange
// modes: 0 tagged 16-byte slots, every thread polls its own slot(s), slots interleaved over CTAs (row i -> CTA i % nc)
//        1 as 0, with a CTA barrier between the arithmetic and the polling
//        2 as 0, with __nanosleep(200) between polls
//        3 tagged 8-byte slots
//        4 per-CTA flags (release/acquire), plain data, thread t polls CTA t's flag
//        5 one atomic counter, thread 0 polls
//        6 cooperative groups grid.sync()
//        7 as 0, but only warp 0 polls (all slots), the rest wait at the CTA barrier
#include <cooperative_groups.h>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

__device__ __forceinline__ void publish16(double* slot, double p, double a, unsigned tag) {
  const unsigned long long pb = ((unsigned long long)__double_as_longlong(p) & ~1ull) | tag;
  const unsigned long long ab = ((unsigned long long)__double_as_longlong(a) & ~1ull) | tag;
  asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(slot), "l"(pb), "l"(ab) : "memory");
}
__device__ __forceinline__ double await16(const double* slot, unsigned tag, int sleep) {
  unsigned long long pb, ab;
  unsigned spins = 0;
  for (;;) {
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(pb), "=l"(ab) : "l"(slot) : "memory");
    if ((pb & 1ull) == tag && (ab & 1ull) == tag) break;
    if (++spins > (1u << 24)) __trap();
    if (sleep) __nanosleep(sleep);
  }
  return __longlong_as_double((long long)(pb & ~1ull));
}
__device__ __forceinline__ void publish8(double* slot, double p, unsigned tag) {
  const unsigned long long pb = ((unsigned long long)__double_as_longlong(p) & ~1ull) | tag;
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(slot), "l"(pb) : "memory");
}
__device__ __forceinline__ double await8(const double* slot, unsigned tag) {
  unsigned long long pb;
  unsigned spins = 0;
  for (;;) {
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(pb) : "l"(slot) : "memory");
    if ((pb & 1ull) == tag) break;
    if (++spins > (1u << 24)) __trap();
  }
  return __longlong_as_double((long long)(pb & ~1ull));
}

__global__ void __launch_bounds__(512, 1) exchange_kernel(double* buf, unsigned* flags, int m, int steps, int mode, int delay,
                                                         double* sink, long long* cycles) {
  __shared__ double s_acc[1024];
  const int tid = threadIdx.x, nt = blockDim.x, nc = gridDim.x, c = blockIdx.x;
  const int R = (m + nc - 1) / nc;
  cg::grid_group grid = cg::this_grid();
  double acc = 0.0;
  const long long t0 = clock64();
  for (int k = 0; k < steps; ++k) {
    double* set = buf + (size_t)(k & 1) * 2 * 1024;
    const unsigned tag = (unsigned)(k >> 1) & 1u;
    auto slot_of = [&](int i) { return i; };
    if (delay) {  // the arithmetic between two exchanges: everybody `delay` cycles, the publishing warps 400 more
      const long long until = clock64() + delay + ((tid >> 5) < R ? 400 : 0);
      while (clock64() < until) {
      }
      if (mode == 1) __syncthreads();  // mode 1: nobody polls before this CTA has published
    }
    // publish this CTA's rows: warp w -> local row w (as tridiag.cu: lane 0 of the row's warp)
    for (int lr = tid >> 5; lr < R; lr += nt >> 5) {
      const int i = c + nc * lr;
      if (i >= m || (tid & 31)) continue;
      const double v = (double)(i + k);
      if (mode == 3) publish8(set + slot_of(i), v, tag);
      else if (mode <= 2 || mode == 7) publish16(set + 2 * (size_t)slot_of(i), v, v, tag);
      else set[i] = v;
    }
    if (mode == 4) {
      __syncthreads();
      if (tid == 0) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(flags + c), "r"((unsigned)k + 1) : "memory");
      if (tid < nc) {
        unsigned v, spins = 0;
        do {
          asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + tid) : "memory");
          if (++spins > (1u << 24)) __trap();
        } while (v < (unsigned)k + 1);
      }
      __syncthreads();
    } else if (mode == 5) {
      __syncthreads();
      if (tid == 0) {
        __threadfence();
        atomicAdd(flags, 1u);
        unsigned v, spins = 0;
        do {
          asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flags) : "memory");
          if (++spins > (1u << 24)) __trap();
        } while (v < (unsigned)(k + 1) * nc);
      }
      __syncthreads();
    } else if (mode == 6) {
      grid.sync();
    }
    if (mode == 7) {
      if (tid < 32) for (int i = tid; i < m; i += 32) s_acc[i] = await16(set + 2 * (size_t)i, tag, 0);
      __syncthreads();
      for (int i = tid; i < m; i += nt) acc += s_acc[i];
    } else {
      for (int i = tid; i < m; i += nt) {
        if (mode == 3) acc += await8(set + slot_of(i), tag);
        else if (mode <= 2) acc += await16(set + 2 * (size_t)slot_of(i), tag, mode == 2 ? 200 : 0);
        else acc += __ldcg(set + i);
      }
    }
    __syncthreads();  // the block-wide reduction that follows in the real kernel
  }
  if (c == 0 && tid == 0) *cycles = clock64() - t0;
  sink[(size_t)c * nt + tid] = acc;
}

int main() {
  setvbuf(stdout, nullptr, _IONBF, 0);
  double *buf, *sink;
  unsigned* flags;
  long long* cyc;
  cudaMalloc(&buf, sizeof(double) * 4 * 1024);
  cudaMalloc(&sink, sizeof(double) * 148 * 512);
  cudaMalloc(&flags, sizeof(unsigned) * 1024);
  cudaMalloc(&cyc, sizeof(long long));
  const int steps = 1000;
  const int cfg[][3] = {{64, 255, 256}, {148, 999, 512}, {74, 999, 512}};
  for (auto& cf : cfg) {
   for (int delay = 0; delay <= 3000; delay += 3000) {
    for (int mode = 0; mode < 8; ++mode) {
      if (delay && mode >= 4) continue;
      cudaMemset(buf, 0xFF, sizeof(double) * 4 * 1024);
      cudaMemset(flags, 0, sizeof(unsigned) * 1024);
      int m = cf[1], st = steps, md = mode;
      int dl = delay;
      void* args[] = {&buf, &flags, &m, &st, &md, &dl, &sink, &cyc};
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0);
      cudaEventCreate(&e1);
      cudaEventRecord(e0);
      cudaError_t e = cudaLaunchCooperativeKernel((const void*)exchange_kernel, dim3(cf[0]), dim3(cf[2]), args, 0, 0);
      cudaEventRecord(e1);
      cudaError_t e2 = cudaDeviceSynchronize();
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      long long h = 0;
      cudaMemcpy(&h, cyc, sizeof h, cudaMemcpyDeviceToHost);
      printf("ctas %3d m %4d threads %3d delay %4d mode %d: %7.3f us/step (%6.0f cycles)  %s %s\n", cf[0], cf[1], cf[2], delay, mode,
             1e3 * ms / steps, (double)h / steps, cudaGetErrorString(e), cudaGetErrorString(e2));
      if (e != cudaSuccess || e2 != cudaSuccess) return 1;
    }
   }
  }
  return 0;
}
