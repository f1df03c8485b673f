// ub_lds.cu — shared-memory access patterns and packed-float32 throughput that decide the shape of the
// lanes=pods scoring kernel (B200).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ub_lds ub_lds.cu
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 2048

// PAT: 0 LDS.128 all lanes one address   1 LDS.128 4 addresses (lane/8), 16 B apart   2 same, 160 B apart
//      3 LDS.128 per-lane distinct        4 LDS.64 per-lane distinct                   5 LDS.32 one address
//      6 LDS.64 two addresses inside one 16 B record (lane parity)                    7 LDS.32 4 addresses 160 B apart
//      8 LDS.128 2 addresses (lane/16) 160 B apart
template <int PAT, bool USE = true>
__global__ void lds_k(float* out) {
  __shared__ __align__(16) float sm[8192];
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) sm[i] = (float)i * 1e-3f;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  unsigned base = (unsigned)__cvta_generic_to_shared(sm);
  unsigned off;
  if (PAT == 0 || PAT == 5) off = 0;
  else if (PAT == 1) off = (lane >> 3) * 16;
  else if (PAT == 2 || PAT == 7) off = (lane >> 3) * 160;
  else if (PAT == 3) off = lane * 16;
  else if (PAT == 4) off = lane * 8;
  else if (PAT == 6) off = (lane & 1) * 8;
  else off = (lane >> 4) * 160;
  float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
  unsigned a = base + off;
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const unsigned addr = a + (unsigned)(u * 640);
      if (PAT == 0 || PAT == 1 || PAT == 2 || PAT == 3 || PAT == 8) {
        float x, y, z, w;
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(x), "=f"(y), "=f"(z), "=f"(w) : "r"(addr));
        if (USE) { acc0 += x; acc1 += y; acc2 += z; acc3 += w; }
      } else if (PAT == 4 || PAT == 6) {
        float x, y;
        asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(x), "=f"(y) : "r"(addr));
        if (USE) { acc0 += x; acc1 += y; }
      } else {
        float x;
        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(addr));
        if (USE) acc0 += x;
      }
    }
    a = base + ((a - base + 5120u) & 16383u) ;
    if (a - base + 8 * 640 + 32 * 16 > 32768u) a = base + off;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc0 + acc1 + acc2 + acc3;
}

// OP: 0 FFMA  1 fma.rn.f32x2  2 add.rn.f32x2  3 FMNMX  4 3-input max  5 FFMA + FMNMX interleaved  6 FSEL
//     7 add.rz.f32x2  8 FFMA2 + FMNMX + LOP3 mix  9 MUFU.RCP + 6 FFMA
template <int OP>
__global__ void alu_k(float* out, float seed) {
  float f[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) f[i] = seed + threadIdx.x * 1e-3f + i;
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < 8; i += 2) {
      if (OP == 0) { f[i] = fmaf(f[i], 1.0000001f, 1e-9f); f[i + 1] = fmaf(f[i + 1], 1.0000001f, 1e-9f); }
      if (OP == 1) {
        unsigned long long v, m = 0x3F8000013F800001ull, c = 0x3089705F3089705Full;
        asm volatile("mov.b64 %0, {%1,%2};" : "=l"(v) : "f"(f[i]), "f"(f[i + 1]));
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v) : "l"(m), "l"(c));
        asm volatile("mov.b64 {%0,%1}, %2;" : "=f"(f[i]), "=f"(f[i + 1]) : "l"(v));
      }
      if (OP == 2 || OP == 7) {
        unsigned long long v, c = 0x3089705F3089705Full;
        asm volatile("mov.b64 %0, {%1,%2};" : "=l"(v) : "f"(f[i]), "f"(f[i + 1]));
        if (OP == 2) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(v) : "l"(c));
        else asm volatile("add.rz.f32x2 %0, %0, %1;" : "+l"(v) : "l"(c));
        asm volatile("mov.b64 {%0,%1}, %2;" : "=f"(f[i]), "=f"(f[i + 1]) : "l"(v));
      }
      if (OP == 3) { f[i] = fmaxf(f[i], f[i + 1] * 1.0f); f[i + 1] = fminf(f[i + 1], f[i]); }
      if (OP == 4) { f[i] = fmaxf(fmaxf(f[i], f[i + 1]), f[(i + 2) & 7]); f[i + 1] = fminf(fminf(f[i + 1], f[i]), f[(i + 3) & 7]); }
      if (OP == 5) { f[i] = fmaf(f[i], 1.0000001f, 1e-9f); f[i + 1] = fmaxf(f[i + 1], f[(i + 3) & 7]); }
      if (OP == 6) { f[i] = f[i + 1] > 1.5f ? f[i] : f[(i + 2) & 7]; f[i + 1] = f[i] > 2.5f ? f[i + 1] : f[(i + 3) & 7]; }
      if (OP == 8) {
        unsigned long long v, m = 0x3F8000013F800001ull, c = 0x3089705F3089705Full;
        asm volatile("mov.b64 %0, {%1,%2};" : "=l"(v) : "f"(f[i]), "f"(f[i + 1]));
        asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v) : "l"(m), "l"(c));
        asm volatile("mov.b64 {%0,%1}, %2;" : "=f"(f[i]), "=f"(f[i + 1]) : "l"(v));
        f[i] = fmaxf(f[i], f[(i + 2) & 7]);
        f[i + 1] = __uint_as_float((__float_as_uint(f[i + 1]) & 0xFFFFFFF8u) | 3u);
      }
      if (OP == 9) {
        float r;
        asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(f[i]));
        f[i] = fmaf(r, 1.0000001f, 1.5f);
        f[i + 1] = fmaf(f[i + 1], 1.0000001f, 1e-9f); f[i + 1] = fmaf(f[i + 1], 1.0000001f, 1e-9f);
        f[i + 1] = fmaf(f[i + 1], 1.0000001f, 1e-9f); f[i + 1] = fmaf(f[i + 1], 1.0000001f, 1e-9f);
        f[i + 1] = fmaf(f[i + 1], 1.0000001f, 1e-9f);
      }
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += f[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static void timeit(const char* name, F launch, double instr_per_thread, int blocks, int threads) {
  int dev; cudaGetDevice(&dev); cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
  launch(); cudaDeviceSynchronize();
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a); launch(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, dev);
  double warp_instr = (double)blocks * threads / 32 * instr_per_thread;
  printf("%-44s %8.3f ms  %7.3f warp-instr/clk/SM (at %d MHz nominal)  err=%s\n", name, ms,
         warp_instr / (ms * 1e-3) / p.multiProcessorCount / (clk * 1e3), clk / 1000, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  int dev = 0; cudaDeviceProp p; cudaGetDeviceProperties(&p, dev);
  int blocks = p.multiProcessorCount * 4, threads = 128;
  float* out; cudaMalloc(&out, sizeof(float) * blocks * 512);
#define L(P, NAME) timeit(NAME, [&] { lds_k<P><<<blocks, threads>>>(out); }, (double)ITERS * 8, blocks, threads)
  L(0, "LDS.128 one address (broadcast)");
  L(1, "LDS.128 4 addresses, 16 B apart");
  L(2, "LDS.128 4 addresses, 160 B apart");
  L(8, "LDS.128 2 addresses, 160 B apart");
  L(3, "LDS.128 32 addresses (lane*16)");
  L(4, "LDS.64 32 addresses (lane*8)");
  L(6, "LDS.64 2 addresses in one 16 B record");
  L(5, "LDS.32 one address (broadcast)");
  L(7, "LDS.32 4 addresses, 160 B apart");
  // the loads alone (results unused, 32 warps per SM): the shared-memory pipe's own rate
  threads = 256;
#define LN(P, NAME) timeit(NAME, [&] { lds_k<P, false><<<blocks, threads>>>(out); }, (double)ITERS * 8, blocks, threads)
  LN(0, "[loads only] LDS.128 one address");
  LN(2, "[loads only] LDS.128 4 addresses, 160 B apart");
  LN(3, "[loads only] LDS.128 32 addresses");
  LN(4, "[loads only] LDS.64 32 addresses");
  LN(6, "[loads only] LDS.64 2 addresses in 16 B");
  LN(5, "[loads only] LDS.32 one address");
  LN(7, "[loads only] LDS.32 4 addresses");
  threads = 512;
#define A(O, NAME, N) timeit(NAME, [&] { alu_k<O><<<blocks, threads>>>(out, 1.0f); }, (double)ITERS * N, blocks, threads)
  A(0, "FFMA", 8);
  A(1, "fma.rn.f32x2 (4 packed = 8 lanes-ops)", 4);
  A(2, "add.rn.f32x2", 4);
  A(7, "add.rz.f32x2", 4);
  A(3, "FMNMX (+FMUL)", 12);
  A(4, "3-input max/min", 8);
  A(5, "FFMA + FMNMX interleaved", 8);
  A(6, "FSETP+FSEL", 16);
  A(8, "FFMA2 + FMNMX + LOP3 mix", 12);
  A(9, "MUFU.RCP + 6 FFMA", 28);
  return 0;
}
