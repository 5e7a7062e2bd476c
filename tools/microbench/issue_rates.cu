// issue_rates.cu — per-SM instruction throughput of the pipes the rasteriser leans on (B200, sm_100a): scalar FFMA,
// packed FFMA2 (fma.rn.f32x2), MUFU.EX2, SHFL, LDS.128 and fp32 RED.  Each kernel runs ITER iterations of 8 independent
// chains per thread, 1024 threads per CTA, 2 CTAs per SM; the result is warp-instructions per clock per SM (from the
// kernel's own clock64 span) — the ceiling a kernel bound by that pipe can reach.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o issue_rates issue_rates.cu && ./issue_rates
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdint.h>

constexpr int ITER = 4096;

__device__ __forceinline__ unsigned long long pk(float a, float b) {
  unsigned long long r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}

template <int MODE>
__global__ void __launch_bounds__(1024) k_rate(float *out, long long *clocks, float seed) {
  __shared__ float4 sm[1024];
  sm[threadIdx.x] = make_float4(seed, seed, seed, seed);
  __syncthreads();
  float a[8];
  unsigned long long p[8];
  for (int i = 0; i < 8; ++i) { a[i] = seed + i + threadIdx.x; p[i] = pk(a[i], a[i] + 1.f); }
  const float m = 0.999f + seed * 1e-9f, c = 1e-3f + seed;
  const unsigned long long m2 = pk(m, m), c2 = pk(c, c);
  long long t0 = clock64();
  for (int it = 0; it < ITER; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) a[i] = fmaf(a[i], m, c);
      if (MODE == 1) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(m2), "l"(c2));
      if (MODE == 2) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
      if (MODE == 3) a[i] = __shfl_xor_sync(0xffffffffu, a[i], 1 + (i & 15));
      if (MODE == 4) { float4 v = sm[(threadIdx.x + i * 32 + it) & 1023]; a[i] += v.x + v.w; }
      if (MODE == 5) a[i] = fminf(a[i] + c, m);  // FADD + FMNMX (alu pipe share)
    }
  }
  long long t1 = clock64();
  float s = 0.f;
  for (int i = 0; i < 8; ++i) {
    float lo, hi;
    asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[i]));
    s += a[i] + lo + hi;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) clocks[blockIdx.x] = t1 - t0;
}

__global__ void __launch_bounds__(256) k_red(float *acc, int rows, int lanes_active, long long *clocks) {
  // one RED.ADD.F32 vector per warp per iteration: `lanes_active` lanes add to consecutive floats of a pseudo-random row
  const int lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  uint32_t r = warp * 2654435761u + 12345u;
  long long t0 = clock64();
  for (int it = 0; it < 2048; ++it) {
    r = r * 1664525u + 1013904223u;
    if (lane < lanes_active) atomicAdd(acc + (size_t)(r % (uint32_t)rows) * 12 + lane, 1.0f);
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) clocks[blockIdx.x] = t1 - t0;
}

template <int MODE>
static void run(const char *name, int per_iter_instr, int sms) {
  int blocks = sms * 2;
  float *out; long long *clk;
  cudaMalloc(&out, sizeof(float) * blocks * 1024);
  cudaMalloc(&clk, sizeof(long long) * blocks);
  k_rate<MODE><<<blocks, 1024>>>(out, clk, 0.5f);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k_rate<MODE><<<blocks, 1024>>>(out, clk, 0.5f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long h[1024]; cudaMemcpy(h, clk, sizeof(long long) * blocks, cudaMemcpyDeviceToHost);
  double mean = 0; for (int i = 0; i < blocks; ++i) mean += h[i]; mean /= blocks;
  // per SM: 2 CTAs x 32 warps, each ITER * 8 * per_iter_instr warp-instructions in `mean` clocks
  double wi_per_clk = 2.0 * 32.0 * ITER * 8.0 * per_iter_instr / mean;
  printf("{\"pipe\": \"%s\", \"warp_instr_per_clk_per_sm\": %.3f, \"clocks\": %.0f, \"ms\": %.3f}\n", name, wi_per_clk, mean, ms);
  cudaFree(out); cudaFree(clk);
}

int main() {
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  printf("{\"sms\": %d}\n", sms);
  run<0>("FFMA", 1, sms);
  run<1>("FFMA2 (2 fp32 FMA per lane)", 1, sms);
  run<2>("MUFU.EX2", 1, sms);
  run<3>("SHFL", 1, sms);
  run<4>("LDS.128 (+2 FADD)", 1, sms);
  run<5>("FADD+FMNMX", 2, sms);
  for (int lanes : {1, 12, 32}) {
    int rows = 50000, blocks = sms * 8;
    float *acc; long long *clk;
    cudaMalloc(&acc, sizeof(float) * 12 * rows + 256); cudaMemset(acc, 0, sizeof(float) * 12 * rows + 256);
    cudaMalloc(&clk, sizeof(long long) * blocks);
    k_red<<<blocks, 256>>>(acc, rows, lanes, clk);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k_red<<<blocks, 256>>>(acc, rows, lanes, clk);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double reds = (double)blocks * 8 * 2048;
    printf("{\"pipe\": \"RED.ADD.F32 %d lanes/warp, %d rows\", \"warp_reds_per_us\": %.1f, \"ms\": %.3f}\n", lanes, rows, reds / (ms * 1e3), ms);
    cudaFree(acc); cudaFree(clk);
  }
  return 0;
}
