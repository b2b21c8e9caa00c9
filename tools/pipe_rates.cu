// pipe_rates.cu — micro-benchmark of the SM pipes the hot path is bound by ("micro-benchmark it"):
// issue rate in thread-instructions per clock per SM of POPC (XU pipe: the matcher's bound), REDUX.MIN (the
// cross-check column minimum), LOP3 / VIMNMX.U16x2 (ALU pipe: matcher adders, FAST pre-test), IMAD (FMA pipe: key
// arithmetic), DFMA / DADD+DMUL (FP64 pipe: five-point RANSAC).  One CTA of 1024 threads per SM, 8 independent
// dependency chains per thread, cycles from clock64() inside the CTA, so the result does not depend on the clock.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/build/pipe_rates tools/pipe_rates.cu
//   tools/build/pipe_rates > gpurun_out/pipe_rates.json
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kIters = 4096, kChains = 8, kThreads = 1024;

template <int OP>
__global__ void __launch_bounds__(kThreads) k_rate(uint32_t seed, unsigned long long* cycles, uint32_t* sink) {
  uint32_t x[kChains];
  double d[kChains];
#pragma unroll
  for (int i = 0; i < kChains; ++i) { x[i] = seed * (threadIdx.x + 1) + i * 0x9e3779b9u; d[i] = 1.0 + 1e-9 * (x[i] & 1023); }
  const uint32_t a = seed | 1u, b = seed ^ 0x5bd1e995u;
  const double da = 1.0 + 1e-12 * seed, db = 1e-13 * seed;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < kIters; ++it) {
#pragma unroll
    for (int i = 0; i < kChains; ++i) {
      if (OP == 0) x[i] = __popc(x[i]) + a;                                   // POPC (+ IADD on the ALU pipe)
      if (OP == 1) x[i] = __reduce_min_sync(0xffffffffu, x[i]) + a;           // REDUX.MIN
      if (OP == 2) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(a), "r"(b));   // LOP3
      if (OP == 3) x[i] = __vminu2(x[i] + a, b);                              // VIMNMX.U16x2 (+ IADD)
      if (OP == 4) x[i] = x[i] * a + b;                                       // IMAD
      if (OP == 5) d[i] = fma(d[i], da, db);                                  // DFMA
      if (OP == 6) d[i] = __dadd_rn(__dmul_rn(d[i], da), db);                 // DMUL + DADD
    }
  }
  const long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < kChains; ++i) s += x[i] + (uint32_t)d[i];
  if (s == 0x12345678u) sink[0] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(t1 - t0);
}

template <int OP>
static double run(int n_sm, unsigned long long* d_cyc, uint32_t* d_sink) {
  k_rate<OP><<<n_sm, kThreads>>>(12345u, d_cyc, d_sink);          // warm-up
  k_rate<OP><<<n_sm, kThreads>>>(12345u, d_cyc, d_sink);
  cudaDeviceSynchronize();
  unsigned long long h[256];
  cudaMemcpy(h, d_cyc, n_sm * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
  double worst = 0;
  for (int i = 0; i < n_sm; ++i) worst = h[i] > worst ? (double)h[i] : worst;
  return (double)kThreads * kIters * kChains / worst;             // thread-instructions per clock per SM
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int n_sm = p.multiProcessorCount;
  unsigned long long* d_cyc; uint32_t* d_sink;
  cudaMalloc(&d_cyc, 256 * sizeof(unsigned long long));
  cudaMalloc(&d_sink, 16);
  const double popc = run<0>(n_sm, d_cyc, d_sink), redux = run<1>(n_sm, d_cyc, d_sink), lop3 = run<2>(n_sm, d_cyc, d_sink);
  const double vmn = run<3>(n_sm, d_cyc, d_sink), imad = run<4>(n_sm, d_cyc, d_sink), dfma = run<5>(n_sm, d_cyc, d_sink);
  const double dmuladd = run<6>(n_sm, d_cyc, d_sink);
  printf("{\"device\": \"%s\", \"sms\": %d, \"unit\": \"thread-instructions per clock per SM\", \"popc\": %.2f, "
         "\"redux_min\": %.2f, \"lop3\": %.2f, \"vimnmx_u16x2_plus_iadd_pairs\": %.2f, \"imad\": %.2f, \"dfma\": %.2f, "
         "\"dmul_plus_dadd_pairs\": %.2f, \"err\": \"%s\"}\n", p.name, n_sm, popc, redux, lop3, vmn, imad, dfma, dmuladd,
         cudaGetErrorString(cudaGetLastError()));
  return 0;
}
