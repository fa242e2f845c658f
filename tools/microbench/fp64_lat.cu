// FP64 pipe microbenchmark: dependent-issue latency and throughput of DFMA / DADD / DMUL on sm_100a as a function of
// ILP (independent chains per thread) and resident warps per SM.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cuda_runtime.h>
template <int ILP, int OP>
__global__ void k(double* out, double a, double b, int iters) {
  double x[ILP];
  for (int i = 0; i < ILP; ++i) x[i] = a + i + threadIdx.x;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      if (OP == 0) x[i] = fma(x[i], a, b);
      else if (OP == 1) x[i] = __dadd_rn(x[i], b);
      else if (OP == 2) x[i] = __dmul_rn(x[i], a);
      else x[i] = x[i]/a;          // IEEE divide: dependent chain of MUFU + DFMAs
    }
  }
  long long t1 = clock64();
  double s = 0; for (int i = 0; i < ILP; ++i) s += x[i];
  out[blockIdx.x*blockDim.x + threadIdx.x] = s + (double)(t1 - t0)*1e-300;
  if (threadIdx.x == 0 && blockIdx.x == 0) out[gridDim.x*blockDim.x] = (double)(t1 - t0);
}
template <int ILP, int OP> void run(const char* name, int warpsPerSM, double* d) {
  const int iters = 4096;
  int threads = 32*(warpsPerSM < 4 ? warpsPerSM : 4); int blocksPerSM = (warpsPerSM + 3)/4;
  int blocks = 148*blocksPerSM;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<ILP, OP><<<blocks, threads>>>(d, 1.0000001, 1e-9, iters);
  cudaEventRecord(e0);
  k<ILP, OP><<<blocks, threads>>>(d, 1.0000001, 1e-9, iters);
  cudaEventRecord(e1); cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double cyc; cudaMemcpy(&cyc, d + blocks*threads, 8, cudaMemcpyDeviceToHost);
  double ops = (double)blocks*threads/32*iters*ILP;   // warp instructions
  printf("%-5s ILP %2d warps/SM %2d : %7.2f cyc per op per warp-chain step, %8.1f G warp-instr/s (%5.1f%% of 581.6)\n",
         name, ILP, blocksPerSM*threads/32, cyc/iters, ops/ms/1e6, ops/ms/1e6/581.6*100);
}
int main() {
  double* d; cudaMalloc(&d, sizeof(double)*(148*16*128 + 8));
  run<1, 0>("DFMA", 1, d); run<1, 1>("DADD", 1, d); run<1, 2>("DMUL", 1, d); run<1, 3>("DDIV", 1, d);
  run<2, 0>("DFMA", 4, d); run<4, 0>("DFMA", 4, d); run<8, 0>("DFMA", 4, d);
  run<1, 0>("DFMA", 8, d); run<1, 0>("DFMA", 16, d); run<1, 0>("DFMA", 32, d); run<1, 0>("DFMA", 64, d);
  run<2, 0>("DFMA", 8, d); run<2, 0>("DFMA", 16, d); run<4, 0>("DFMA", 8, d); run<4, 0>("DFMA", 16, d); run<4, 0>("DFMA", 32, d);
  run<8, 0>("DFMA", 8, d); run<8, 0>("DFMA", 16, d);
  // no-FMA issue peak (the kernels are built with -fmad=false): DADD / DMUL with 8 independent chains, 16 and 32 warps per SM
  run<8, 1>("DADD", 16, d); run<8, 1>("DADD", 32, d); run<8, 2>("DMUL", 16, d); run<8, 2>("DMUL", 32, d);
  run<4, 1>("DADD", 8, d); run<4, 2>("DMUL", 8, d); run<2, 1>("DADD", 8, d); run<1, 1>("DADD", 8, d);
  run<1, 3>("DDIV", 4, d); run<1, 3>("DDIV", 8, d); run<1, 3>("DDIV", 16, d); run<4, 3>("DDIV", 8, d); run<4, 3>("DDIV", 16, d);
  return 0;
}
