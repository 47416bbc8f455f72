// Micro-benchmark: throughput of RED.E.ADD.F64 (atomicAdd(double) without return) on B200 for the access shapes a
// CRS scatter can produce.  Used to decide the scatter layout of the assembly kernels (DESIGN.md "Scatter").
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/red_microbench tools/red_microbench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %s\n", cudaGetErrorString(e), #x); return 1; } } while (0)

// each warp instruction: lanes grouped in runs of G consecutive doubles; runs are `stride` doubles apart
template <int G, bool STORE>
__global__ void red_pattern(double* buf, int64_t ndoubles, int stride, int reps) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t run = t / G; const int in = (int)(t % G);
  for (int r = 0; r < reps; ++r) {
    const int64_t a = ((run * stride + (int64_t)r * 4 * G) + in) % ndoubles;
    if (STORE) buf[a] = 1.0; else atomicAdd(buf + a, 1.0);
  }
}

template <int G, bool STORE>
double run(double* buf, int64_t nd, int stride, int64_t threads, int reps) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  red_pattern<G, STORE><<<(unsigned)((threads + 255) / 256), 256>>>(buf, nd, stride, 1);
  float best = 1e30f;
  for (int k = 0; k < 3; ++k) {
    cudaEventRecord(e0); red_pattern<G, STORE><<<(unsigned)((threads + 255) / 256), 256>>>(buf, nd, stride, reps); cudaEventRecord(e1);
    cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  return (double)threads * reps / (best * 1e-3) / 1e9;
}

int main() {
  const int64_t nd = 454756609LL;           // doubles: the CRS value array of the 256^3 Q1 case (3.6 GB)
  double* buf; CK(cudaMalloc(&buf, nd * sizeof(double))); CK(cudaMemset(buf, 0, nd * sizeof(double)));
  const int64_t threads = 16LL * 1024 * 1024; const int reps = 16;
  printf("pattern, G lanes per run, run stride (doubles), Gops/s\n");
  printf("RED  G=1  stride 27   %8.1f\n", run<1, false>(buf, nd, 27, threads, reps));
  printf("RED  G=2  stride 27   %8.1f\n", run<2, false>(buf, nd, 27, threads, reps));
  printf("RED  G=4  stride 28   %8.1f\n", run<4, false>(buf, nd, 28, threads, reps));
  printf("RED  G=8  stride 32   %8.1f\n", run<8, false>(buf, nd, 32, threads, reps));
  printf("RED  G=32 stride 32   %8.1f   (fully coalesced)\n", run<32, false>(buf, nd, 32, threads, reps));
  printf("RED  G=1  stride 4    %8.1f   (one lane per 32B sector, dense)\n", run<1, false>(buf, nd, 4, threads, reps));
  printf("ST   G=1  stride 27   %8.1f\n", run<1, true>(buf, nd, 27, threads, reps));
  printf("ST   G=32 stride 32   %8.1f\n", run<32, true>(buf, nd, 32, threads, reps));
  // working set small enough to stay in L2 (64 MB)
  const int64_t nd2 = 8LL * 1024 * 1024;
  printf("RED  G=1  stride 27  L2-resident  %8.1f\n", run<1, false>(buf, nd2, 27, threads, reps));
  printf("RED  G=4  stride 28  L2-resident  %8.1f\n", run<4, false>(buf, nd2, 28, threads, reps));
  printf("RED  G=32 stride 32  L2-resident  %8.1f\n", run<32, false>(buf, nd2, 32, threads, reps));
  return 0;
}
