// DMMA (mma.sync m8n8k4 f64) versus DFMA throughput on one GPU: decides whether the element contraction
// K = sum_q T^T D T runs on the FP64 tensor path.   build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/dmma_microbench tools/dmma_microbench.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int CH>
__global__ void dmma_kernel(double* out, int iters) {
  double c[CH][2];
  for (int k = 0; k < CH; ++k) { c[k][0] = threadIdx.x * 1e-9 + k; c[k][1] = k; }
  const double a = 1.0 + 1e-9 * threadIdx.x, b = 1e-3;
  for (int i = 0; i < iters; ++i)
#pragma unroll
    for (int k = 0; k < CH; ++k) dmma(c[k][0], c[k][1], a, b);
  double s = 0; for (int k = 0; k < CH; ++k) s += c[k][0] + c[k][1];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int CH>
__global__ void dfma_kernel(double* out, int iters) {
  double c[CH];
  for (int k = 0; k < CH; ++k) c[k] = threadIdx.x * 1e-9 + k;
  const double a = 1.0000001, b = 1e-9;
  for (int i = 0; i < iters; ++i)
#pragma unroll
    for (int k = 0; k < CH; ++k) c[k] = c[k] * a + b;
  double s = 0; for (int k = 0; k < CH; ++k) s += c[k];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// mixed: CH dmma chains + CH dfma chains interleaved (do the two share the pipe?)
template <int CH>
__global__ void mixed_kernel(double* out, int iters) {
  double c[CH][2], f[CH];
  for (int k = 0; k < CH; ++k) { c[k][0] = threadIdx.x * 1e-9 + k; c[k][1] = k; f[k] = k; }
  const double a = 1.0 + 1e-9 * threadIdx.x, b = 1e-3;
  for (int i = 0; i < iters; ++i)
#pragma unroll
    for (int k = 0; k < CH; ++k) { dmma(c[k][0], c[k][1], a, b); f[k] = f[k] * a + b; }
  double s = 0; for (int k = 0; k < CH; ++k) s += c[k][0] + c[k][1] + f[k];
  out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <class F> float time_ms(F f) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main() {
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* out; cudaMalloc(&out, sizeof(double) * sms * 8 * 1024);
  const int iters = 20000;
  for (int warps : {1, 2, 4, 8, 16}) {
    const int threads = 32 * (warps < 4 ? warps * 4 : 16), blocks = sms * (warps < 4 ? 1 : warps / 4);   // warps per SMSP
    float m1 = time_ms([&] { dmma_kernel<8><<<blocks, threads>>>(out, iters); });
    float m2 = time_ms([&] { dfma_kernel<8><<<blocks, threads>>>(out, iters); });
    float m3 = time_ms([&] { mixed_kernel<8><<<blocks, threads>>>(out, iters); });
    float m4 = time_ms([&] { dmma_kernel<1><<<blocks, threads>>>(out, iters); });
    const double nw = (double)blocks * threads / 32;
    printf("warps/SMSP %2d: DMMA x8 chains %.2f TFLOP/s | DFMA x8 %.2f TFLOP/s | mixed %.2f TFLOP/s (dmma %.2f + dfma %.2f) | DMMA 1 chain %.2f TFLOP/s (latency %.1f clk @1.965GHz)\n", warps,
           nw * iters * 8 * 512.0 / (m1 * 1e9), nw * iters * 8 * 64.0 / (m2 * 1e9), nw * iters * 8 * 576.0 / (m3 * 1e9),
           nw * iters * 8 * 512.0 / (m3 * 1e9), nw * iters * 8 * 64.0 / (m3 * 1e9), nw * iters * 512.0 / (m4 * 1e9), m4 * 1e-3 * 1.965e9 / iters);
  }
  return 0;
}
