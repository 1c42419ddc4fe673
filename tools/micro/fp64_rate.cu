// DFMA / FFMA issue-rate probe (one GPU): prints achieved TFLOP/s of dependent-chain-free FMA loops.
#include <cstdio>
#include <cuda_runtime.h>
template <typename T>
__global__ void fma_loop(T* out, int iters, T a, T b)
{
  T x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; ++i) {
    x0 = x0 * a + b; x1 = x1 * a + b; x2 = x2 * a + b; x3 = x3 * a + b;
    x4 = x4 * a + b; x5 = x5 * a + b; x6 = x6 * a + b; x7 = x7 * a + b;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}
template <typename T>
void run(const char* name)
{
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int blocks = sms * 8, threads = 256, iters = 20000;
  T* out; cudaMalloc(&out, sizeof(T) * blocks * threads);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  fma_loop<T><<<blocks, threads>>>(out, 100, (T)1.0000001, (T)1e-9);
  cudaEventRecord(e0);
  fma_loop<T><<<blocks, threads>>>(out, iters, (T)1.0000001, (T)1e-9);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double flop = 2.0 * 8 * (double)iters * blocks * threads;
  printf("%s: %.3f ms, %.2f TFLOP/s (%d SMs)\n", name, ms, flop / (ms * 1e-3) / 1e12, sms);
  cudaFree(out);
}
int main() { run<float>("FFMA fp32"); run<double>("DFMA fp64"); return 0; }
