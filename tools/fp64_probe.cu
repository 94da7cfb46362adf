// GPU probe (timing only): latency of a dependent fp64 FMA chain and fp64 FMA throughput per SM on B200,
// beside the fp32 numbers.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/build/fp64_probe tools/fp64_probe.cu
#include <cuda_runtime.h>
#include <cstdio>

template <typename T, int ILP>
__global__ void chain(T* out, long long* cyc, int iters, T a, T b) {
  T x[ILP];
#pragma unroll
  for (int i = 0; i < ILP; ++i) x[i] = (T)(threadIdx.x + i);
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = x[i] * a + b;
  }
  const long long t1 = clock64();
  T s = 0;
#pragma unroll
  for (int i = 0; i < ILP; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <typename T, int ILP>
void run(const char* name, int threads) {
  T* out;
  long long* cyc;
  cudaMalloc(&out, 1 << 20);
  cudaMalloc(&cyc, 8);
  const int iters = 4096;
  chain<T, ILP><<<1, threads>>>(out, cyc, iters, (T)1.0000001, (T)1e-9);
  chain<T, ILP><<<1, threads>>>(out, cyc, iters, (T)1.0000001, (T)1e-9);
  long long h;
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  const double per = (double)h / iters;
  printf("%-6s ILP %2d threads %4d: %7.2f clk per round of %d FMA/thread -> %6.2f FMA/clk/SM\n", name, ILP, threads, per, ILP,
         (double)ILP * threads / per);
  cudaFree(out);
  cudaFree(cyc);
}

int main() {
  run<double, 1>("fp64", 32);
  run<double, 2>("fp64", 32);
  run<double, 4>("fp64", 32);
  run<double, 8>("fp64", 32);
  run<double, 1>("fp64", 128);
  run<double, 4>("fp64", 128);
  run<double, 8>("fp64", 256);
  run<double, 8>("fp64", 1024);
  run<float, 1>("fp32", 32);
  run<float, 8>("fp32", 128);
  run<float, 8>("fp32", 1024);
  return 0;
}
