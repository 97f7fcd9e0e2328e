// Microbenchmark: FP64 mma.sync throughput on sm_100a next to the DFMA vector pipe.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o dmma dmma.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void mma884(double& d, double a, double b, double c) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%4,%5};"
               : "+d"(d), "+d"(c) : "d"(a), "d"(b), "d"(d), "d"(c));
}
__device__ __forceinline__ void mma1688(double (&d)[4], const double (&a)[4], const double (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+d"(d[0]), "+d"(d[1]), "+d"(d[2]), "+d"(d[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
__device__ __forceinline__ void mma1684(double (&d)[4], const double (&a)[2], double b) {
  asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
               : "+d"(d[0]), "+d"(d[1]), "+d"(d[2]), "+d"(d[3]) : "d"(a[0]), "d"(a[1]), "d"(b));
}

// ILP independent accumulators per warp, iters rounds
template <int ILP>
__global__ void k_1688(double* out, int iters) {
  double d[ILP][4], a[4], b[2];
  for (int i = 0; i < 4; ++i) a[i] = 1.0 + threadIdx.x * 1e-9 + i;
  b[0] = 1e-3 * threadIdx.x; b[1] = 2e-3;
  for (int j = 0; j < ILP; ++j) for (int i = 0; i < 4; ++i) d[j][i] = j + i;
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int j = 0; j < ILP; ++j) mma1688(d[j], a, b);
  double s = 0;
  for (int j = 0; j < ILP; ++j) for (int i = 0; i < 4; ++i) s += d[j][i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int ILP>
__global__ void k_1684(double* out, int iters) {
  double d[ILP][4], a[2], b;
  for (int i = 0; i < 2; ++i) a[i] = 1.0 + threadIdx.x * 1e-9 + i;
  b = 1e-3 * threadIdx.x;
  for (int j = 0; j < ILP; ++j) for (int i = 0; i < 4; ++i) d[j][i] = j + i;
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int j = 0; j < ILP; ++j) mma1684(d[j], a, b);
  double s = 0;
  for (int j = 0; j < ILP; ++j) for (int i = 0; i < 4; ++i) s += d[j][i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// pruning-shaped chain: D = A*B; A' = D * sibling (elementwise), next level uses A'
template <int ILP>
__global__ void k_chain(double* out, int iters) {
  double a[ILP][4], b[2], sib[4];
  for (int j = 0; j < ILP; ++j) for (int i = 0; i < 4; ++i) a[j][i] = 0.25 + 1e-3 * (threadIdx.x & 7) + 1e-4 * i;
  b[0] = 0.24 + 1e-3 * (threadIdx.x & 3); b[1] = 0.26;
  for (int i = 0; i < 4; ++i) sib[i] = 1.0 + 1e-6 * i;
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int j = 0; j < ILP; ++j) {
      double d[4] = {0, 0, 0, 0};
      mma1688(d, a[j], b);
      for (int i = 0; i < 4; ++i) a[j][i] = d[i] * sib[i];
    }
  double s = 0;
  for (int j = 0; j < ILP; ++j) for (int i = 0; i < 4; ++i) s += a[j][i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int ILP>
__global__ void k_dfma(double* out, int iters) {
  double d[ILP], a = 1.0 + 1e-9 * threadIdx.x, b = 1e-3;
  for (int j = 0; j < ILP; ++j) d[j] = j;
  for (int it = 0; it < iters; ++it)
#pragma unroll
    for (int j = 0; j < ILP; ++j) d[j] = fma(d[j], a, b);
  double s = 0;
  for (int j = 0; j < ILP; ++j) s += d[j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
float timeit(F f) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  return ms;
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount;
  double* out; cudaMalloc(&out, sizeof(double) * sms * 64 * 1024);
  const int iters = 20000;
  printf("device %s, %d SMs\n", p.name, sms);
  for (int wps : {4, 8, 12, 16, 32}) {     // warps per SM (one block per SM... blocks of wps warps)
    const int thr = wps * 32;
    auto rep = [&](const char* name, float ms, double fma_per_warp_iter) {
      double tf = 2.0 * fma_per_warp_iter * iters * wps * sms / (ms * 1e-3) / 1e12;
      printf("%-22s warps/SM %2d  %8.3f ms  %7.2f TFLOP/s\n", name, wps, ms, tf);
    };
    rep("m16n8k8 ilp4", timeit([&] { k_1688<4><<<sms, thr>>>(out, iters); }), 4 * 1024.0);
    rep("m16n8k8 ilp8", timeit([&] { k_1688<8><<<sms, thr>>>(out, iters); }), 8 * 1024.0);
    rep("m16n8k4 ilp8", timeit([&] { k_1684<8><<<sms, thr>>>(out, iters); }), 8 * 512.0);
    rep("chain m16n8k8+4dmul i8", timeit([&] { k_chain<8><<<sms, thr>>>(out, iters); }), 8 * 1024.0);
    rep("dfma ilp8", timeit([&] { k_dfma<8><<<sms, thr>>>(out, iters); }), 8 * 32.0);
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("status %s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
