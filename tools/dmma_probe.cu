// FP64 tensor-core (DMMA, mma.sync f64) throughput probe vs plain DFMA, registers only.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/dmma_probe tools/dmma_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma16816(double (&c)[4], const double (&a)[8], const double (&b)[4]) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                 "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}

template <int NACC>
__global__ void k884(double* out, int iters, double x) {
  double c[NACC][2];
  for (int i = 0; i < NACC; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; }
  double a = x + threadIdx.x, b = x - threadIdx.x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) dmma884(c[i], a, b);
  }
  double s = 0; for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NACC>
__global__ void k16816(double* out, int iters, double x) {
  double c[NACC][4];
  for (int i = 0; i < NACC; ++i) for (int j = 0; j < 4; ++j) c[i][j] = threadIdx.x + i + j;
  double a[8], b[4];
  for (int j = 0; j < 8; ++j) a[j] = x + threadIdx.x + j;
  for (int j = 0; j < 4; ++j) b[j] = x - threadIdx.x - j;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) dmma16816(c[i], a, b);
  }
  double s = 0; for (int i = 0; i < NACC; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NACC>
__global__ void kdfma(double* out, int iters, double x) {
  double c[NACC];
  for (int i = 0; i < NACC; ++i) c[i] = threadIdx.x + i;
  double a = x + threadIdx.x, b = x - threadIdx.x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) c[i] = fma(a, b, c[i]);
  }
  double s = 0; for (int i = 0; i < NACC; ++i) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// The register pattern of a real GEMM warp tile: MI x NJ accumulator tiles, MI distinct A and NJ distinct B fragments
// per k-step (no operand shared by all DMMAs as in k884).  LDS = 1: the fragments are re-read from shared memory
// every k-step (conflict-free pitch), as the GEMM kernels of em_batch.cu / em_var.cu do.
template <int MI, int NJ, int LDS>
__global__ void k884_grid(double* out, int iters, double x) {
  __shared__ double sa[4 * 132], sb[4 * 132];
  for (int e = threadIdx.x; e < 4 * 132; e += blockDim.x) { sa[e] = x + e; sb[e] = x - e; }
  __syncthreads();
  const int lane = threadIdx.x & 31, gid = lane >> 2, tig = lane & 3;
  double c[MI][NJ][2];
  for (int i = 0; i < MI; ++i) for (int j = 0; j < NJ; ++j) { c[i][j][0] = i; c[i][j][1] = j; }
  double a[MI], b[NJ];
  for (int i = 0; i < MI; ++i) a[i] = x + threadIdx.x + i;
  for (int j = 0; j < NJ; ++j) b[j] = x - threadIdx.x - j;
  const volatile double* pa = sa + tig * 132 + gid;
  const volatile double* pb = sb + tig * 132 + gid;
  for (int it = 0; it < iters; ++it) {
    if (LDS) {
#pragma unroll
      for (int i = 0; i < MI; ++i) a[i] = pa[i * 8];
#pragma unroll
      for (int j = 0; j < NJ; ++j) b[j] = pb[j * 8];
    }
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
      for (int j = 0; j < NJ; ++j) dmma884(c[i][j], a[i], b[j]);
  }
  double s = 0;
  for (int i = 0; i < MI; ++i) for (int j = 0; j < NJ; ++j) s += c[i][j][0] + c[i][j][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
static double timeit(F launch) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  launch(); cudaDeviceSynchronize();
  cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); return ms * 1e-3;
}

int main() {
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* out; cudaMalloc(&out, sizeof(double) * sms * 8 * 1024);
  const int iters = 20000;
  for (int warps : {4, 8, 16}) {
    const int threads = warps * 32, blocks = sms;
    double t;
    t = timeit([&] { k884<8><<<blocks, threads>>>(out, iters, 1e-9); });
    printf("warps/SM %2d  DMMA m8n8k4  x8 acc : %.2f TFLOP/s\n", warps, 2.0 * 8 * 8 * 4 * 8 * iters * (double)warps * blocks / t / 1e12);
    t = timeit([&] { k16816<4><<<blocks, threads>>>(out, iters, 1e-9); });
    printf("warps/SM %2d  DMMA m16n8k16 x4 acc : %.2f TFLOP/s\n", warps, 2.0 * 16 * 8 * 16 * 4 * iters * (double)warps * blocks / t / 1e12);
    t = timeit([&] { k16816<8><<<blocks, threads>>>(out, iters, 1e-9); });
    printf("warps/SM %2d  DMMA m16n8k16 x8 acc : %.2f TFLOP/s\n", warps, 2.0 * 16 * 8 * 16 * 8 * iters * (double)warps * blocks / t / 1e12);
    t = timeit([&] { kdfma<32><<<blocks, threads>>>(out, iters, 1e-9); });
    printf("warps/SM %2d  DFMA x32 acc         : %.2f TFLOP/s\n", warps, 2.0 * 32 * 32 * iters * (double)warps * blocks / t / 1e12);
  }
  for (int warps : {4, 8}) {
    const int threads = warps * 32, blocks = sms, it2 = 4000;
    auto tf = [&](double t, int mi, int nj) { return 2.0 * 8 * 8 * 4 * mi * nj * it2 * (double)warps * blocks / t / 1e12; };
    double t;
    t = timeit([&] { k884_grid<8, 4, 0><<<blocks, threads>>>(out, it2, 1e-9); });
    printf("warps/SM %2d  DMMA m8n8k4 8x4 tile grid, registers : %.2f TFLOP/s\n", warps, tf(t, 8, 4));
    t = timeit([&] { k884_grid<8, 4, 1><<<blocks, threads>>>(out, it2, 1e-9); });
    printf("warps/SM %2d  DMMA m8n8k4 8x4 tile grid, LDS frags : %.2f TFLOP/s\n", warps, tf(t, 8, 4));
    t = timeit([&] { k884_grid<4, 4, 0><<<blocks, threads>>>(out, it2, 1e-9); });
    printf("warps/SM %2d  DMMA m8n8k4 4x4 tile grid, registers : %.2f TFLOP/s\n", warps, tf(t, 4, 4));
    t = timeit([&] { k884_grid<4, 4, 1><<<blocks, threads>>>(out, it2, 1e-9); });
    printf("warps/SM %2d  DMMA m8n8k4 4x4 tile grid, LDS frags : %.2f TFLOP/s\n", warps, tf(t, 4, 4));
    t = timeit([&] { k884_grid<4, 8, 1><<<blocks, threads>>>(out, it2, 1e-9); });
    printf("warps/SM %2d  DMMA m8n8k4 4x8 tile grid, LDS frags : %.2f TFLOP/s\n", warps, tf(t, 4, 8));
  }
  cudaFree(out);
  return 0;
}
