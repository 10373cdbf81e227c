// Read-only HBM stream probe: what a kernel that only READS can sustain on this GPU
// (the roofline denominator in MEASURED_PEAKS.json is a read+write copy).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/hbm_probe tools/hbm_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
__global__ void __launch_bounds__(512) read_sum(const double2* __restrict__ x, size_t n2, double* out) {
  double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, st = (size_t)gridDim.x * blockDim.x;
  for (; i + 3 * st < n2; i += 4 * st) {
    double2 a = x[i], b = x[i + st], c = x[i + 2 * st], d = x[i + 3 * st];
    s0 += a.x + a.y; s1 += b.x + b.y; s2 += c.x + c.y; s3 += d.x + d.y;
  }
  for (; i < n2; i += st) { double2 a = x[i]; s0 += a.x + a.y; }
  double s = (s0 + s1) + (s2 + s3);
  if (s == 1.2345e300) out[0] = s;
}
int main(int argc, char** argv) {
  size_t gb = argc > 1 ? atol(argv[1]) : 32;
  size_t bytes = gb << 30, n2 = bytes / 16;
  double2* x; double* out;
  cudaMalloc(&x, bytes); cudaMalloc(&out, 8); cudaMemset(x, 0, bytes);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int ctas = 2; ctas <= 8; ctas *= 2) {
    int grid = 148 * ctas;
    for (int rep = 0; rep < 2; ++rep) {
      int iters = rep == 0 ? 3 : 60;      // burst, then ~0.3 s sustained
      read_sum<<<grid, 512>>>(x, n2, out);
      cudaDeviceSynchronize();
      cudaEventRecord(e0);
      for (int i = 0; i < iters; ++i) read_sum<<<grid, 512>>>(x, n2, out);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      printf("read-only stream: %zu GiB, %d CTAs/SM x 512 thr, %d launches: %.1f GB/s\n", gb, ctas, iters,
             (double)bytes * iters / (ms * 1e-3) / 1e9);
    }
  }
  return 0;
}
