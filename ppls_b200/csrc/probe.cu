// Measured machine ceilings for the roofline lines of bench.py (include/ppls_b200.h: ppls_probe): what a kernel
// that only READS can stream from HBM on this GPU (MEASURED_PEAKS.json's figure is a read+write copy), and what
// the FP64 tensor path (mma.sync m8n8k4 f64 -> DMMA, the instruction of em_batch.cu / em_var.cu) sustains from
// registers.  Plain probes: no product arithmetic depends on them.
#include "common.cuh"

namespace ppls {

__global__ void __launch_bounds__(512) probe_read_kernel(const double2* __restrict__ x, size_t n2, double* out) {
  double s0 = 0, s1 = 0, s2 = 0, s3 = 0;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t st = (size_t)gridDim.x * blockDim.x;
  for (; i + 3 * st < n2; i += 4 * st) {
    const double2 a = x[i], b = x[i + st], c = x[i + 2 * st], d = x[i + 3 * st];
    s0 += a.x + a.y; s1 += b.x + b.y; s2 += c.x + c.y; s3 += d.x + d.y;
  }
  for (; i < n2; i += st) { const double2 a = x[i]; s0 += a.x + a.y; }
  const double s = (s0 + s1) + (s2 + s3);
  if (s == 1.2345e300) out[0] = s;                 // keeps the loads alive
}

__global__ void __launch_bounds__(256) probe_dmma_kernel(double* out, int iters, double x) {
  double c[8][2];
#pragma unroll
  for (int i = 0; i < 8; ++i) { c[i][0] = threadIdx.x; c[i][1] = i; }
  const double a = x + threadIdx.x, b = x - threadIdx.x;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// GB/s of a read-only stream over `bytes` of device memory at `x` (launches repeated for ~`launches` passes)
cudaError_t probe_read(const void* x, size_t bytes, int sm_count, int launches, double* scratch, double* gbs, cudaStream_t stream) {
  cudaEvent_t e0, e1;
  cudaError_t e = cudaEventCreate(&e0);
  if (e != cudaSuccess) return e;
  e = cudaEventCreate(&e1);
  if (e != cudaSuccess) { cudaEventDestroy(e0); return e; }
  const size_t n2 = bytes / 16;
  const int grid = sm_count * 4;
  probe_read_kernel<<<grid, 512, 0, stream>>>((const double2*)x, n2, scratch);
  cudaEventRecord(e0, stream);
  for (int i = 0; i < launches; ++i) probe_read_kernel<<<grid, 512, 0, stream>>>((const double2*)x, n2, scratch);
  cudaEventRecord(e1, stream);
  e = cudaEventSynchronize(e1);
  float ms = 0.f;
  if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
  if (e == cudaSuccess) e = cudaGetLastError();
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (e == cudaSuccess) *gbs = (double)(n2 * 16) * launches / (ms * 1e-3) / 1e9;
  return e;
}

// TFLOP/s of back-to-back DMMA.8x8x4 from registers, 8 warps per SM
cudaError_t probe_dmma(int sm_count, double* scratch /* >= sm_count*256 doubles */, double* tflops, cudaStream_t stream) {
  cudaEvent_t e0, e1;
  cudaError_t e = cudaEventCreate(&e0);
  if (e != cudaSuccess) return e;
  e = cudaEventCreate(&e1);
  if (e != cudaSuccess) { cudaEventDestroy(e0); return e; }
  const int iters = 20000;
  probe_dmma_kernel<<<sm_count, 256, 0, stream>>>(scratch, 200, 1e-9);
  cudaEventRecord(e0, stream);
  probe_dmma_kernel<<<sm_count, 256, 0, stream>>>(scratch, iters, 1e-9);
  cudaEventRecord(e1, stream);
  e = cudaEventSynchronize(e1);
  float ms = 0.f;
  if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
  if (e == cudaSuccess) e = cudaGetLastError();
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (e == cudaSuccess) *tflops = 2.0 * 8 * 8 * 4 * 8 * (double)iters * 8.0 * sm_count / (ms * 1e-3) / 1e12;
  return e;
}

}  // namespace ppls
