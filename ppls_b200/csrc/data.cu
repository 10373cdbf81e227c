// Data-side kernels: re-tiling R's column-major matrices into the engine's row-major shard,
// ||X||_F^2 / ||Y||_F^2 (loop-invariant in the reference, yet recomputed on every call at
// Package/PPLS/src/loglC.cpp:334), and the device-side synthetic generator (model of simulC,
// Package/PPLS/src/loglC.cpp:288-311, orthogonal parts zero, r components).
// Compiled with -fmad=false so the generator produces the same bits as the host restatement.
#include "common.cuh"

namespace ppls {

// ---- column-major (R) -> row-major shard ---------------------------------------------------
// src: nrows x ncols column-major with leading dimension lds; dst rows row0.., columns col0..
__global__ void retile_kernel(const double* __restrict__ src, int64_t lds, int64_t nrows, int ncols,
                              double* __restrict__ D, const DataShape s, int col0, int64_t row0) {
  __shared__ double tile[32][33];
  const int64_t rb = (int64_t)blockIdx.x * 32;
  const int cb = blockIdx.y * 32;
  for (int c = threadIdx.y; c < 32; c += 8) {   // coalesced along rows of the column-major source
    const int64_t i = rb + threadIdx.x;
    const int j = cb + c;
    tile[c][threadIdx.x] = (i < nrows && j < ncols) ? src[i + (int64_t)j * lds] : 0.0;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += 8) {   // coalesced along columns of the row-major shard
    const int64_t i = rb + r;
    const int j = cb + threadIdx.x;
    if (i < nrows && j < ncols) D[data_index(s, row0 + i, col0 + j)] = tile[threadIdx.x][r];
  }
}

__global__ void untile_kernel(const double* __restrict__ D, const DataShape s, int col0, int64_t row0, int64_t nrows,
                              int ncols, double* __restrict__ dst, int64_t ldd) {
  __shared__ double tile[32][33];
  const int64_t rb = (int64_t)blockIdx.x * 32;
  const int cb = blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += 8) {
    const int64_t i = rb + r;
    const int j = cb + threadIdx.x;
    tile[r][threadIdx.x] = (i < nrows && j < ncols) ? D[data_index(s, row0 + i, col0 + j)] : 0.0;
  }
  __syncthreads();
  for (int c = threadIdx.y; c < 32; c += 8) {
    const int64_t i = rb + threadIdx.x;
    const int j = cb + c;
    if (i < nrows && j < ncols) dst[i + (int64_t)j * ldd] = tile[threadIdx.x][c];
  }
}

cudaError_t launch_retile(const double* colmajor, int64_t ldc, int64_t nrows, int ncols, double* D, const DataShape& s,
                          int col0, int64_t row0, cudaStream_t stream) {
  if (nrows <= 0 || ncols <= 0) return cudaSuccess;
  const int64_t gx = (nrows + 31) / 32, gy = (ncols + 31) / 32;
  if (gy > 65535) return cudaErrorInvalidValue;
  dim3 grid((unsigned)gx, (unsigned)gy), block(32, 8);
  retile_kernel<<<grid, block, 0, stream>>>(colmajor, ldc, nrows, ncols, D, s, col0, row0);
  return cudaGetLastError();
}

cudaError_t launch_untile(const double* D, const DataShape& s, int col0, int64_t row0, int64_t nrows, int ncols,
                          double* colmajor, int64_t ldc, cudaStream_t stream) {
  if (nrows <= 0 || ncols <= 0) return cudaSuccess;
  const int64_t gx = (nrows + 31) / 32, gy = (ncols + 31) / 32;
  if (gy > 65535) return cudaErrorInvalidValue;
  dim3 grid((unsigned)gx, (unsigned)gy), block(32, 8);
  untile_kernel<<<grid, block, 0, stream>>>(D, s, col0, row0, nrows, ncols, colmajor, ldc);
  return cudaGetLastError();
}

// pad columns (odd p or q) must be zero: the kernels read them
__global__ void zero_pad_kernel(double* D, int64_t n, int ld, int p, int p2, int q, int q2) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (p2 != p) D[(size_t)i * ld + p] = 0.0;
  if (q2 != q) D[(size_t)i * ld + p2 + q] = 0.0;
}
cudaError_t launch_zero_pad(double* D, const DataShape& s, cudaStream_t stream) {
  if (s.panel || (s.p2 == s.p && s.q2 == s.q)) return cudaSuccess;   // panel-major storage is zero-filled up front
  zero_pad_kernel<<<(unsigned)((s.n + 255) / 256), 256, 0, stream>>>(D, s.n, s.ld, s.p, s.p2, s.q, s.q2);
  return cudaGetLastError();
}

// ---- ||X||_F^2, ||Y||_F^2 -------------------------------------------------------------------
__global__ void ssq_partial_kernel(const double* __restrict__ D, const DataShape s,
                                   double* __restrict__ scratch /*[grid][2]*/) {
  const int64_t n = s.n; const int ld = s.ld, p2 = s.p2;
  double sx = 0.0, sy = 0.0;
  if (!s.panel) {
    // row-major: 16-byte loads (p2 is even, so a pair never straddles the X|Y boundary), 4 rows in flight per thread
    const int np = ld >> 1, px = p2 >> 1;
    const double2* D2 = reinterpret_cast<const double2*>(D);
    for (int64_t i = (int64_t)blockIdx.x * 4; i < n; i += (int64_t)gridDim.x * 4) {
      const int rows = (n - i < 4) ? (int)(n - i) : 4;
      for (int j = threadIdx.x; j < np; j += blockDim.x) {
        double2 v[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) v[r] = (r < rows) ? D2[(size_t)(i + r) * np + j] : make_double2(0.0, 0.0);
        double t = 0.0;
#pragma unroll
        for (int r = 0; r < 4; ++r) t += v[r].x * v[r].x + v[r].y * v[r].y;
        if (j < px) sx += t; else sy += t;
      }
    }
  } else {
    for (int64_t i = blockIdx.x; i < n; i += gridDim.x) {
      for (int j = threadIdx.x; j < ld; j += blockDim.x) {
        const double v = D[data_index(s, i, j)];
        if (j < p2) sx += v * v; else sy += v * v;
      }
    }
  }
  __shared__ double shx[8], shy[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sx += __shfl_xor_sync(0xffffffffu, sx, o);
    sy += __shfl_xor_sync(0xffffffffu, sy, o);
  }
  if ((threadIdx.x & 31) == 0) { shx[threadIdx.x >> 5] = sx; shy[threadIdx.x >> 5] = sy; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double ax = 0.0, ay = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { ax += shx[w]; ay += shy[w]; }
    scratch[2 * blockIdx.x] = ax; scratch[2 * blockIdx.x + 1] = ay;
  }
}
__global__ void ssq_final_kernel(const double* scratch, int nblocks, double* out2) {
  if (threadIdx.x < 2) {
    double s = 0.0;
    for (int b = 0; b < nblocks; ++b) s += scratch[2 * b + threadIdx.x];
    out2[threadIdx.x] = s;
  }
}
cudaError_t launch_ssq(const double* D, const DataShape& s, double* out2, double* scratch, int nblocks,
                       cudaStream_t stream) {
  ssq_partial_kernel<<<nblocks, 256, 0, stream>>>(D, s, scratch);
  ssq_final_kernel<<<1, 32, 0, stream>>>(scratch, nblocks, out2);
  return cudaGetLastError();
}

// ---- synthetic generator ----------------------------------------------------------------------
// Philox4x32-10 + AS241 inverse normal + an exactly reproducible log (only + - * / sqrt):
// same bits on the device as on any IEEE-754 host.
__device__ __forceinline__ void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int round = 0; round < 10; ++round) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
    const uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}

__device__ __forceinline__ double det_log(double x) {
  unsigned long long bits = (unsigned long long)__double_as_longlong(x);
  int e = (int)((bits >> 52) & 0x7FF) - 1022;
  bits = (bits & 0x000FFFFFFFFFFFFFull) | 0x3FE0000000000000ull;
  double m = __longlong_as_double((long long)bits);
  if (m < 0.70710678118654752440) { m = m * 2.0; e -= 1; }
  const double s = (m - 1.0) / (m + 1.0);
  const double s2 = s * s;
  double P = 1.0 / 27.0;
  P = P * s2 + 1.0 / 25.0; P = P * s2 + 1.0 / 23.0; P = P * s2 + 1.0 / 21.0;
  P = P * s2 + 1.0 / 19.0; P = P * s2 + 1.0 / 17.0; P = P * s2 + 1.0 / 15.0;
  P = P * s2 + 1.0 / 13.0; P = P * s2 + 1.0 / 11.0; P = P * s2 + 1.0 / 9.0;
  P = P * s2 + 1.0 / 7.0;  P = P * s2 + 1.0 / 5.0;  P = P * s2 + 1.0 / 3.0;
  P = P * s2 + 1.0;
  return 2.0 * s * P + (double)e * 0.69314718055994530942;
}

__device__ double det_ndtri(double u) {
  const double q = u - 0.5;
  double r, num, den;
  if (fabs(q) <= 0.425) {
    r = 0.180625 - q * q;
    num = 2.5090809287301226727e+3;
    num = num * r + 3.3430575583588128105e+4; num = num * r + 6.7265770927008700853e+4;
    num = num * r + 4.5921953931549871457e+4; num = num * r + 1.3731693765509461125e+4;
    num = num * r + 1.9715909503065514427e+3; num = num * r + 1.3314166789178437745e+2;
    num = num * r + 3.3871328727963666080e0;
    den = 5.2264952788528545610e+3;
    den = den * r + 2.8729085735721942674e+4; den = den * r + 3.9307895800092710610e+4;
    den = den * r + 2.1213794301586595867e+4; den = den * r + 5.3941960214247511077e+3;
    den = den * r + 6.8718700749205790830e+2; den = den * r + 4.2313330701600911252e+1;
    den = den * r + 1.0;
    return q * num / den;
  }
  r = (q < 0.0) ? u : 1.0 - u;
  r = sqrt(-det_log(r));
  if (r <= 5.0) {
    r = r - 1.6;
    num = 7.74545014278341407640e-4;
    num = num * r + 2.27238449892691845833e-2; num = num * r + 2.41780725177450611770e-1;
    num = num * r + 1.27045825245236838258e0;  num = num * r + 3.64784832476320460504e0;
    num = num * r + 5.76949722146069140550e0;  num = num * r + 4.63033784615654529590e0;
    num = num * r + 1.42343711074968357734e0;
    den = 1.05075007164441684324e-9;
    den = den * r + 5.47593808499534494600e-4; den = den * r + 1.51986665636164571966e-2;
    den = den * r + 1.48103976427480074590e-1; den = den * r + 6.89767334985100004550e-1;
    den = den * r + 1.67638483018380384940e0;  den = den * r + 2.05319162663775882187e0;
    den = den * r + 1.0;
  } else {
    r = r - 5.0;
    num = 2.01033439929228813265e-7;
    num = num * r + 2.71155556874348757815e-5; num = num * r + 1.24266094738807843860e-3;
    num = num * r + 2.65321895265761230930e-2; num = num * r + 2.96560571828504891230e-1;
    num = num * r + 1.78482653991729133580e0;  num = num * r + 5.46378491116411436990e0;
    num = num * r + 6.65790464350110377720e0;
    den = 2.04426310338993978564e-15;
    den = den * r + 1.42151175831644588870e-7; den = den * r + 1.84631831751005468180e-5;
    den = den * r + 7.86869131145613259100e-4; den = den * r + 1.48753612908506148525e-2;
    den = den * r + 1.36929880922735805310e-1; den = den * r + 5.99832206555887937690e-1;
    den = den * r + 1.0;
  }
  const double z = num / den;
  return (q < 0.0) ? -z : z;
}

// streams: 0 Z_T, 1 Z_H, 2 Z_E, 3 Z_F, 4 Z_To (Y-orthogonal part of X), 5 Z_Uo (X-orthogonal part of Y)
__device__ __forceinline__ double gen_normal(uint64_t seed, uint32_t stream, uint64_t row, uint32_t col) {
  uint32_t c[4] = {(uint32_t)row, (uint32_t)(row >> 32), col, stream};
  philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
  const double u = ((double)(c[0] >> 5) * 67108864.0 + (double)(c[1] >> 6) + 0.5) / 9007199254740992.0;
  return det_ndtri(u);
}

constexpr int SYN_ROWS = 8;
__global__ void __launch_bounds__(256) synth_kernel(double* __restrict__ D, const DataShape s, int64_t row0,
                                                    uint32_t cx0, uint32_t cy0, int R, uint64_t seed,
                                                    const double* __restrict__ Lt, const double* __restrict__ b,
                                                    const double* __restrict__ t, double sigE, double sigF,
                                                    double sigH, const SynthOrth o) {
  __shared__ double Ts[SYN_ROWS][MAXR], Us[SYN_ROWS][MAXR], Tos[SYN_ROWS][MAXR], Uos[SYN_ROWS][MAXR];
  const int64_t n = s.n; const int ld = s.ld, p = s.p, p2 = s.p2, q = s.q;
  const int64_t ngroups = (n + SYN_ROWS - 1) / SYN_ROWS;
  for (int64_t g = blockIdx.x; g < ngroups; g += gridDim.x) {
    const int64_t i0 = g * SYN_ROWS;
    __syncthreads();
    if ((int)threadIdx.x < SYN_ROWS * R) {
      const int i = threadIdx.x / R, k = threadIdx.x - i * R;
      if (i0 + i < n) {
        const uint64_t row = (uint64_t)(row0 + i0 + i);
        const double T = t[k] * gen_normal(seed, 0u, row, (uint32_t)k);
        Ts[i][k] = T;
        Us[i][k] = T * b[k] + sigH * gen_normal(seed, 1u, row, (uint32_t)k);
      }
    }
    // simulC's orthogonal parts (loglC.cpp:290-291,304-305): TYo = Z_To diag(sigTo), UXo = Z_Uo diag(sigUo)
    if ((int)threadIdx.x < SYN_ROWS * MAXR) {
      const int i = threadIdx.x / MAXR, k = threadIdx.x - i * MAXR;
      if (i0 + i < n) {
        const uint64_t row = (uint64_t)(row0 + i0 + i);
        if (k < o.nx) Tos[i][k] = o.sigTo[k] * gen_normal(seed, 4u, row, (uint32_t)k);
        if (k < o.ny) Uos[i][k] = o.sigUo[k] * gen_normal(seed, 5u, row, (uint32_t)k);
      }
    }
    __syncthreads();
    for (int i = 0; i < SYN_ROWS && i0 + i < n; ++i) {
      const uint64_t row = (uint64_t)(row0 + i0 + i);
      for (int j = threadIdx.x; j < ld; j += blockDim.x) {
        double v = 0.0;
        if (j < p) {
          double acc = 0.0;
          for (int k = 0; k < R; ++k) acc = acc + Ts[i][k] * Lt[(size_t)j * R + k];
          if (o.nx > 0) {
            double ao = 0.0;
            for (int k = 0; k < o.nx; ++k) ao = ao + Tos[i][k] * o.PYo[(size_t)j * o.nx + k];
            acc = acc + ao;
          }
          v = acc + sigE * gen_normal(seed, 2u, row, cx0 + (uint32_t)j);
        } else if (j >= p2 && j - p2 < q) {
          double acc = 0.0;
          for (int k = 0; k < R; ++k) acc = acc + Us[i][k] * Lt[(size_t)j * R + k];
          if (o.ny > 0) {
            double ao = 0.0;
            for (int k = 0; k < o.ny; ++k) ao = ao + Uos[i][k] * o.PXo[(size_t)(j - p2) * o.ny + k];
            acc = acc + ao;
          }
          v = acc + sigF * gen_normal(seed, 3u, row, cy0 + (uint32_t)(j - p2));
        }
        D[data_index(s, i0 + i, j)] = v;
      }
    }
  }
}

cudaError_t launch_synth(double* D, const DataShape& s, int64_t row0, uint32_t colX0, uint32_t colY0, int R, uint64_t seed, const double* Ltruth,
                         const double* b, const double* t, double sigE, double sigF, double sigH,
                         cudaStream_t stream, const SynthOrth* orth) {
  SynthOrth o{};
  if (orth) o = *orth;
  int64_t ngroups = (s.n + SYN_ROWS - 1) / SYN_ROWS;
  int grid = (int)(ngroups < 148 * 8 ? (ngroups < 1 ? 1 : ngroups) : 148 * 8);
  synth_kernel<<<grid, 256, 0, stream>>>(D, s, row0, colX0, colY0, R, seed, Ltruth, b, t, sigE, sigF, sigH, o);
  return cudaGetLastError();
}

// ---- initialGuess = "random" of PPLSi (Package/PPLS/R/EM_W_multi.R:133) on the counter-based generator ------------------
//   Wnw = orth(runif(p)); Cnw = orth(runif(q)); Bnw = rchisq(1,1); siglatnw = rchisq(2,100)/100; signw = rchisq(2,10)/100
// (orth of one column = the normalised vector; a chi-square with k degrees of freedom = the sum of k squared normals.)
// R's RNG stream cannot exist outside R: the draws come from Philox (seed, component), streams 16 (W), 17 (C), 18 (scalars),
// so only the DISTRIBUTION of the reference's start is reproduced, bit-exactly repeatable for a given seed.
__device__ __forceinline__ double gen_uniform(uint64_t seed, uint32_t stream, uint64_t row, uint32_t col) {
  uint32_t c[4] = {(uint32_t)row, (uint32_t)(row >> 32), col, stream};
  philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
  return ((double)(c[0] >> 5) * 67108864.0 + (double)(c[1] >> 6) + 0.5) / 9007199254740992.0;
}
__global__ void __launch_bounds__(256) guess_kernel(double* __restrict__ W, int p, double* __restrict__ C, int q,
                                                    double* __restrict__ sc /*B, sigE, sigF, sigH, sigT*/, uint64_t seed,
                                                    uint32_t comp) {
  __shared__ double red[256];
  for (int side = 0; side < 2; ++side) {
    double* v = side ? C : W;
    const int n = side ? q : p;
    double s = 0.0;
    for (int j = threadIdx.x; j < n; j += 256) { const double u = gen_uniform(seed, 16u + side, comp, (uint32_t)j); v[j] = u; s += u * u; }
    red[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) { if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o]; __syncthreads(); }
    const double nrm = sqrt(red[0]);
    __syncthreads();
    for (int j = threadIdx.x; j < n; j += 256) v[j] = v[j] / nrm;
  }
  if (threadIdx.x == 0) {
    auto chisq = [&](uint32_t first, int k) {
      double s = 0.0;
      for (int e = 0; e < k; ++e) { const double z = gen_normal(seed, 18u, comp, first + (uint32_t)e); s += z * z; }
      return s;
    };
    sc[0] = chisq(0u, 1);                  // Bnw
    sc[3] = chisq(1u, 100) / 100.0;        // siglatnw[1] = sigH
    sc[4] = chisq(101u, 100) / 100.0;      // siglatnw[2] = sigT
    sc[1] = chisq(201u, 10) / 100.0;       // signw[1] = sigE
    sc[2] = chisq(211u, 10) / 100.0;       // signw[2] = sigF
  }
}
cudaError_t launch_guess(double* W, int p, double* C, int q, double* sc, uint64_t seed, uint32_t comp, cudaStream_t stream) {
  guess_kernel<<<1, 256, 0, stream>>>(W, p, C, q, sc, seed, comp);
  return cudaGetLastError();
}

}  // namespace ppls
