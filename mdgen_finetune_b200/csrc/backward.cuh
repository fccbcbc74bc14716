// backward.cuh -- fp32 adjoint kernels of the velocity network (row N4: gradients for RTB posterior fine-tuning,
// samplers.py:120-146, gfn_diffusion.py:459-460).  Each kernel is the adjoint of one forward op of train.cuh's tape and
// ACCUMULATES (+=) into its gradient outputs, so shared weights (the IPA layers run on two streams) and fan-out sum up.
// Sizes on this path are small (RTB runs T = 1: 128 tokens per call), so the kernels favour clarity over tiling.
#pragma once
#include <math.h>

#include "attention_simt.cuh"
#include "common.cuh"
#include "ipa.cuh"

namespace mdgen {
namespace bwd {

// ---- generic strided GEMM: C[i, j] (+)= sum_k A[i * sa0 + k * sa1] * B[k * sb0 + j * sb1] -----------------------------
constexpr int GT = 32;
__global__ void __launch_bounds__(GT * 8)
gemm_strided_kernel(const float* __restrict__ A, long sa0, long sa1, const float* __restrict__ B, long sb0, long sb1,
                    float* __restrict__ C, long ldc, int M, int N, int K, int accumulate, float alpha) {
  __shared__ float As[GT][GT + 1], Bs[GT][GT + 1];
  const int tx = threadIdx.x % GT, ty = threadIdx.x / GT;     // ty in [0, 8): each thread owns 4 rows
  const int i0 = blockIdx.y * GT, j0 = blockIdx.x * GT;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int k0 = 0; k0 < K; k0 += GT) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int rr = ty + 8 * r;
      const int i = i0 + rr, k = k0 + tx;
      As[rr][tx] = (i < M && k < K) ? A[(long)i * sa0 + (long)k * sa1] : 0.f;
      const int kb = k0 + rr, j = j0 + tx;
      Bs[rr][tx] = (kb < K && j < N) ? B[(long)kb * sb0 + (long)j * sb1] : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < GT; ++k) {
      const float b = Bs[k][tx];
#pragma unroll
      for (int r = 0; r < 4; ++r) acc[r] = fmaf(As[ty + 8 * r][k], b, acc[r]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int i = i0 + ty + 8 * r, j = j0 + tx;
    if (i < M && j < N) {
      float* c = C + (long)i * ldc + j;
      *c = accumulate ? *c + alpha * acc[r] : alpha * acc[r];
    }
  }
}
inline int gemm_strided(cudaStream_t st, const float* A, long sa0, long sa1, const float* B, long sb0, long sb1, float* C,
                        long ldc, int M, int N, int K, bool accumulate, float alpha = 1.f) {
  if (M <= 0 || N <= 0) return 0;
  dim3 grid((N + GT - 1) / GT, (M + GT - 1) / GT);
  gemm_strided_kernel<<<grid, GT * 8, 0, st>>>(A, sa0, sa1, B, sb0, sb1, C, ldc, M, N, K, accumulate ? 1 : 0, alpha);
  MD_LAUNCHED();
  MD_CUDA(cudaGetLastError());
  return 0;
}

// db[j] += alpha * sum_i dY[i, j]
__global__ void colsum_kernel(const float* __restrict__ dY, long ld, int M, int N, float* __restrict__ db, float alpha) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= N) return;
  float s = 0.f;
  for (int i = blockIdx.y; i < M; i += gridDim.y) s += dY[(long)i * ld + j];
  atomicAdd(db + j, alpha * s);
}
inline int colsum(cudaStream_t st, const float* dY, long ld, int M, int N, float* db, float alpha = 1.f) {
  if (M <= 0 || N <= 0) return 0;
  dim3 grid((N + 127) / 128, M < 64 ? 1 : 64);
  colsum_kernel<<<grid, 128, 0, st>>>(dY, ld, M, N, db, alpha);
  MD_LAUNCHED();
  MD_CUDA(cudaGetLastError());
  return 0;
}

// ---- element-wise adjoints -------------------------------------------------------------------------------------
// y = x * sigmoid(x):  dx += dy * (s + x s (1 - s))
__global__ void silu_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dx, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float v = x[i], s = 1.0f / (1.0f + expf(-v));
  dx[i] += dy[i] * (s + v * s * (1.0f - s));
}
// y = x Phi(x) (exact erf GELU, layers.py:78-85):  dx += dy * (Phi(x) + x phi(x))
__global__ void gelu_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dx, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float v = x[i];
  const float Phi = 0.5f * (1.0f + erff(v * 0.70710678118654752440f));
  const float phi = 0.39894228040143267794f * expf(-0.5f * v * v);
  dx[i] += dy[i] * (Phi + v * phi);
}
__global__ void gelu_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = gelu_erf(x[i]);
}
__global__ void add_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ y, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = a[i] + b[i];
}
__global__ void acc_kernel(float* __restrict__ y, const float* __restrict__ x, float a, size_t n) {   // y += a x
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] += a * x[i];
}

// ---- LayerNorm (+ adaLN modulate or affine), one warp per row ---------------------------------------------------
// forward (elementwise.cuh ln_modulate_kernel): h = xhat * (1 + scale[b]) + shift[b]   or   h = xhat * gamma + beta,
// xhat = (x - mean) * rstd, batch b = (row / rows_per_batch) % nbatch, shift / scale rows have pitch `ada_ld`.
// adjoint: g = dh * (1 + scale) (or dh * gamma);  dx += rstd * (g - mean(g) - xhat * mean(g * xhat));
//          dshift[b] += dh, dscale[b] += dh * xhat (or dbeta += dh, dgamma += dh * xhat), atomically over the rows.
__global__ void __launch_bounds__(256)
ln_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dh, float* __restrict__ dx, int M, int C,
              const float* __restrict__ scale, float* __restrict__ dshift, float* __restrict__ dscale, int ada_ld,
              int rows_per_batch, int nbatch, const float* __restrict__ gamma, float* __restrict__ dgamma,
              float* __restrict__ dbeta, float eps) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= M) return;
  const float* xr = x + (size_t)row * C;
  const float* dr = dh + (size_t)row * C;
  float s = 0.f;
  for (int c = lane; c < C; c += 32) s += xr[c];
  const float mean = warp_sum(s) / C;
  float v = 0.f;
  for (int c = lane; c < C; c += 32) {
    const float d = xr[c] - mean;
    v = fmaf(d, d, v);
  }
  const float rstd = rsqrtf(warp_sum(v) / C + eps);
  const int b = (row / rows_per_batch) % nbatch;
  const float* sc = scale ? scale + (size_t)b * ada_ld : nullptr;
  float sg = 0.f, sgx = 0.f;
  for (int c = lane; c < C; c += 32) {
    const float xh = (xr[c] - mean) * rstd;
    const float g = dr[c] * (sc ? 1.0f + sc[c] : (gamma ? gamma[c] : 1.0f));
    sg += g;
    sgx = fmaf(g, xh, sgx);
  }
  sg = warp_sum(sg) / C;
  sgx = warp_sum(sgx) / C;
  for (int c = lane; c < C; c += 32) {
    const float xh = (xr[c] - mean) * rstd;
    const float d = dr[c];
    const float g = d * (sc ? 1.0f + sc[c] : (gamma ? gamma[c] : 1.0f));
    dx[(size_t)row * C + c] += rstd * (g - sg - xh * sgx);
    if (dshift) atomicAdd(dshift + (size_t)b * ada_ld + c, d);
    if (dscale) atomicAdd(dscale + (size_t)b * ada_ld + c, d * xh);
    if (dbeta) atomicAdd(dbeta + c, d);
    if (dgamma) atomicAdd(dgamma + c, d * xh);
  }
}

// ---- gated residual: xo = x + gate[b] * y  (gate = 1 when gate == null) ------------------------------------------
__global__ void gate_resid_fwd_kernel(const float* __restrict__ x, const float* __restrict__ y, const float* __restrict__ gate,
                                      int gate_ld, int rows_per_batch, int nbatch, float* __restrict__ xo, int M, int C) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)M * C) return;
  const int row = (int)(i / C), c = (int)(i % C);
  const float g = gate ? gate[(size_t)((row / rows_per_batch) % nbatch) * gate_ld + c] : 1.0f;
  xo[i] = x[i] + g * y[i];
}
// dx += dxo;  dy += gate * dxo;  dgate[b] += sum_rows dxo * y
__global__ void gate_resid_bwd_kernel(const float* __restrict__ dxo, const float* __restrict__ y, const float* __restrict__ gate,
                                      float* __restrict__ dgate, int gate_ld, int rows_per_batch, int nbatch,
                                      float* __restrict__ dx, float* __restrict__ dy, int M, int C) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)M * C) return;
  const int row = (int)(i / C), c = (int)(i % C);
  const int b = (row / rows_per_batch) % nbatch;
  const float d = dxo[i];
  dx[i] += d;
  if (gate) {
    dy[i] += gate[(size_t)b * gate_ld + c] * d;
    atomicAdd(dgate + (size_t)b * gate_ld + c, d * y[i]);
  } else {
    dy[i] += d;
  }
}

}  // namespace bwd
}  // namespace mdgen
