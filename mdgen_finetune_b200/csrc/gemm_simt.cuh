// gemm_simt.cuh -- fp32-accumulate SIMT GEMM  C[M,N] = A[M,K] . W[N,K]^T (+ epilogue).
// Used (a) for every projection in the fp32 check mode (north_star: 1e-4 rel-L2 mode),
// (b) for the fp32 IPA projections (ipa.py:113-138,250-254 keep geometry in fp32, see
// primitives.py:201-204) and (c) as the reference the tcgen05 GEMM is unit-tested against.
#pragma once
#include "common.cuh"

namespace mdgen {

template <typename OT>
__device__ __forceinline__ void apply_epilogue(int epi, const EpiParams& p, int m, int n, float acc) {
  float v = acc + (p.bias ? p.bias[n] : 0.f);
  if (epi == EPI_STORE) {
    reinterpret_cast<OT*>(p.out)[(size_t)m * p.ldo + n] = from_f<OT>(v);
  } else if (epi == EPI_GELU) {
    reinterpret_cast<OT*>(p.out)[(size_t)m * p.ldo + n] = from_f<OT>(gelu_erf(v));
  } else if (epi == EPI_GATE_RESID) {
    int b = (m / p.rows_per_batch) % p.nbatch;
    float g = p.gate[(size_t)b * p.gate_stride + n];
    p.resid[(size_t)m * p.ldr + n] += g * v;
  } else if (epi == EPI_RESID) {
    p.resid[(size_t)m * p.ldr + n] += v;
  } else {  // EPI_EULER
    reinterpret_cast<float*>(p.out)[(size_t)m * p.ldo + n] = v;
    if (p.resid) p.resid[(size_t)m * p.ldr + n] += p.dt * v;
  }
}

constexpr int SG_BM = 64, SG_BN = 64, SG_BK = 16;

template <typename AT, typename OT>
__global__ void __launch_bounds__(256)
gemm_simt_kernel(const AT* __restrict__ A, int lda, const float* __restrict__ W, int ldw, int M, int N,
                 int K, int epi, EpiParams p) {
  __shared__ float As[SG_BK][SG_BM + 4];
  __shared__ float Ws[SG_BK][SG_BN + 4];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int m0 = blockIdx.y * SG_BM, n0 = blockIdx.x * SG_BN;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  // each thread loads 4 elements of A and 4 of W per k-tile: row = tid/4, kk = (tid%4)*4..+3
  const int lr = threadIdx.x / 4, lk = (threadIdx.x % 4) * 4;
  for (int k0 = 0; k0 < K; k0 += SG_BK) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      int k = k0 + lk + e;
      int m = m0 + lr, n = n0 + lr;
      As[lk + e][lr] = (m < M && k < K) ? to_f(A[(size_t)m * lda + k]) : 0.f;
      Ws[lk + e][lr] = (n < N && k < K) ? W[(size_t)n * ldw + k] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < SG_BK; ++kk) {
      float a[4], w[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) w[j] = Ws[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + tx * 4 + j;
      if (n < N) apply_epilogue<OT>(epi, p, m, n, acc[i][j]);
    }
  }
}

template <typename AT, typename OT>
inline int launch_gemm_simt(const AT* A, int lda, const float* W, int ldw, int M, int N, int K, int epi,
                            const EpiParams& p, cudaStream_t st) {
  if (M <= 0 || N <= 0) return 0;
  ProfScope ps(PC_GEMM_SIMT, st);
  dim3 grid((N + SG_BN - 1) / SG_BN, (M + SG_BM - 1) / SG_BM);
  gemm_simt_kernel<AT, OT><<<grid, 256, 0, st>>>(A, lda, W, ldw, M, N, K, epi, p);
  MD_LAUNCHED();
  MD_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace mdgen
