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

constexpr int SG_BK = 32;

// BT x BT output tile per block of 256 threads (16 x 16 thread grid, (BT / 16)^2 outputs per thread), BT = 64 or 32.
// The operand tiles of step k + 1 are loaded from global memory into registers BEFORE the FMAs of step k (register double
// buffering), and problems too small to fill the machine with 64 x 64 tiles run on 32 x 32 tiles: the skinny GEMMs this
// kernel serves in the product path (IPA projections of a few hundred rows, K = 384) were 44 blocks on 148 SMs paying the
// load latency serially at every k step (38 us per launch on config 4, whose design sampler issues 15 of them per step).
template <typename AT, typename OT, int BT>
__global__ void __launch_bounds__(256)
gemm_simt_kernel(const AT* __restrict__ A, int lda, const float* __restrict__ W, int ldw, int M, int N,
                 int K, int epi, EpiParams p) {
  constexpr int TM = BT / 16;                      // outputs per thread along m and n
  __shared__ __align__(16) float As[SG_BK][BT + 4];
  __shared__ __align__(16) float Ws[SG_BK][BT + 4];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int m0 = blockIdx.y * BT, n0 = blockIdx.x * BT;
  float acc[TM][TM];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TM; ++j) acc[i][j] = 0.f;

  // each thread loads PT consecutive k of one A row and one W row per step
  constexpr int PT = BT * SG_BK / 256, TPR = SG_BK / PT;
  const int lr = threadIdx.x / TPR, lk = (threadIdx.x % TPR) * PT;
  const int am = m0 + lr, wn = n0 + lr;
  const AT* arow = A + (size_t)(am < M ? am : 0) * lda;
  const float* wrow = W + (size_t)(wn < N ? wn : 0) * ldw;
  float ra[PT], rw[PT];
  auto fetch = [&](int k0) {
#pragma unroll
    for (int e = 0; e < PT; ++e) {
      const int k = k0 + lk + e;
      ra[e] = (am < M && k < K) ? to_f(arow[k]) : 0.f;
      rw[e] = (wn < N && k < K) ? wrow[k] : 0.f;
    }
  };
  fetch(0);
  for (int k0 = 0; k0 < K; k0 += SG_BK) {
#pragma unroll
    for (int e = 0; e < PT; ++e) {
      As[lk + e][lr] = ra[e];
      Ws[lk + e][lr] = rw[e];
    }
    __syncthreads();
    if (k0 + SG_BK < K) fetch(k0 + SG_BK);          // in flight during the FMAs below
#pragma unroll
    for (int kk = 0; kk < SG_BK; ++kk) {
      float a[TM], w[TM];
#pragma unroll
      for (int i = 0; i < TM; ++i) a[i] = As[kk][ty * TM + i];
#pragma unroll
      for (int j = 0; j < TM; ++j) w[j] = Ws[kk][tx * TM + j];
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TM; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    int m = m0 + ty * TM + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < TM; ++j) {
      int n = n0 + tx * TM + j;
      if (n < N) apply_epilogue<OT>(epi, p, m, n, acc[i][j]);
    }
  }
}

template <typename AT, typename OT>
inline int launch_gemm_simt(const AT* A, int lda, const float* W, int ldw, int M, int N, int K, int epi,
                            const EpiParams& p, cudaStream_t st) {
  if (M <= 0 || N <= 0) return 0;
  ProfScope ps(PC_GEMM_SIMT, st);
  const long t64 = (long)((N + 63) / 64) * ((M + 63) / 64);
  if (t64 >= 148) {
    dim3 grid((N + 63) / 64, (M + 63) / 64);
    gemm_simt_kernel<AT, OT, 64><<<grid, 256, 0, st>>>(A, lda, W, ldw, M, N, K, epi, p);
  } else {                                           // too few 64 x 64 tiles to fill the machine
    dim3 grid((N + 31) / 32, (M + 31) / 32);
    gemm_simt_kernel<AT, OT, 32><<<grid, 256, 0, st>>>(A, lda, W, ldw, M, N, K, epi, p);
  }
  MD_LAUNCHED();
  MD_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace mdgen
