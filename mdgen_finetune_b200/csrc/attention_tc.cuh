// attention_tc.cuh -- tcgen05 attention core for the long axes (frame axis T <= 1023 keys of the 4AA
// configs, 257 / 251 keys of ATLAS).  Semantics = attention_simt.cuh (mha.py:258-397): bias-KV token as
// key S, RoPE on q/k, key-padding -> -inf, fp32 softmax statistics, P.V.
//
// One CTA = one (sequence, head, chunk of <= 4 query tiles of 128).  The whole K (RoPE'd, d 24 -> 32
// zero padded) and V^T of the head live in shared memory as bf16 UMMA operands (no-swizzle K-major
// core-matrix layout).  Per query tile, two passes over the key blocks of 128:
//   pass 1: S = Q.K^T (tcgen05.mma, fp32 in TMEM)  -> row max            (no O rescaling needed later)
//   pass 2: S again -> P = exp2((S - max) log2e) -> bf16 P tile in smem -> O += P.V (tcgen05.mma)
// S is double buffered in TMEM so the MMA of block k+1 overlaps the softmax of block k.
// Warps 0-3: softmax / epilogue (thread = query row = TMEM lane), warp 4: MMA issuer, all 8 warps load K/V.
#pragma once
#include "attention_simt.cuh"
#include "gemm_tc.cuh"

namespace mdgen {
namespace atc {

using tc::mbar_arrive;
using tc::mbar_init;
using tc::mbar_wait;
using tc::smem_u32;
using tc::tcgen05_commit;
using tc::tcgen05_fence_after;
using tc::tcgen05_fence_before;
using tc::tmem_ld_32x32b_x32;
using tc::tmem_ld_wait;
using tc::umma_bf16;

constexpr int HD = 24, DP = 32;          // head dim and its MMA-padded width
constexpr int QT = 128, KB = 128;        // query tile rows, key block width
constexpr int MAX_KEYS = 1024;           // S + 1 <= 1024
constexpr int QT_PER_ITEM = 4;
constexpr int THREADS = 256;
// shared memory map (bytes)
constexpr int K_BYTES = MAX_KEYS * DP * 2;        // 65536  rows = keys,  LBO 128, SBO 512
constexpr int VT_BYTES = DP * MAX_KEYS * 2;       // 65536  rows = d,     LBO 128, SBO 16384
constexpr int Q_BYTES = QT * DP * 2;              // 8192   rows = q,     LBO 128, SBO 512
constexpr int P_BYTES = QT * KB * 2;              // 32768  rows = q,     LBO 128, SBO 2048   (x2 buffers)
constexpr int BIAS_BYTES = MAX_KEYS * 4;          // additive key bias: 0 or -inf
constexpr int OFF_K = 0, OFF_VT = OFF_K + K_BYTES, OFF_Q = OFF_VT + VT_BYTES, OFF_P = OFF_Q + Q_BYTES,
              OFF_BIAS = OFF_P + 2 * P_BYTES, OFF_BAR = OFF_BIAS + BIAS_BYTES, SMEM_BYTES = OFF_BAR + 256 + 128;
constexpr uint32_t TMEM_COLS = 512;
constexpr uint32_t TM_S0 = 0, TM_S1 = 128, TM_O = 256;

// no-swizzle K-major descriptor: start>>4 | LBO>>4 << 16 | SBO>>4 << 32 | version 1 << 46 | layout 0
__device__ __forceinline__ uint64_t desc_nosw(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
         ((uint64_t)1 << 46);
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
// core-matrix offset of element (row, col) in a K-major no-swizzle tile with the given SBO (LBO = 128)
__device__ __forceinline__ uint32_t cm_off(int row, int col, int sbo) {
  return (uint32_t)((row >> 3) * sbo + (col >> 3) * 128 + (row & 7) * 16 + (col & 7) * 2);
}

struct Bars {
  uint64_t s_full[2], s_empty[2], p_full[2], p_empty[2], q_full, o_full, o_empty;
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(THREADS, 1)
attention_tc_kernel(AttnParams<bf16> p, int chunks_per_seq) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
  Bars* bars = reinterpret_cast<Bars*>(smem + OFF_BAR);
  float* kbias = reinterpret_cast<float*>(smem + OFF_BIAS);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, tid = threadIdx.x;
  const int item = blockIdx.x;
  const int chunk = item % chunks_per_seq, h = (item / chunks_per_seq) % p.H, nbi = item / (chunks_per_seq * p.H);
  const int C = p.H * HD, S = p.S, nkeys = S + 1;
  const int nkb = (nkeys + KB - 1) / KB;
  const int q_begin = chunk * QT_PER_ITEM * QT;
  const int q_end = min(S, q_begin + QT_PER_ITEM * QT);
  const int ntiles = (q_end - q_begin + QT - 1) / QT;

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->s_full[i], 1);
      mbar_init(&bars->s_empty[i], 4);
      mbar_init(&bars->p_full[i], 4);
      mbar_init(&bars->p_empty[i], 1);
    }
    mbar_init(&bars->q_full, 4);
    mbar_init(&bars->o_full, 1);
    mbar_init(&bars->o_empty, 4);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 4) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&bars->tmem_base)), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // ---- load K (RoPE'd) and V^T of this head into shared memory: one key per thread-iteration -------
  for (int key = tid; key < nkb * KB; key += THREADS) {
    float kf[HD], vf[HD];
    bool valid = false;
    if (key < S) {
      const size_t r = attn_row(nbi, key, S, p.inner);
      const uint4* src = reinterpret_cast<const uint4*>(p.qkv + r * p.ld + C + h * HD);
      const uint4* srcv = reinterpret_cast<const uint4*>(p.qkv + r * p.ld + 2 * C + h * HD);
      uint4 kr[3] = {src[0], src[1], src[2]}, vr[3] = {srcv[0], srcv[1], srcv[2]};
      const bf16* kb16 = reinterpret_cast<const bf16*>(kr);
      const bf16* vb16 = reinterpret_cast<const bf16*>(vr);
#pragma unroll
      for (int i = 0; i < HD; ++i) { kf[i] = __bfloat162float(kb16[i]); vf[i] = __bfloat162float(vb16[i]); }
      valid = p.key_valid[r % p.mask_mod] != 0.f;
    } else if (key == S) {
#pragma unroll
      for (int i = 0; i < HD; ++i) { kf[i] = p.bias_k[h * HD + i]; vf[i] = p.bias_v[h * HD + i]; }
      valid = true;
    } else {
#pragma unroll
      for (int i = 0; i < HD; ++i) { kf[i] = 0.f; vf[i] = 0.f; }
    }
    if (p.cs && key <= S) {
#pragma unroll
      for (int i = 0; i < HD / 2; ++i) {
        const float c = p.cs[key * (HD / 2) + i], s = p.sn[key * (HD / 2) + i];
        const float a = kf[i], b = kf[i + HD / 2];
        kf[i] = a * c - b * s;
        kf[i + HD / 2] = b * c + a * s;
      }
    }
    uint8_t* krow = smem + OFF_K;
#pragma unroll
    for (int g = 0; g < 3; ++g) {
      uint4 v = make_uint4(pack_bf16(kf[g * 8], kf[g * 8 + 1]), pack_bf16(kf[g * 8 + 2], kf[g * 8 + 3]),
                           pack_bf16(kf[g * 8 + 4], kf[g * 8 + 5]), pack_bf16(kf[g * 8 + 6], kf[g * 8 + 7]));
      *reinterpret_cast<uint4*>(krow + cm_off(key, g * 8, 512)) = v;
    }
    *reinterpret_cast<uint4*>(krow + cm_off(key, 24, 512)) = make_uint4(0, 0, 0, 0);
    bf16* vt = reinterpret_cast<bf16*>(smem + OFF_VT);
#pragma unroll
    for (int d = 0; d < HD; ++d) vt[cm_off(d, key, 16384) >> 1] = __float2bfloat16_rn(vf[d]);
#pragma unroll
    for (int d = HD; d < DP; ++d) vt[cm_off(d, key, 16384) >> 1] = __float2bfloat16_rn(0.f);
    kbias[key] = valid ? 0.f : -INFINITY;
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic smem writes -> visible to tcgen05.mma
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = bars->tmem_base;
  const uint32_t sK = smem_u32(smem + OFF_K), sVT = smem_u32(smem + OFF_VT), sQ = smem_u32(smem + OFF_Q),
                 sP = smem_u32(smem + OFF_P);

  if (warp == 4) {
    // =============================== MMA issuer ===============================
    if (lane == 0) {
      constexpr uint32_t idesc_s = tc::make_idesc_bf16(QT, KB);   // S[128 x 128] = Q[128x32] K[128x32]^T
      constexpr uint32_t idesc_o = tc::make_idesc_bf16(QT, DP);   // O[128 x 32] += P[128x128] V^T[32x128]^T
      uint32_t s_it = 0, p_it = 0;
      auto issue_s = [&](int kb) {
        const int b = s_it & 1;
        mbar_wait(&bars->s_empty[b], ((s_it >> 1) & 1) ^ 1);
        tcgen05_fence_after();
        const uint32_t d = tmem + (b ? TM_S1 : TM_S0);
#pragma unroll
        for (int k = 0; k < DP / 16; ++k)
          umma_bf16(d, desc_nosw(sQ + k * 256, 128, 512), desc_nosw(sK + kb * (KB / 8) * 512 + k * 256, 128, 512), idesc_s, k);
        tcgen05_commit(&bars->s_full[b]);
        ++s_it;
      };
      auto issue_pv = [&](int kb, bool first) {
        const int b = p_it & 1;
        mbar_wait(&bars->p_full[b], (p_it >> 1) & 1);
        tcgen05_fence_after();
#pragma unroll
        for (int k = 0; k < KB / 16; ++k)
          umma_bf16(tmem + TM_O, desc_nosw(sP + b * P_BYTES + k * 256, 128, 2048),
                    desc_nosw(sVT + (kb * (KB / 8) + k * 2) * 128, 128, 16384), idesc_o, (first && k == 0) ? 0u : 1u);
        tcgen05_commit(&bars->p_empty[b]);
        ++p_it;
      };
      for (int t = 0; t < ntiles; ++t) {
        mbar_wait(&bars->q_full, t & 1);
        tcgen05_fence_after();
        for (int kb = 0; kb < nkb; ++kb) issue_s(kb);                 // pass 1: row maxima
        mbar_wait(&bars->o_empty, (t & 1) ^ 1);                       // previous tile's O drained
        tcgen05_fence_after();
        for (int kb = 0; kb < nkb; ++kb) {                            // pass 2: probabilities and P.V
          issue_s(kb);
          if (kb > 0) issue_pv(kb - 1, kb == 1);
        }
        issue_pv(nkb - 1, nkb == 1);
        tcgen05_commit(&bars->o_full);
      }
    }
  } else if (warp < 4) {
    // =============================== softmax / epilogue ===============================
    const int r = tid;                                   // query row in tile = TMEM lane
    const uint32_t lane_addr = ((uint32_t)(warp * 32) << 16);
    const float LOG2E = 1.4426950408889634f;
    uint32_t s_it = 0, p_it = 0;
    for (int t = 0; t < ntiles; ++t) {
      const int q = q_begin + t * QT + r;
      // ---- Q row: global -> RoPE -> bf16 -> smem (padded to 32)
      {
        float qf[HD];
        if (q < S) {
          const uint4* src = reinterpret_cast<const uint4*>(p.qkv + attn_row(nbi, q, S, p.inner) * p.ld + h * HD);
          uint4 qr[3] = {src[0], src[1], src[2]};
          const bf16* q16 = reinterpret_cast<const bf16*>(qr);
#pragma unroll
          for (int i = 0; i < HD; ++i) qf[i] = __bfloat162float(q16[i]);
          if (p.cs) {
#pragma unroll
            for (int i = 0; i < HD / 2; ++i) {
              const float c = p.cs[q * (HD / 2) + i], s = p.sn[q * (HD / 2) + i];
              const float a = qf[i], b = qf[i + HD / 2];
              qf[i] = a * c - b * s;
              qf[i + HD / 2] = b * c + a * s;
            }
          }
        } else {
#pragma unroll
          for (int i = 0; i < HD; ++i) qf[i] = 0.f;
        }
        uint8_t* qrow = smem + OFF_Q;
#pragma unroll
        for (int g = 0; g < 3; ++g)
          *reinterpret_cast<uint4*>(qrow + cm_off(r, g * 8, 512)) =
              make_uint4(pack_bf16(qf[g * 8], qf[g * 8 + 1]), pack_bf16(qf[g * 8 + 2], qf[g * 8 + 3]),
                         pack_bf16(qf[g * 8 + 4], qf[g * 8 + 5]), pack_bf16(qf[g * 8 + 6], qf[g * 8 + 7]));
        *reinterpret_cast<uint4*>(qrow + cm_off(r, 24, 512)) = make_uint4(0, 0, 0, 0);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->q_full);
      }
      // ---- pass 1: row max
      float mrow = -INFINITY;
      for (int kb = 0; kb < nkb; ++kb) {
        const int b = s_it & 1;
        mbar_wait(&bars->s_full[b], (s_it >> 1) & 1);
        tcgen05_fence_after();
        const uint32_t ta = tmem + lane_addr + (b ? TM_S1 : TM_S0);
#pragma unroll
        for (int c = 0; c < KB / 32; ++c) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(ta + c * 32, v);
          tmem_ld_wait();
          const float* kbp = kbias + kb * KB + c * 32;
#pragma unroll
          for (int j = 0; j < 32; ++j) mrow = fmaxf(mrow, __uint_as_float(v[j]) + kbp[j]);
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->s_empty[b]);
        ++s_it;
      }
      const float mneg = -mrow * LOG2E;
      // ---- pass 2: P = exp2(S log2e - m log2e), row sum, P -> smem
      float lsum = 0.f;
      for (int kb = 0; kb < nkb; ++kb) {
        const int b = s_it & 1, pb = p_it & 1;
        mbar_wait(&bars->s_full[b], (s_it >> 1) & 1);
        mbar_wait(&bars->p_empty[pb], ((p_it >> 1) & 1) ^ 1);
        tcgen05_fence_after();
        const uint32_t ta = tmem + lane_addr + (b ? TM_S1 : TM_S0);
        uint8_t* prow = smem + OFF_P + pb * P_BYTES + (r >> 3) * 2048 + (r & 7) * 16;
#pragma unroll
        for (int c = 0; c < KB / 32; ++c) {
          uint32_t v[32];
          tmem_ld_32x32b_x32(ta + c * 32, v);
          tmem_ld_wait();
          const float* kbp = kbias + kb * KB + c * 32;
          float e[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            e[j] = ex2(fmaf(__uint_as_float(v[j]) + kbp[j], LOG2E, mneg));
            lsum += e[j];
          }
#pragma unroll
          for (int g = 0; g < 4; ++g)
            *reinterpret_cast<uint4*>(prow + (c * 4 + g) * 128) =
                make_uint4(pack_bf16(e[g * 8], e[g * 8 + 1]), pack_bf16(e[g * 8 + 2], e[g * 8 + 3]),
                           pack_bf16(e[g * 8 + 4], e[g * 8 + 5]), pack_bf16(e[g * 8 + 6], e[g * 8 + 7]));
        }
        tcgen05_fence_before();
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&bars->s_empty[b]);
          mbar_arrive(&bars->p_full[pb]);
        }
        ++s_it;
        ++p_it;
      }
      // ---- epilogue: O / rowsum -> bf16 -> global
      mbar_wait(&bars->o_full, t & 1);
      tcgen05_fence_after();
      {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem + lane_addr + TM_O, v);
        tmem_ld_wait();
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->o_empty);
        if (q < S) {
          const float inv = 1.0f / lsum;
          uint4* dst = reinterpret_cast<uint4*>(p.out + attn_row(nbi, q, S, p.inner) * p.ldo + h * HD);
#pragma unroll
          for (int g = 0; g < 3; ++g)
            dst[g] = make_uint4(pack_bf16(__uint_as_float(v[g * 8]) * inv, __uint_as_float(v[g * 8 + 1]) * inv),
                                pack_bf16(__uint_as_float(v[g * 8 + 2]) * inv, __uint_as_float(v[g * 8 + 3]) * inv),
                                pack_bf16(__uint_as_float(v[g * 8 + 4]) * inv, __uint_as_float(v[g * 8 + 5]) * inv),
                                pack_bf16(__uint_as_float(v[g * 8 + 6]) * inv, __uint_as_float(v[g * 8 + 7]) * inv));
        }
      }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 4) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
  }
}

}  // namespace atc

inline bool attention_tc_supported(const AttnParams<bf16>& p, int head_dim) {
  return head_dim == atc::HD && p.S + 1 <= atc::MAX_KEYS && p.S >= 64 && (p.ld % 8) == 0 && (p.ldo % 8) == 0;
}

inline int launch_attention_tc(const AttnParams<bf16>& p, int head_dim, cudaStream_t st) {
  MD_CHECK(attention_tc_supported(p, head_dim), "attention_tc: unsupported shape");
  static bool configured = false;
  if (!configured) {
    MD_CUDA(cudaFuncSetAttribute(atc::attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, atc::SMEM_BYTES));
    configured = true;
  }
  const int chunks = (p.S + atc::QT_PER_ITEM * atc::QT - 1) / (atc::QT_PER_ITEM * atc::QT);
  ProfScope ps(PC_ATTN_TC, st);
  atc::attention_tc_kernel<<<p.nb * p.H * chunks, atc::THREADS, atc::SMEM_BYTES, st>>>(p, chunks);
  MD_LAUNCHED();
  MD_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace mdgen
