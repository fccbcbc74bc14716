// attention_simt.cuh -- fp32 flash-style attention core of MultiheadAttention as reached from
// AttentionWithRoPE (mha.py:258-397, latent_model.py:334-338): learned bias-KV token appended as
// key S (mha.py:265-268), RoPE on q (positions 0..S-1) and k (0..S) with rotate_half pairing
// (fair-esm), -inf key-padding fill (mha.py:370-376), fp32 softmax (mha.py:381), P.V.
// One thread = one query; keys streamed through shared memory in chunks with online softmax.
// This is the fp32-check-mode core and the short-sequence (residue axis, L = 4) core.
#pragma once
#include <math.h>

#include "common.cuh"

namespace mdgen {

template <typename AT>
struct AttnParams {
  const AT* qkv;  // [rows, ld]: q | k | v, each C = H*HD wide; q pre-scaled by HD^-0.5
  int ld;
  AT* out;        // [rows, ldo]
  int ldo;
  int nb, S, inner, H;
  const float* bias_k;     // [C]
  const float* bias_v;     // [C]
  const float* key_valid;  // [mask_mod] 1 = valid, indexed by row % mask_mod
  int mask_mod;
  const float* cs;  // [S+1, HD/2] or null (no rope)
  const float* sn;
};

__host__ __device__ __forceinline__ size_t attn_row(int nbi, int s, int S, int inner) {
  return (size_t)(nbi / inner) * S * inner + (size_t)s * inner + (nbi % inner);
}

// 24 contiguous elements (16-byte aligned) -> fp32 registers with 128-bit loads
__device__ __forceinline__ void load24(const float* p, float* o) {
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(p) + i);
    o[4 * i] = v.x; o[4 * i + 1] = v.y; o[4 * i + 2] = v.z; o[4 * i + 3] = v.w;
  }
}
__device__ __forceinline__ void load24(const bf16* p, float* o) {
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(p) + i);
    const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = __bfloat1622float2(h2[j]);
      o[8 * i + 2 * j] = f.x; o[8 * i + 2 * j + 1] = f.y;
    }
  }
}
__device__ __forceinline__ void store24(float* p, const float* o) {
#pragma unroll
  for (int i = 0; i < 6; ++i) reinterpret_cast<float4*>(p)[i] = make_float4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
}
__device__ __forceinline__ void store24(bf16* p, const float* o) {
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    uint4 v;
    __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
    for (int j = 0; j < 4; ++j) h2[j] = __floats2bfloat162_rn(o[8 * i + 2 * j], o[8 * i + 2 * j + 1]);
    reinterpret_cast<uint4*>(p)[i] = v;
  }
}

constexpr int AS_THREADS = 128;
constexpr int AS_CHUNK = 64;

template <typename AT, int HD>
__global__ void __launch_bounds__(AS_THREADS)
attention_simt_kernel(AttnParams<AT> p) {
  constexpr int HALF = HD / 2;
  __shared__ float Ks[AS_CHUNK][HD + 1];
  __shared__ float Vs[AS_CHUNK][HD + 1];
  __shared__ float valid[AS_CHUNK];
  const int nbi = blockIdx.x / p.H, h = blockIdx.x % p.H;
  const int C = p.H * HD;
  const int qi = blockIdx.y * AS_THREADS + threadIdx.x;
  const bool active = qi < p.S;

  float q[HD], acc[HD];
#pragma unroll
  for (int i = 0; i < HD; ++i) acc[i] = 0.f;
  if (active) {
    const AT* qr = p.qkv + attn_row(nbi, qi, p.S, p.inner) * p.ld + h * HD;
    float raw[HD];
#pragma unroll
    for (int i = 0; i < HD; ++i) raw[i] = to_f(qr[i]);
    if (p.cs) {
#pragma unroll
      for (int i = 0; i < HALF; ++i) {
        float c = p.cs[qi * HALF + i], s = p.sn[qi * HALF + i];
        q[i] = raw[i] * c - raw[i + HALF] * s;
        q[i + HALF] = raw[i + HALF] * c + raw[i] * s;
      }
    } else {
#pragma unroll
      for (int i = 0; i < HD; ++i) q[i] = raw[i];
    }
  }
  float mrun = -INFINITY, lrun = 0.f;
  const int nkeys = p.S + 1;
  for (int k0 = 0; k0 < nkeys; k0 += AS_CHUNK) {
    __syncthreads();
    // cooperative load: (key, pair) items
    for (int it = threadIdx.x; it < AS_CHUNK * HALF; it += AS_THREADS) {
      int kk = it / HALF, i = it % HALF, key = k0 + kk;
      float a = 0.f, b = 0.f, va = 0.f, vb = 0.f;
      if (key < p.S) {
        const AT* kr = p.qkv + attn_row(nbi, key, p.S, p.inner) * p.ld + C + h * HD;
        a = to_f(kr[i]);
        b = to_f(kr[i + HALF]);
        va = to_f(kr[C + i]);
        vb = to_f(kr[C + i + HALF]);
      } else if (key == p.S) {
        a = p.bias_k[h * HD + i];
        b = p.bias_k[h * HD + i + HALF];
        va = p.bias_v[h * HD + i];
        vb = p.bias_v[h * HD + i + HALF];
      }
      if (p.cs && key <= p.S) {
        float c = p.cs[key * HALF + i], s = p.sn[key * HALF + i];
        float ra = a * c - b * s, rb = b * c + a * s;
        a = ra;
        b = rb;
      }
      Ks[kk][i] = a;
      Ks[kk][i + HALF] = b;
      Vs[kk][i] = va;
      Vs[kk][i + HALF] = vb;
    }
    for (int kk = threadIdx.x; kk < AS_CHUNK; kk += AS_THREADS) {
      int key = k0 + kk;
      float v = 0.f;
      if (key < p.S) v = p.key_valid[attn_row(nbi, key, p.S, p.inner) % p.mask_mod];
      else if (key == p.S) v = 1.f;
      valid[kk] = v;
    }
    __syncthreads();
    if (!active) continue;
    const int kend = min(AS_CHUNK, nkeys - k0);
    for (int kk = 0; kk < kend; ++kk) {
      if (valid[kk] == 0.f) continue;  // -inf logit: contributes exactly 0
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < HD; ++i) s = fmaf(q[i], Ks[kk][i], s);
      if (s > mrun) {
        float corr = expf(mrun - s);  // mrun = -inf -> 0
        lrun *= corr;
#pragma unroll
        for (int i = 0; i < HD; ++i) acc[i] *= corr;
        mrun = s;
      }
      float pr = expf(s - mrun);
      lrun += pr;
#pragma unroll
      for (int i = 0; i < HD; ++i) acc[i] = fmaf(pr, Vs[kk][i], acc[i]);
    }
  }
  if (active) {
    AT* o = p.out + attn_row(nbi, qi, p.S, p.inner) * p.ldo + h * HD;
    const float inv = 1.0f / lrun;
#pragma unroll
    for (int i = 0; i < HD; ++i) o[i] = from_f<AT>(acc[i] * inv);
  }
}

// Short sequences (residue axis of the tetrapeptide configs: S = L = 4, 5 keys): one thread per
// (sequence, head, query).  Each thread loads and rotates ITS OWN token's q, k, v once; the S threads of a
// (sequence, head) group exchange the rotated keys / values through shared memory (layout [key][float4][group],
// conflict-free), so every K/V element is read from HBM and rotated exactly once.
constexpr int ASH_THREADS = 128;
template <typename AT, int HD, int S_>
__global__ void __launch_bounds__(ASH_THREADS, 6)
attention_short_kernel(AttnParams<AT> p, long long total_groups) {
  constexpr int HALF = HD / 2, G = ASH_THREADS / S_, F4 = HD / 4;
  __shared__ float4 ks[S_][F4][G], vs[S_][F4][G];
  __shared__ float kvalid[S_][G];
  __shared__ __align__(16) float bks[16][HD + 4], bvs[16][HD + 4];   // bias token per head, key rotated to position S
  const int g = threadIdx.x / S_, qi = threadIdx.x % S_;
  const long long grp = (long long)blockIdx.x * G + g;     // (sequence, head) with head fastest
  const bool live = g < G && grp < total_groups;
  const int h = live ? (int)(grp % p.H) : 0;
  const int nbi = live ? (int)(grp / p.H) : 0;
  const int C = p.H * HD;
  for (int i = threadIdx.x; i < p.H * HALF; i += ASH_THREADS) {
    const int hh = i / HALF, ii = i % HALF;
    const float a = p.bias_k[hh * HD + ii], b = p.bias_k[hh * HD + ii + HALF];
    const float c = p.cs ? p.cs[S_ * HALF + ii] : 1.f, sn = p.cs ? p.sn[S_ * HALF + ii] : 0.f;
    bks[hh][ii] = a * c - b * sn;
    bks[hh][ii + HALF] = b * c + a * sn;
    bvs[hh][ii] = p.bias_v[hh * HD + ii];
    bvs[hh][ii + HALF] = p.bias_v[hh * HD + ii + HALF];
  }
  float q[HD];
  if (live) {
    const size_t r = attn_row(nbi, qi, S_, p.inner);
    const AT* row = p.qkv + r * p.ld + h * HD;
    float k[HD], v[HD];
    load24(row, q);
    load24(row + C, k);
    load24(row + 2 * C, v);
    if (p.cs) {
      float cc[HALF], ss[HALF];
#pragma unroll
      for (int j = 0; j < HALF / 4; ++j) {
        const float4 c4 = __ldg(reinterpret_cast<const float4*>(p.cs + qi * HALF) + j);
        const float4 s4 = __ldg(reinterpret_cast<const float4*>(p.sn + qi * HALF) + j);
        cc[4 * j] = c4.x; cc[4 * j + 1] = c4.y; cc[4 * j + 2] = c4.z; cc[4 * j + 3] = c4.w;
        ss[4 * j] = s4.x; ss[4 * j + 1] = s4.y; ss[4 * j + 2] = s4.z; ss[4 * j + 3] = s4.w;
      }
#pragma unroll
      for (int i = 0; i < HALF; ++i) {
        const float c = cc[i], sn = ss[i];
        const float a = q[i], b = q[i + HALF], ka = k[i], kb = k[i + HALF];
        q[i] = a * c - b * sn;
        q[i + HALF] = b * c + a * sn;
        k[i] = ka * c - kb * sn;
        k[i + HALF] = kb * c + ka * sn;
      }
    }
#pragma unroll
    for (int j = 0; j < F4; ++j) {
      ks[qi][j][g] = make_float4(k[4 * j], k[4 * j + 1], k[4 * j + 2], k[4 * j + 3]);
      vs[qi][j][g] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
    kvalid[qi][g] = p.key_valid[r % p.mask_mod];
  }
  __syncthreads();
  if (!live) return;
  float acc[HD];
#pragma unroll
  for (int i = 0; i < HD; ++i) acc[i] = 0.f;
  float mrun = -INFINITY, lrun = 0.f;
#pragma unroll
  for (int key = 0; key <= S_; ++key) {
    float kv[HD], vv[HD];
    if (key < S_) {
      if (kvalid[key][g] == 0.f) continue;
#pragma unroll
      for (int j = 0; j < F4; ++j) {
        const float4 a = ks[key][j][g], b = vs[key][j][g];
        kv[4 * j] = a.x; kv[4 * j + 1] = a.y; kv[4 * j + 2] = a.z; kv[4 * j + 3] = a.w;
        vv[4 * j] = b.x; vv[4 * j + 1] = b.y; vv[4 * j + 2] = b.z; vv[4 * j + 3] = b.w;
      }
    } else {   // learned bias token at position S (mha.py:265-268), rotated like any key
#pragma unroll
      for (int j = 0; j < F4; ++j) {
        const float4 a = *reinterpret_cast<const float4*>(&bks[h][4 * j]);
        const float4 b = *reinterpret_cast<const float4*>(&bvs[h][4 * j]);
        kv[4 * j] = a.x; kv[4 * j + 1] = a.y; kv[4 * j + 2] = a.z; kv[4 * j + 3] = a.w;
        vv[4 * j] = b.x; vv[4 * j + 1] = b.y; vv[4 * j + 2] = b.z; vv[4 * j + 3] = b.w;
      }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < HD; ++i) s = fmaf(q[i], kv[i], s);
    if (s > mrun) {
      const float corr = __expf(mrun - s);
      lrun *= corr;
#pragma unroll
      for (int i = 0; i < HD; ++i) acc[i] *= corr;
      mrun = s;
    }
    const float pr = __expf(s - mrun);
    lrun += pr;
#pragma unroll
    for (int i = 0; i < HD; ++i) acc[i] = fmaf(pr, vv[i], acc[i]);
  }
  const float inv = 1.0f / lrun;
#pragma unroll
  for (int i = 0; i < HD; ++i) acc[i] *= inv;
  store24(p.out + attn_row(nbi, qi, S_, p.inner) * p.ldo + h * HD, acc);
}

// S <= 4 (the tetrapeptide residue axis): one thread per (sequence, head) owns ALL S queries.  Nothing is exchanged
// between threads, every q/k/v element is loaded, converted and rotated once, scores for the S x (S+1) pairs are kept in
// registers and the softmax is taken without running-max rescaling.  ~2.4x fewer instructions per query than the
// thread-per-query kernel above, and consecutive lanes read consecutive 48-byte head slices of the same token row.
// rotate one 24-wide head vector to position pos (rotate_half pairing i, i+12); tables read as 6 float4
__device__ __forceinline__ void rope24(float* x, const float* cs, const float* sn, int pos) {
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const float4 c4 = __ldg(reinterpret_cast<const float4*>(cs + pos * 12) + j);
    const float4 s4 = __ldg(reinterpret_cast<const float4*>(sn + pos * 12) + j);
    const float cc[4] = {c4.x, c4.y, c4.z, c4.w}, ss[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int i = 4 * j + e;
      const float a = x[i], b = x[i + 12];
      x[i] = a * cc[e] - b * ss[e];
      x[i + 12] = b * cc[e] + a * ss[e];
    }
  }
}
constexpr int ASQ_THREADS = 128;
template <typename AT, int HD, int S_>
__global__ void __launch_bounds__(ASQ_THREADS)
attention_seq_kernel(AttnParams<AT> p, long long total_groups) {
  constexpr int HALF = HD / 2;
  const long long grp = (long long)blockIdx.x * ASQ_THREADS + threadIdx.x;   // (sequence, head), head fastest
  if (grp >= total_groups) return;
  const int h = (int)(grp % p.H), nbi = (int)(grp / p.H);
  const int C = p.H * HD;
  float q[S_][HD];
  float sc[S_][S_ + 1];
  const AT* rows[S_];
  bool valid[S_];
#pragma unroll
  for (int t = 0; t < S_; ++t) {
    const size_t r = attn_row(nbi, t, S_, p.inner);
    rows[t] = p.qkv + r * p.ld + h * HD;
    valid[t] = p.key_valid[r % p.mask_mod] != 0.f;
    load24(rows[t], q[t]);
  }
  static_assert(HD == 24, "rope24 is written for head_dim 24");
  if (p.cs) {
#pragma unroll
    for (int t = 0; t < S_; ++t) rope24(q[t], p.cs, p.sn, t);
  }
  // ---- scores: keys 0..S-1 from the tokens, key S = learned bias token (mha.py:265-268), all rotated by position
#pragma unroll
  for (int j = 0; j <= S_; ++j) {
    float k[HD];
    if (j < S_) {
      load24(rows[j] + C, k);
    } else {
#pragma unroll
      for (int i = 0; i < HD; ++i) k[i] = __ldg(p.bias_k + h * HD + i);
    }
    if (p.cs) rope24(k, p.cs, p.sn, j);
#pragma unroll
    for (int t = 0; t < S_; ++t) {
      float s0 = 0.f, s1 = 0.f;       // two partial sums: halves the dependent FMA chain
#pragma unroll
      for (int i = 0; i < HALF; ++i) {
        s0 = fmaf(q[t][i], k[i], s0);
        s1 = fmaf(q[t][i + HALF], k[i + HALF], s1);
      }
      sc[t][j] = (j < S_ && !valid[j]) ? -INFINITY : s0 + s1;
    }
  }
  // ---- softmax over the S+1 keys (the bias key is never masked, so the maximum is finite)
#pragma unroll
  for (int t = 0; t < S_; ++t) {
    float m = sc[t][S_];
#pragma unroll
    for (int j = 0; j < S_; ++j) m = fmaxf(m, sc[t][j]);
    float l = 0.f;
#pragma unroll
    for (int j = 0; j <= S_; ++j) {
      sc[t][j] = __expf(sc[t][j] - m);
      l += sc[t][j];
    }
    const float inv = 1.0f / l;
#pragma unroll
    for (int j = 0; j <= S_; ++j) sc[t][j] *= inv;
  }
  // ---- P.V (q's registers are dead: the accumulators take their place)
#pragma unroll
  for (int t = 0; t < S_; ++t)
#pragma unroll
    for (int i = 0; i < HD; ++i) q[t][i] = 0.f;
#pragma unroll
  for (int j = 0; j <= S_; ++j) {
    float v[HD];
    if (j < S_) {
      load24(rows[j] + 2 * C, v);
    } else {
#pragma unroll
      for (int i = 0; i < HD; ++i) v[i] = __ldg(p.bias_v + h * HD + i);
    }
#pragma unroll
    for (int t = 0; t < S_; ++t)
#pragma unroll
      for (int i = 0; i < HD; ++i) q[t][i] = fmaf(sc[t][j], v[i], q[t][i]);
  }
#pragma unroll
  for (int t = 0; t < S_; ++t) store24(p.out + attn_row(nbi, t, S_, p.inner) * p.ldo + h * HD, q[t]);
}

// ---- residue-axis attention at S = 4 on contiguous sequences (inner == 1, ld == 3C): the block's rows are ONE
// contiguous byte range, so a single bulk copy (cp.async.bulk, mbarrier completion) brings q | k | v of ASB_SEQ
// sequences into shared memory at full memory-level parallelism.  K is rotated in place (bf16, as the tensor-core
// kernel stages it), then thread = (sequence, query, head) works out of shared memory (48-byte slices at a 48-byte
// lane stride: conflict-free 128-bit LDS, the two queries of a warp share their K / V reads) and writes one
// coalesced output row slice.  64 registers -> 40 warps per SM; attention_seq_kernel (thread = (sequence, head),
// three dependent global-load phases at ~11 warps per SM) took 36 us for the 98 MB of config 3.
__device__ __forceinline__ void load24_smem(const bf16* p, float* o) {
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const uint4 v = reinterpret_cast<const uint4*>(p)[i];
    const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = __bfloat1622float2(h2[j]);
      o[8 * i + 2 * j] = f.x; o[8 * i + 2 * j + 1] = f.y;
    }
  }
}
__device__ __forceinline__ void rope24s(float* x, const float* cs, const float* sn, int pos) {   // tables in shared memory
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const float4 c4 = reinterpret_cast<const float4*>(cs + pos * 12)[j];
    const float4 s4 = reinterpret_cast<const float4*>(sn + pos * 12)[j];
    const float cc[4] = {c4.x, c4.y, c4.z, c4.w}, ss[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int i = 4 * j + e;
      const float a = x[i], b = x[i + 12];
      x[i] = a * cc[e] - b * ss[e];
      x[i + 12] = b * cc[e] + a * ss[e];
    }
  }
}
// dot product of 24 fp32 values with 24 bf16 values in shared memory
__device__ __forceinline__ float dot24_smem(const float* q, const bf16* k) {
  float s0 = 0.f, s1 = 0.f;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const uint4 v = reinterpret_cast<const uint4*>(k)[i];
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      s0 = fmaf(q[8 * i + 2 * j], __uint_as_float(w[j] << 16), s0);
      s1 = fmaf(q[8 * i + 2 * j + 1], __uint_as_float(w[j] & 0xffff0000u), s1);
    }
  }
  return s0 + s1;
}
// acc[0..23] += pr * v[0..23] (bf16 values in shared memory)
__device__ __forceinline__ void axpy24_smem(float* acc, float pr, const bf16* vp) {
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const uint4 v = reinterpret_cast<const uint4*>(vp)[i];
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      acc[8 * i + 2 * j] = fmaf(pr, __uint_as_float(w[j] << 16), acc[8 * i + 2 * j]);
      acc[8 * i + 2 * j + 1] = fmaf(pr, __uint_as_float(w[j] & 0xffff0000u), acc[8 * i + 2 * j + 1]);
    }
  }
}
template <int HD, int ASB_SEQ>   // ASB_SEQ sequences per block; thread = (sequence, query, head), head fastest
__global__ void __launch_bounds__(ASB_SEQ * 64)
attention_seq4_bulk_kernel(AttnParams<bf16> p) {
  constexpr int S_ = 4, HALF = HD / 2, ASB_THREADS = ASB_SEQ * 64;
  static_assert(HD == 24, "rope24 is written for head_dim 24");
  extern __shared__ __align__(128) uint8_t asb_smem[];
  const int C = p.H * HD, ld = 3 * C;
  bf16* tile = reinterpret_cast<bf16*>(asb_smem);                       // [ASB_SEQ * 4 rows][3C]
  uint64_t* bar = reinterpret_cast<uint64_t*>(asb_smem + (size_t)ASB_SEQ * S_ * ld * sizeof(bf16));
  float* rope_cs = reinterpret_cast<float*>(bar + 2);                     // [5][12] cos, [5][12] sin
  float* rope_sn = rope_cs + (S_ + 1) * HALF;
  float* bias_kr = rope_sn + (S_ + 1) * HALF;                             // [H][24] bias key rotated to position S
  float* bias_vs = bias_kr + 16 * HD;                                     // [H][24]
  const int tid = threadIdx.x;
  const int seq0 = blockIdx.x * ASB_SEQ;
  const int nseq = min(ASB_SEQ, p.nb - seq0);
  const uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(bar);
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const uint32_t bytes = (uint32_t)nseq * S_ * ld * (uint32_t)sizeof(bf16);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(tile)),
                 "l"(p.qkv + (size_t)seq0 * S_ * ld), "r"(bytes), "r"(bar_a)
                 : "memory");
  }
  for (int i = tid; i < (S_ + 1) * HALF; i += ASB_THREADS) {
    rope_cs[i] = p.cs ? p.cs[i] : 1.f;
    rope_sn[i] = p.cs ? p.sn[i] : 0.f;
  }
  for (int i = tid; i < 16 * HD; i += ASB_THREADS) bias_vs[i] = p.bias_v[i];
  __syncthreads();                                                       // barrier initialised, rope tables visible
  if (tid < 16) {                                                        // bias key of head tid, rotated by position S
    float k[HD];
#pragma unroll
    for (int i = 0; i < HD; ++i) k[i] = p.bias_k[tid * HD + i];
    rope24s(k, rope_cs, rope_sn, S_);
#pragma unroll
    for (int i = 0; i < HD; ++i) bias_kr[tid * HD + i] = k[i];
  }
  const int h = tid & 15, t = (tid >> 4) & 3, sl = tid >> 6;
  const int nbi = seq0 + sl;
  float valid[S_];
#pragma unroll
  for (int j = 0; j < S_; ++j) valid[j] = (sl < nseq) ? p.key_valid[((size_t)nbi * S_ + j) % p.mask_mod] : 0.f;
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "ASB_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], 0;\n\t"
      "@P1 bra ASB_DONE;\n\t"
      "bra ASB_WAIT;\n\t"
      "ASB_DONE:\n\t}" ::"r"(bar_a)
      : "memory");
  // ---- rotate K in place (bf16, like the tensor-core kernel's staged K): this thread owns row (sl, t), head h
  bf16* rows = tile + (size_t)sl * S_ * ld + h * HD;
  if (sl < nseq) {
    float k[HD];
    load24_smem(rows + t * ld + C, k);
    rope24s(k, rope_cs, rope_sn, t);
    store24(rows + t * ld + C, k);
  }
  __syncthreads();
  if (sl >= nseq) return;
  float q[HD];
  load24_smem(rows + t * ld, q);
  rope24s(q, rope_cs, rope_sn, t);
  // ---- scores: keys 0..3 from the tokens, key 4 = learned bias token (mha.py:265-268); fp32 softmax
  float sc[S_ + 1];
#pragma unroll
  for (int j = 0; j < S_; ++j) {
    const float d = dot24_smem(q, rows + j * ld + C);
    sc[j] = valid[j] != 0.f ? d : -INFINITY;
  }
  {
    float s0 = 0.f;
#pragma unroll
    for (int i = 0; i < HD; ++i) s0 = fmaf(q[i], bias_kr[h * HD + i], s0);
    sc[S_] = s0;
  }
  float m = sc[S_];
#pragma unroll
  for (int j = 0; j < S_; ++j) m = fmaxf(m, sc[j]);
  float l = 0.f;
#pragma unroll
  for (int j = 0; j <= S_; ++j) {
    sc[j] = __expf(sc[j] - m);
    l += sc[j];
  }
  const float inv = 1.0f / l;
  // ---- P.V (q's registers are dead: the accumulators take their place)
#pragma unroll
  for (int i = 0; i < HD; ++i) q[i] = sc[S_] * inv * bias_vs[h * HD + i];
#pragma unroll
  for (int j = 0; j < S_; ++j) axpy24_smem(q, sc[j] * inv, rows + j * ld + 2 * C);
  store24(p.out + ((size_t)nbi * S_ + t) * p.ldo + h * HD, q);
}
inline bool attention_seq4_bulk_supported(const AttnParams<bf16>& p) {
  return p.S == 4 && p.inner == 1 && p.H == 16 && p.ld == 3 * p.H * 24 && (((uintptr_t)p.qkv) & 15) == 0 && !g_sw.no_attn_bulk;
}
inline bool attention_seq4_bulk_supported(const AttnParams<float>&) { return false; }
template <int NSEQ>
inline cudaError_t launch_attention_seq4_bulk_n(const AttnParams<bf16>& p, cudaStream_t st) {
  const int smem = NSEQ * 4 * p.ld * (int)sizeof(bf16) + 16 + 2 * 5 * 12 * 4 + 2 * 16 * 24 * 4;
  static DevOnce once;
  const int dev = cur_device();
  if (once.need(dev)) {
    cudaError_t e = cudaFuncSetAttribute(attention_seq4_bulk_kernel<24, NSEQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    once.done(dev);
  }
  MD_VARIANT(V_ATTN_SEQ4);
  attention_seq4_bulk_kernel<24, NSEQ><<<(p.nb + NSEQ - 1) / NSEQ, NSEQ * 64, smem, st>>>(p);
  return cudaSuccess;
}
inline cudaError_t launch_attention_seq4_bulk(const AttnParams<bf16>& p, cudaStream_t st) {
  const int nseq = g_sw.asb_seq;
  return nseq == 1 ? launch_attention_seq4_bulk_n<1>(p, st) : nseq == 2 ? launch_attention_seq4_bulk_n<2>(p, st) : launch_attention_seq4_bulk_n<4>(p, st);
}
inline cudaError_t launch_attention_seq4_bulk(const AttnParams<float>&, cudaStream_t) { return cudaErrorNotSupported; }

template <typename AT, int S_>
inline void launch_attention_short(const AttnParams<AT>& p, cudaStream_t st) {
  const long long groups = (long long)p.nb * p.H;
  const int G = ASH_THREADS / S_;
  attention_short_kernel<AT, 24, S_><<<(unsigned)((groups + G - 1) / G), ASH_THREADS, 0, st>>>(p, groups);
}

template <typename AT>
inline int launch_attention_simt(const AttnParams<AT>& p, int head_dim, cudaStream_t st) {
  MD_CHECK(head_dim == 24, "attention_simt: head_dim must be 24");
  if (p.nb <= 0 || p.S <= 0) return 0;
  const bool is_short = (p.S <= 4 && p.S >= 2) || ((p.S == 8 || p.S == 16) && p.H <= 16);
  ProfScope ps(is_short ? PC_ATTN_SHORT : PC_ATTN_SIMT, st);
  if (attention_seq4_bulk_supported(p)) {
    MD_CUDA(launch_attention_seq4_bulk(p, st));
  } else if (p.S == 4 || p.S == 2 || p.S == 3) {
    const long long groups = (long long)p.nb * p.H;
    const unsigned grid = (unsigned)((groups + ASQ_THREADS - 1) / ASQ_THREADS);
    if (p.S == 4) attention_seq_kernel<AT, 24, 4><<<grid, ASQ_THREADS, 0, st>>>(p, groups);
    else if (p.S == 3) attention_seq_kernel<AT, 24, 3><<<grid, ASQ_THREADS, 0, st>>>(p, groups);
    else attention_seq_kernel<AT, 24, 2><<<grid, ASQ_THREADS, 0, st>>>(p, groups);
  }
  else if (is_short && p.S == 4) launch_attention_short<AT, 4>(p, st);
  else if (is_short && p.S == 8) launch_attention_short<AT, 8>(p, st);
  else if (is_short) launch_attention_short<AT, 16>(p, st);
  else {
    dim3 grid(p.nb * p.H, (p.S + AS_THREADS - 1) / AS_THREADS);
    attention_simt_kernel<AT, 24><<<grid, AS_THREADS, 0, st>>>(p);
  }
  MD_LAUNCHED();
  MD_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace mdgen
