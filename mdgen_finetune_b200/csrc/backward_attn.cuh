// backward_attn.cuh -- adjoint of the MultiheadAttention core (mha.py:258-397 as reached from AttentionWithRoPE):
// bias-KV token as key S, RoPE on q / k (rotate_half pairing), key-padding mask, fp32 softmax, P.V.
// Two passes, no atomics except on the bias token that all sequences share:
//   attention_bwd_q_kernel : thread = query.  Recomputes the row's softmax statistics (max m, sum l), D = dO . O, and
//                            dq = sum_j ds_ij k_j  with ds_ij = p_ij (dO_i . v_j - D_i); un-rotates dq; stores (m, l, D).
//   attention_bwd_kv_kernel: thread = key.  dk_j = sum_i ds_ij q_i (un-rotated at position j), dv_j = sum_i p_ij dO_i;
//                            the thread of key S adds its result to dbias_k / dbias_v.
#pragma once
#include "attention_simt.cuh"
#include "common.cuh"

namespace mdgen {
namespace bwd {

constexpr int AB_THREADS = 128, AB_CHUNK = 64;

template <int HD>
__device__ __forceinline__ void rope_fwd(float* v, const float* cs, const float* sn, int pos) {
  constexpr int HALF = HD / 2;
#pragma unroll
  for (int i = 0; i < HALF; ++i) {
    const float c = cs[pos * HALF + i], s = sn[pos * HALF + i];
    const float a = v[i], b = v[i + HALF];
    v[i] = a * c - b * s;
    v[i + HALF] = b * c + a * s;
  }
}
// adjoint of rope_fwd = rotation by the opposite angle
template <int HD>
__device__ __forceinline__ void rope_bwd(float* g, const float* cs, const float* sn, int pos) {
  constexpr int HALF = HD / 2;
#pragma unroll
  for (int i = 0; i < HALF; ++i) {
    const float c = cs[pos * HALF + i], s = sn[pos * HALF + i];
    const float ga = g[i], gb = g[i + HALF];
    g[i] = ga * c + gb * s;
    g[i + HALF] = gb * c - ga * s;
  }
}

// key `key` (0..S; S = the bias token) of (sequence nbi, head h): rotated k, v, validity
template <int HD>
__device__ __forceinline__ void load_key(const AttnParams<float>& p, int nbi, int h, int key, float* k, float* v, float* valid) {
  const int C = p.H * HD;
  if (key < p.S) {
    const size_t row = attn_row(nbi, key, p.S, p.inner);
    const float* kr = p.qkv + row * p.ld + C + h * HD;
#pragma unroll
    for (int i = 0; i < HD; ++i) {
      k[i] = kr[i];
      v[i] = kr[C + i];
    }
    *valid = p.key_valid[row % p.mask_mod];
  } else {
#pragma unroll
    for (int i = 0; i < HD; ++i) {
      k[i] = p.bias_k[h * HD + i];
      v[i] = p.bias_v[h * HD + i];
    }
    *valid = 1.f;
  }
  if (p.cs) rope_fwd<HD>(k, p.cs, p.sn, key);
}

template <int HD>
__global__ void __launch_bounds__(AB_THREADS)
attention_bwd_q_kernel(AttnParams<float> p, const float* __restrict__ out, const float* __restrict__ dout, int ldo,
                       float* __restrict__ dqkv, float* __restrict__ stats) {
  __shared__ float Ks[AB_CHUNK][HD + 1], Vs[AB_CHUNK][HD + 1], valid[AB_CHUNK];
  const int nbi = blockIdx.x / p.H, h = blockIdx.x % p.H;
  const int qi = blockIdx.y * AB_THREADS + threadIdx.x;
  const bool active = qi < p.S;
  const int nkeys = p.S + 1;
  float q[HD], dO[HD], dq[HD];
  float D = 0.f;
  size_t row = 0;
  if (active) {
    row = attn_row(nbi, qi, p.S, p.inner);
    const float* qr = p.qkv + row * p.ld + h * HD;
#pragma unroll
    for (int i = 0; i < HD; ++i) {
      q[i] = qr[i];
      dO[i] = dout[row * ldo + h * HD + i];
      D = fmaf(dO[i], out[row * ldo + h * HD + i], D);
      dq[i] = 0.f;
    }
    if (p.cs) rope_fwd<HD>(q, p.cs, p.sn, qi);
  }
  float m = -INFINITY, l = 0.f;
  for (int pass = 0; pass < 2; ++pass) {
    for (int k0 = 0; k0 < nkeys; k0 += AB_CHUNK) {
      __syncthreads();
      for (int kk = threadIdx.x; kk < AB_CHUNK; kk += AB_THREADS) {
        const int key = k0 + kk;
        float kt[HD], vt[HD], vv = 0.f;
        if (key < nkeys) load_key<HD>(p, nbi, h, key, kt, vt, &vv);
#pragma unroll
        for (int i = 0; i < HD; ++i) {
          Ks[kk][i] = key < nkeys ? kt[i] : 0.f;
          Vs[kk][i] = key < nkeys ? vt[i] : 0.f;
        }
        valid[kk] = vv;
      }
      __syncthreads();
      if (!active) continue;
      const int kn = min(AB_CHUNK, nkeys - k0);
      for (int kk = 0; kk < kn; ++kk) {
        if (valid[kk] == 0.f) continue;
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < HD; ++i) s = fmaf(q[i], Ks[kk][i], s);
        if (pass == 0) {
          const float mn = fmaxf(m, s);
          l = l * expf(m - mn) + expf(s - mn);
          m = mn;
        } else {
          const float pj = expf(s - m) / l;
          float dp = 0.f;
#pragma unroll
          for (int i = 0; i < HD; ++i) dp = fmaf(dO[i], Vs[kk][i], dp);
          const float ds = pj * (dp - D);
#pragma unroll
          for (int i = 0; i < HD; ++i) dq[i] = fmaf(ds, Ks[kk][i], dq[i]);
        }
      }
    }
  }
  if (!active) return;
  if (p.cs) rope_bwd<HD>(dq, p.cs, p.sn, qi);
  float* g = dqkv + row * p.ld + h * HD;
#pragma unroll
  for (int i = 0; i < HD; ++i) g[i] += dq[i];
  float* st = stats + (row * p.H + h) * 3;
  st[0] = m;
  st[1] = l;
  st[2] = D;
}

template <int HD>
__global__ void __launch_bounds__(AB_THREADS)
attention_bwd_kv_kernel(AttnParams<float> p, const float* __restrict__ dout, int ldo, const float* __restrict__ stats,
                        float* __restrict__ dqkv, float* __restrict__ dbias_k, float* __restrict__ dbias_v) {
  __shared__ float Qs[AB_CHUNK][HD + 1], Os[AB_CHUNK][HD + 1], St[AB_CHUNK][3];
  const int nbi = blockIdx.x / p.H, h = blockIdx.x % p.H;
  const int C = p.H * HD;
  const int key = blockIdx.y * AB_THREADS + threadIdx.x;
  const bool active = key <= p.S;
  float k[HD], v[HD], dk[HD], dv[HD], kvalid = 0.f;
  if (active) load_key<HD>(p, nbi, h, key, k, v, &kvalid);
#pragma unroll
  for (int i = 0; i < HD; ++i) dk[i] = dv[i] = 0.f;
  for (int q0 = 0; q0 < p.S; q0 += AB_CHUNK) {
    __syncthreads();
    for (int qq = threadIdx.x; qq < AB_CHUNK; qq += AB_THREADS) {
      const int qi = q0 + qq;
      if (qi < p.S) {
        const size_t row = attn_row(nbi, qi, p.S, p.inner);
        float qt[HD];
#pragma unroll
        for (int i = 0; i < HD; ++i) {
          qt[i] = p.qkv[row * p.ld + h * HD + i];
          Os[qq][i] = dout[row * ldo + h * HD + i];
        }
        if (p.cs) rope_fwd<HD>(qt, p.cs, p.sn, qi);
#pragma unroll
        for (int i = 0; i < HD; ++i) Qs[qq][i] = qt[i];
        const float* st = stats + (row * p.H + h) * 3;
        St[qq][0] = st[0];
        St[qq][1] = st[1];
        St[qq][2] = st[2];
      }
    }
    __syncthreads();
    if (!active || kvalid == 0.f) continue;
    const int qn = min(AB_CHUNK, p.S - q0);
    for (int qq = 0; qq < qn; ++qq) {
      float s = 0.f, dp = 0.f;
#pragma unroll
      for (int i = 0; i < HD; ++i) {
        s = fmaf(Qs[qq][i], k[i], s);
        dp = fmaf(Os[qq][i], v[i], dp);
      }
      const float pj = expf(s - St[qq][0]) / St[qq][1];
      const float ds = pj * (dp - St[qq][2]);
#pragma unroll
      for (int i = 0; i < HD; ++i) {
        dk[i] = fmaf(ds, Qs[qq][i], dk[i]);
        dv[i] = fmaf(pj, Os[qq][i], dv[i]);
      }
    }
  }
  if (!active || kvalid == 0.f) return;
  if (p.cs) rope_bwd<HD>(dk, p.cs, p.sn, key);
  if (key < p.S) {
    float* g = dqkv + attn_row(nbi, key, p.S, p.inner) * p.ld + C + h * HD;
#pragma unroll
    for (int i = 0; i < HD; ++i) {
      g[i] += dk[i];
      g[C + i] += dv[i];
    }
  } else {      // the learned bias token (mha.py:265-268) is shared by every sequence
#pragma unroll
    for (int i = 0; i < HD; ++i) {
      if (dbias_k) atomicAdd(dbias_k + h * HD + i, dk[i]);
      if (dbias_v) atomicAdd(dbias_v + h * HD + i, dv[i]);
    }
  }
}

// stats: [rows * H * 3] scratch.  dqkv [rows, p.ld] accumulates; dbias_k / dbias_v [C] accumulate (may be null).
inline int attention_backward(const AttnParams<float>& p, const float* out, const float* dout, int ldo, float* dqkv,
                              float* stats, float* dbias_k, float* dbias_v, cudaStream_t st) {
  if (p.nb <= 0 || p.S <= 0) return 0;
  dim3 gq(p.nb * p.H, (p.S + AB_THREADS - 1) / AB_THREADS), gk(p.nb * p.H, (p.S + 1 + AB_THREADS - 1) / AB_THREADS);
  attention_bwd_q_kernel<24><<<gq, AB_THREADS, 0, st>>>(p, out, dout, ldo, dqkv, stats);
  MD_LAUNCHED();
  attention_bwd_kv_kernel<24><<<gk, AB_THREADS, 0, st>>>(p, dout, ldo, stats, dqkv, dbias_k, dbias_v);
  MD_LAUNCHED();
  MD_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace bwd
}  // namespace mdgen
