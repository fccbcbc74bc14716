// backward_ipa.cuh -- adjoint of InvariantPointAttention with c_z = 0 (ipa.py:92-255) as computed by ipa.cuh:
//   P [M, 672] = [q | kv | q_pts | kv_pts]  ->  ipa_points_kernel (Rigid.apply)  ->  ipa_attention_kernel  ->  cat [M, 256].
// ipa_attention_bwd_kernel: one block per query row i, warp = head.  Recomputes the row's logits / softmax / o_pt, then
//   dcat -> (d o, d local points, d norm) -> d o_pt (global frame) -> d probs -> d logits -> dq, dk, dv (into dP),
//   d q-points (dGq), d k-points / v-points (dGkv), d head_weights.  Contributions to OTHER rows (keys / values) are
//   atomics.  ipa_points_bwd_kernel maps dGq / dGkv back through x_g = R x + t into the point columns of dP.
#pragma once
#include "common.cuh"
#include "ipa.cuh"

namespace mdgen {
namespace bwd {

__global__ void __launch_bounds__(128)
ipa_attention_bwd_kernel(const float* __restrict__ P, int ldp, IpaDims d, const float* __restrict__ Gq,
                         const float* __restrict__ Gkv, const float* __restrict__ rot, const float* __restrict__ trans,
                         int ntBL, const float* __restrict__ head_w, const float* __restrict__ mask, int mask_mod, int L,
                         const float* __restrict__ dcat, float inf, float eps, float* __restrict__ dP,
                         float* __restrict__ dGq, float* __restrict__ dGkv, float* __restrict__ dhead_w) {
  extern __shared__ float smem[];  // per warp: probs[L] | d2[L] | dprob[L] | q[32] | q-points[24] | o_pt / d o_pt[24]
  const int h = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x;
  const int g0 = (i / L) * L;
  float* lg = smem + (size_t)h * (3 * L + 80);
  float* d2s = lg + L;
  float* dps = d2s + L;
  float* qs = dps + L;
  float* qp = qs + 32;
  float* op = qp + 24;
  const int nq = d.H * d.Pq, nkv = d.H * (d.Pq + d.Pv);
  qs[lane] = P[(size_t)i * ldp + h * d.c + lane];
  if (lane < 24) qp[lane] = Gq[((size_t)i * nq + h * d.Pq) * 3 + lane];
  __syncwarp();
  const float hwk = sqrtf(1.0f / (3.0f * (d.Pq * 9.0f / 2.0f)));
  const float w = head_w[h];
  const float hw = log1pf(expf(w)) * hwk;
  const float qk_scale = sqrtf(1.0f / (3.0f * d.c));
  const float mi = mask[i % mask_mod];
  // ---- forward recompute: logits, softmax, o_pt
  float mx = -INFINITY;
  for (int j = lane; j < L; j += 32) {
    const int rj = g0 + j;
    const float* kr = P + (size_t)rj * ldp + d.off_kv() + h * 2 * d.c;
    float s = 0.f;
#pragma unroll 8
    for (int c = 0; c < 32; ++c) s = fmaf(qs[c], kr[c], s);
    const float* kp = Gkv + ((size_t)rj * nkv + h * (d.Pq + d.Pv)) * 3;
    float dd = 0.f;
#pragma unroll
    for (int e = 0; e < 24; ++e) {
      const float df = qp[e] - kp[e];
      dd = fmaf(df, df, dd);
    }
    const float a = s * qk_scale - 0.5f * hw * dd + inf * (mi * mask[rj % mask_mod] - 1.0f);
    lg[j] = a;
    d2s[j] = dd;
    mx = fmaxf(mx, a);
  }
  mx = warp_max(mx);
  float sum = 0.f;
  for (int j = lane; j < L; j += 32) {
    const float e = expf(lg[j] - mx);
    lg[j] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  __syncwarp();
  const float inv = 1.0f / sum;
  for (int j = lane; j < L; j += 32) lg[j] *= inv;      // probabilities
  __syncwarp();
  float opt = 0.f;
  for (int j = 0; j < L; ++j)
    if (lane < 24) opt = fmaf(lg[j], Gkv[((size_t)(g0 + j) * nkv + h * (d.Pq + d.Pv) + d.Pq) * 3 + lane], opt);
  if (lane < 24) op[lane] = opt;
  __syncwarp();
  // ---- d cat -> d o (lane = channel), d o_pt in the global frame (op[] reused)
  const float* dc = dcat + (size_t)i * d.cat_width();
  const float dO = dc[h * d.c + lane];
  float gx = 0.f, gy = 0.f, gz = 0.f;
  if (lane < 8) {
    const int fr = (i / ntBL) * mask_mod + (i % mask_mod);
    const float* R = rot + (size_t)fr * 9;
    const float* t = trans + (size_t)fr * 3;
    const float vx = op[lane * 3 + 0] - t[0], vy = op[lane * 3 + 1] - t[1], vz = op[lane * 3 + 2] - t[2];
    const float lx = R[0] * vx + R[3] * vy + R[6] * vz;
    const float ly = R[1] * vx + R[4] * vy + R[7] * vz;
    const float lz = R[2] * vx + R[5] * vy + R[8] * vz;
    const float nrm = sqrtf(lx * lx + ly * ly + lz * lz + eps);
    const int base = d.H * d.c, np = d.H * d.Pv, pi = h * d.Pv + lane;
    const float dn = dc[base + 3 * np + pi] / nrm;
    const float dlx = dc[base + pi] + dn * lx, dly = dc[base + np + pi] + dn * ly, dlz = dc[base + 2 * np + pi] + dn * lz;
    gx = R[0] * dlx + R[1] * dly + R[2] * dlz;           // local = R^T (g - t)  =>  d g = R d local
    gy = R[3] * dlx + R[4] * dly + R[5] * dlz;
    gz = R[6] * dlx + R[7] * dly + R[8] * dlz;
  }
  __syncwarp();
  if (lane < 8) {
    op[lane * 3 + 0] = gx;
    op[lane * 3 + 1] = gy;
    op[lane * 3 + 2] = gz;
  }
  __syncwarp();
  const float dopt = lane < 24 ? op[lane] : 0.f;
  // ---- d probs_j = dO . v_j + d o_pt . v_pts_j ;  D = sum_j p_j dprob_j
  float Dl = 0.f;
  for (int j = 0; j < L; ++j) {
    const int rj = g0 + j;
    float part = dO * P[(size_t)rj * ldp + d.off_kv() + h * 2 * d.c + d.c + lane];
    if (lane < 24) part = fmaf(dopt, Gkv[((size_t)rj * nkv + h * (d.Pq + d.Pv) + d.Pq) * 3 + lane], part);
    part = warp_sum(part);
    if (lane == 0) dps[j] = part;
    Dl = fmaf(lg[j], part, Dl);
  }
  __syncwarp();
  // ---- scatter
  float dq = 0.f, dqp = 0.f, dhw = 0.f;
  const float qv = qs[lane];
  for (int j = 0; j < L; ++j) {
    const int rj = g0 + j;
    const float pj = lg[j];
    const float ds = pj * (dps[j] - Dl);
    float* dPj = dP + (size_t)rj * ldp + d.off_kv() + h * 2 * d.c;
    atomicAdd(dPj + d.c + lane, pj * dO);                                          // d v_j
    const float kv = P[(size_t)rj * ldp + d.off_kv() + h * 2 * d.c + lane];
    dq = fmaf(ds * qk_scale, kv, dq);
    atomicAdd(dPj + lane, ds * qk_scale * qv);                                     // d k_j
    if (lane < 24) {
      float* dG = dGkv + ((size_t)rj * nkv + h * (d.Pq + d.Pv)) * 3;
      atomicAdd(dG + d.Pq * 3 + lane, pj * dopt);                                  // d v_pts_j
      const float df = qp[lane] - Gkv[((size_t)rj * nkv + h * (d.Pq + d.Pv)) * 3 + lane];
      const float gpt = -ds * hw * df;                                             // d/d q_pt of -0.5 hw |q_pt - k_pt|^2
      dqp += gpt;
      atomicAdd(dG + lane, -gpt);                                                  // d k_pts_j
    }
    if (lane == 0) dhw = fmaf(ds, -0.5f * d2s[j], dhw);
  }
  atomicAdd(dP + (size_t)i * ldp + h * d.c + lane, dq);
  if (lane < 24) atomicAdd(dGq + ((size_t)i * nq + h * d.Pq) * 3 + lane, dqp);
  if (lane == 0 && dhead_w) atomicAdd(dhead_w + h, dhw * hwk / (1.0f + expf(-w)));  // d softplus = sigmoid
}

// x_g = R x + t  =>  d x = R^T d x_g, scattered to the (xyz, point) layout of the projection columns (ipa.py:130-143)
__global__ void ipa_points_bwd_kernel(const float* __restrict__ dGq, const float* __restrict__ dGkv, IpaDims d,
                                      const float* __restrict__ rot, int M, int ntBL, int BL, float* __restrict__ dP, int ldp) {
  const int nq = d.H * d.Pq, nkv = d.H * (d.Pq + d.Pv);
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= M * (nq + nkv)) return;
  const int m = idx / (nq + nkv), pt = idx % (nq + nkv);
  const int fr = (m / ntBL) * BL + (m % BL);
  const float* R = rot + (size_t)fr * 9;
  const float* g = pt < nq ? dGq + ((size_t)m * nq + pt) * 3 : dGkv + ((size_t)m * nkv + (pt - nq)) * 3;
  const float x = R[0] * g[0] + R[3] * g[1] + R[6] * g[2];
  const float y = R[1] * g[0] + R[4] * g[1] + R[7] * g[2];
  const float z = R[2] * g[0] + R[5] * g[1] + R[8] * g[2];
  float* row = dP + (size_t)m * ldp;
  if (pt < nq) {
    row[d.off_qp() + pt] += x;
    row[d.off_qp() + nq + pt] += y;
    row[d.off_qp() + 2 * nq + pt] += z;
  } else {
    const int q = pt - nq;
    row[d.off_kvp() + q] += x;
    row[d.off_kvp() + nkv + q] += y;
    row[d.off_kvp() + 2 * nkv + q] += z;
  }
}

}  // namespace bwd
}  // namespace mdgen
