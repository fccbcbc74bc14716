// ipa.cuh -- InvariantPointAttention with c_z = 0 (src/mdgen/model/ipa.py:92-255), fp32 throughout
// (the reference keeps IPA geometry out of AMP: primitives.py:201-204, rigid_utils.py:28-30,70-71).
// Pipeline per IPA call:  proj GEMM  n[M,C] -> P[M, 672] = [q 128 | kv 256 | q_pts 96 | kv_pts 192]
//   -> ipa_points_kernel (Rigid.apply of the points, ipa.py:130-143)
//   -> ipa_attention_kernel (logits ipa.py:161-198, softmax :203, o / o_pt / inverse rigid / norm :209-234)
//   -> linear_out GEMM with residual epilogue (ipa.py:250-254, latent_model.py:382).
#pragma once
#include <math.h>

#include "common.cuh"

namespace mdgen {

struct IpaDims {
  int H, c, Pq, Pv;  // 4, 32, 8, 8
  __host__ __device__ int proj_width() const { return H * c * 3 + H * Pq * 3 + H * (Pq + Pv) * 3; }
  __host__ __device__ int off_kv() const { return H * c; }
  __host__ __device__ int off_qp() const { return H * c * 3; }
  __host__ __device__ int off_kvp() const { return H * c * 3 + H * Pq * 3; }
  __host__ __device__ int cat_width() const { return H * (c + 4 * Pv); }
};

// Gq[m][h*Pq+p][3], Gkv[m][h*(Pq+Pv)+j][3]: points mapped to the global frame, x_g = R x + t.
// Rows are (stream, step, b, l); frames are per (stream, b, l): frame row = (m / ntBL) * BL + m % BL.
__global__ void ipa_points_kernel(const float* __restrict__ P, int ldp, IpaDims d, const float* __restrict__ rot,
                                  const float* __restrict__ trans, int M, int ntBL, int BL, float* __restrict__ Gq,
                                  float* __restrict__ Gkv) {
  const int nq = d.H * d.Pq, nkv = d.H * (d.Pq + d.Pv);
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= M * (nq + nkv)) return;
  const int m = idx / (nq + nkv), pt = idx % (nq + nkv);
  const int fr = (m / ntBL) * BL + (m % BL);
  const float* R = rot + (size_t)fr * 9;
  const float* t = trans + (size_t)fr * 3;
  const float* row = P + (size_t)m * ldp;
  float x, y, z;
  float* dst;
  if (pt < nq) {  // raw layout (xyz, H*Pq): ipa.py:130-131
    x = row[d.off_qp() + pt];
    y = row[d.off_qp() + nq + pt];
    z = row[d.off_qp() + 2 * nq + pt];
    dst = Gq + ((size_t)m * nq + pt) * 3;
  } else {
    const int q = pt - nq;
    x = row[d.off_kvp() + q];
    y = row[d.off_kvp() + nkv + q];
    z = row[d.off_kvp() + 2 * nkv + q];
    dst = Gkv + ((size_t)m * nkv + q) * 3;
  }
  dst[0] = R[0] * x + R[1] * y + R[2] * z + t[0];  // rot_vec_mul, rigid_utils.py:64-86
  dst[1] = R[3] * x + R[4] * y + R[5] * z + t[1];
  dst[2] = R[6] * x + R[7] * y + R[8] * z + t[2];
}

// One block per query row i; warp h = head h (H warps).  Requires c == 32, Pq == Pv == 8.
// mask index = row % mask_mod (mask is [B,L]; rows are (stream, b, l)).
__global__ void __launch_bounds__(128)
ipa_attention_kernel(const float* __restrict__ P, int ldp, IpaDims d, const float* __restrict__ Gq,
                     const float* __restrict__ Gkv, const float* __restrict__ rot,
                     const float* __restrict__ trans, int ntBL, const float* __restrict__ head_w,
                     const float* __restrict__ mask, int mask_mod, int L, float* __restrict__ cat,
                     float inf, float eps) {
  extern __shared__ float smem[];  // per warp: L logits + 32 q + 24 q-points + 24 o_pt
  const int h = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int i = blockIdx.x;
  const int g0 = (i / L) * L;  // first row of this (stream, b) group
  float* lg = smem + (size_t)h * (L + 80);
  float* qs = lg + L;
  float* qp = qs + 32;
  float* op = qp + 24;
  const int nq = d.H * d.Pq, nkv = d.H * (d.Pq + d.Pv);
  qs[lane] = P[(size_t)i * ldp + h * d.c + lane];
  if (lane < 24) qp[lane] = Gq[((size_t)i * nq + h * d.Pq) * 3 + lane];
  __syncwarp();
  const float sp = log1pf(expf(head_w[h]));  // softplus (ipa.py:176)
  const float hw = sp * sqrtf(1.0f / (3.0f * (d.Pq * 9.0f / 2.0f)));
  const float qk_scale = sqrtf(1.0f / (3.0f * d.c));
  const float mi = mask[i % mask_mod];
  float mx = -INFINITY;
  for (int j = lane; j < L; j += 32) {
    const int rj = g0 + j;
    const float* kr = P + (size_t)rj * ldp + d.off_kv() + h * 2 * d.c;
    float s = 0.f;
#pragma unroll 8
    for (int c = 0; c < 32; ++c) s = fmaf(qs[c], kr[c], s);
    const float* kp = Gkv + ((size_t)rj * nkv + h * (d.Pq + d.Pv)) * 3;
    float d2 = 0.f;
#pragma unroll
    for (int e = 0; e < 24; ++e) {
      float df = qp[e] - kp[e];
      d2 = fmaf(df, df, d2);
    }
    float a = s * qk_scale - 0.5f * hw * d2 + inf * (mi * mask[rj % mask_mod] - 1.0f);
    lg[j] = a;
    mx = fmaxf(mx, a);
  }
  mx = warp_max(mx);
  float sum = 0.f;
  for (int j = lane; j < L; j += 32) {
    float e = expf(lg[j] - mx);
    lg[j] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  __syncwarp();
  const float inv = 1.0f / sum;
  // o[h, c = lane] and o_pt[h, p, x] (lane < 24 -> p = lane / 3, x = lane % 3)
  float o = 0.f, opt = 0.f;
  for (int j = 0; j < L; ++j) {
    const int rj = g0 + j;
    const float a = lg[j];
    o = fmaf(a, P[(size_t)rj * ldp + d.off_kv() + h * 2 * d.c + d.c + lane], o);
    if (lane < 24) opt = fmaf(a, Gkv[((size_t)rj * nkv + h * (d.Pq + d.Pv) + d.Pq) * 3 + lane], opt);
  }
  float* crow = cat + (size_t)i * d.cat_width();
  crow[h * d.c + lane] = o * inv;
  if (lane < 24) op[lane] = opt * inv;
  __syncwarp();
  if (lane < 8) {  // invert_apply (rigid_utils.py:1061-1073): R^T (x - t)
    const int fr = (i / ntBL) * mask_mod + (i % mask_mod);
    const float* R = rot + (size_t)fr * 9;
    const float* t = trans + (size_t)fr * 3;
    float vx = op[lane * 3 + 0] - t[0], vy = op[lane * 3 + 1] - t[1], vz = op[lane * 3 + 2] - t[2];
    float lx = R[0] * vx + R[3] * vy + R[6] * vz;
    float ly = R[1] * vx + R[4] * vy + R[7] * vz;
    float lz = R[2] * vx + R[5] * vy + R[8] * vz;
    const int base = d.H * d.c, np = d.H * d.Pv, pi = h * d.Pv + lane;
    crow[base + pi] = lx;
    crow[base + np + pi] = ly;
    crow[base + 2 * np + pi] = lz;
    crow[base + 3 * np + pi] = sqrtf(lx * lx + ly * ly + lz * lz + eps);  // ipa.py:229-231
  }
}

// Relative rigid 7-vectors of latent_model.py:198-199 with a closed-form quaternion (w >= 0):
//   f7 = to_tensor_7(start^-1 o end), r7 = to_tensor_7(end^-1 o start).
__device__ inline void rot_to_quat_closed(const float* R, float* q) {
  // Shepperd's method: pick the largest of (trace, R00, R11, R22) for conditioning.
  float xx = R[0], yy = R[4], zz = R[8];
  float tr = xx + yy + zz;
  float w, x, y, z;
  if (tr >= xx && tr >= yy && tr >= zz) {
    float s = sqrtf(fmaxf(1.f + tr, 0.f)) * 2.f;
    w = 0.25f * s; x = (R[7] - R[5]) / s; y = (R[2] - R[6]) / s; z = (R[3] - R[1]) / s;
  } else if (xx >= yy && xx >= zz) {
    float s = sqrtf(fmaxf(1.f + xx - yy - zz, 0.f)) * 2.f;
    w = (R[7] - R[5]) / s; x = 0.25f * s; y = (R[1] + R[3]) / s; z = (R[2] + R[6]) / s;
  } else if (yy >= zz) {
    float s = sqrtf(fmaxf(1.f + yy - xx - zz, 0.f)) * 2.f;
    w = (R[2] - R[6]) / s; x = (R[1] + R[3]) / s; y = 0.25f * s; z = (R[5] + R[7]) / s;
  } else {
    float s = sqrtf(fmaxf(1.f + zz - xx - yy, 0.f)) * 2.f;
    w = (R[3] - R[1]) / s; x = (R[2] + R[6]) / s; y = (R[5] + R[7]) / s; z = 0.25f * s;
  }
  float n = rsqrtf(w * w + x * x + y * y + z * z);
  if (w < 0.f) n = -n;
  q[0] = w * n; q[1] = x * n; q[2] = y * n; q[3] = z * n;
}

__device__ inline void rel7(const float* Ra, const float* ta, const float* Rb, const float* tb, float* out) {
  // a^-1 o b : rot = Ra^T Rb ; trans = Ra^T (tb - ta)      (rigid_utils.py:1031-1045,1075-1085)
  float R[9];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) R[r * 3 + c] = Ra[r] * Rb[c] + Ra[3 + r] * Rb[3 + c] + Ra[6 + r] * Rb[6 + c];
  rot_to_quat_closed(R, out);
  float dx = tb[0] - ta[0], dy = tb[1] - ta[1], dz = tb[2] - ta[2];
  out[4] = Ra[0] * dx + Ra[3] * dy + Ra[6] * dz;
  out[5] = Ra[1] * dx + Ra[4] * dy + Ra[7] * dz;
  out[6] = Ra[2] * dx + Ra[5] * dy + Ra[8] * dz;
}

__global__ void relative_frames_kernel(const float* __restrict__ srot, const float* __restrict__ strans,
                                       const float* __restrict__ erot, const float* __restrict__ etrans, int n,
                                       float* __restrict__ f7, float* __restrict__ r7) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  rel7(srot + (size_t)i * 9, strans + (size_t)i * 3, erot + (size_t)i * 9, etrans + (size_t)i * 3, f7 + (size_t)i * 7);
  rel7(erot + (size_t)i * 9, etrans + (size_t)i * 3, srot + (size_t)i * 9, strans + (size_t)i * 3, r7 + (size_t)i * 7);
}

// ---- prep_batch (wrapper.py:283-365): trajectory frames -> sampler latents + conditioning, one thread per (b, t, l).
//   offsets  = to_tensor_7(rigids[:, 0]^-1 o rigids), quaternion real part >= 0 (utils.py:7-14, wrapper.py:304-309)
//   offsets_r = the same against the LAST frame for tps / inpainting (:314-318)
//   latents  = [offsets | offsets_r | torsions(14)] (:320-327);  x_cond = cond_mask ? latents : 0 (:338-346, 361)
// The reference gets the quaternion from eigh and then fixes the sign; rot_to_quat_closed gives that unit quaternion
// directly.  flags: PREP_* below.  design_key_frames (:296-302) reads residues {0,0,3,3} in the first / last frame.
enum { PREP_TWO_REF = 1, PREP_NO_OFFSETS = 2, PREP_NO_TORSION = 4, PREP_NO_DESIGN_TORSION = 8, PREP_KEY_FRAMES = 16,
       PREP_SIM_COND = 32, PREP_TPS_COND = 64, PREP_INPAINT_COND = 128 };
__global__ void prep_batch_kernel(const float* __restrict__ trans, const float* __restrict__ rots,
                                  const float* __restrict__ torsions, int B, int T, int L, int flags, int cond_interval,
                                  float* __restrict__ latents, float* __restrict__ x_cond, long long* __restrict__ cond_mask) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * T * L) return;
  const int l = i % L, t = (i / L) % T, b = i / (L * T);
  const bool two = flags & PREP_TWO_REF;
  const int D = (two ? 14 : 7) + 14;
  auto frame = [&](int tt, int ll) -> size_t {       // design_key_frames: the key frames borrow residues 0 / 3
    if ((flags & PREP_KEY_FRAMES) && (tt == 0 || tt == T - 1)) ll = ll < 2 ? 0 : 3;
    return ((size_t)b * T + tt) * L + ll;
  };
  const size_t me = frame(t, l);
  float out[28];
  const float I3[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, Z3[3] = {0, 0, 0};
  if (flags & PREP_NO_OFFSETS) {
    rel7(I3, Z3, rots + me * 9, trans + me * 3, out);
  } else {
    const size_t r0 = frame(0, l);
    rel7(rots + r0 * 9, trans + r0 * 3, rots + me * 9, trans + me * 3, out);
  }
  int n = 7;
  if (two) {
    const size_t r1 = frame(T - 1, l);
    rel7(rots + r1 * 9, trans + r1 * 3, rots + me * 9, trans + me * 3, out + 7);
    n = 14;
  }
  const bool zero_tors = (flags & PREP_NO_TORSION) || ((flags & PREP_NO_DESIGN_TORSION) && (l == 1 || l == 2));
#pragma unroll
  for (int k = 0; k < 14; ++k) out[n + k] = zero_tors ? 0.f : torsions[(size_t)i * 14 + k];
  bool cond = false;
  if ((flags & PREP_SIM_COND) && t == 0) cond = true;
  if ((flags & PREP_TPS_COND) && (t == 0 || t == T - 1)) cond = true;
  if (cond_interval > 0 && t % cond_interval == 0) cond = true;
  if ((flags & PREP_INPAINT_COND) && (l == 0 || l == 3)) cond = true;
  cond_mask[i] = cond ? 1 : 0;
  for (int k = 0; k < D; ++k) {
    latents[(size_t)i * D + k] = out[k];
    x_cond[(size_t)i * D + k] = cond ? out[k] : 0.f;
  }
}

// ---- decode (wrapper.py:487-508): sampled latent -> atom14, one thread per (b, t, l).
//   frame   = rigids[:, 0] o Rigid.from_tensor_7(latent[:7], normalize_quats=True)        (:497-500)
//   torsion frames: default frame of the residue type o rotation about x by the (normalised) torsion, chi2..chi4
//   chained onto chi1 (geometry.py:273-334); atom = frame of its rigid group applied to the literature position,
//   masked (geometry.py:236-270).  The four residue tables are the caller's (mdgen.residue_constants).
enum { DEC_NO_OFFSETS = 1, DEC_RAW_TORSIONS = 2, DEC_DESIGN = 4 };
struct Rig { float R[9]; float t[3]; };
__device__ inline void rig_compose(const Rig& a, const Rig& b, Rig& o) {      // a o b  (rigid_utils.py:1031-1045)
#pragma unroll
  for (int r = 0; r < 3; ++r) {
#pragma unroll
    for (int c = 0; c < 3; ++c) o.R[r * 3 + c] = a.R[r * 3] * b.R[c] + a.R[r * 3 + 1] * b.R[3 + c] + a.R[r * 3 + 2] * b.R[6 + c];
    o.t[r] = a.R[r * 3] * b.t[0] + a.R[r * 3 + 1] * b.t[1] + a.R[r * 3 + 2] * b.t[2] + a.t[r];
  }
}
__global__ void decode_atom14_kernel(const float* __restrict__ samples, int D, int tors_off, const float* __restrict__ rot0,
                                     const float* __restrict__ trans0, const long long* __restrict__ seqres, int B, int T, int L,
                                     const float* __restrict__ default_frames, const long long* __restrict__ group_idx,
                                     const float* __restrict__ atom_mask, const float* __restrict__ lit_positions, int flags,
                                     float* __restrict__ atom14, long long* __restrict__ aa_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * T * L) return;
  const int l = i % L, b = i / (L * T);
  const float* x = samples + (size_t)i * D;
  const int aa = (int)seqres[b * L + l];
  // ---- backbone frame
  Rig off, fr;
  {
    float w = x[0], qx = x[1], qy = x[2], qz = x[3];
    const float n = 1.0f / sqrtf(w * w + qx * qx + qy * qy + qz * qz);
    w *= n; qx *= n; qy *= n; qz *= n;                                  // rigid_utils.quat_to_rot (:129-165)
    off.R[0] = w * w + qx * qx - qy * qy - qz * qz; off.R[1] = 2 * (qx * qy - w * qz); off.R[2] = 2 * (qx * qz + w * qy);
    off.R[3] = 2 * (qx * qy + w * qz); off.R[4] = w * w - qx * qx + qy * qy - qz * qz; off.R[5] = 2 * (qy * qz - w * qx);
    off.R[6] = 2 * (qx * qz - w * qy); off.R[7] = 2 * (qy * qz + w * qx); off.R[8] = w * w - qx * qx - qy * qy + qz * qz;
    off.t[0] = x[4]; off.t[1] = x[5]; off.t[2] = x[6];
  }
  if (flags & DEC_NO_OFFSETS) {
    fr = off;
  } else {
    Rig f0;
#pragma unroll
    for (int k = 0; k < 9; ++k) f0.R[k] = rot0[(size_t)(b * L + l) * 9 + k];
#pragma unroll
    for (int k = 0; k < 3; ++k) f0.t[k] = trans0[(size_t)(b * L + l) * 3 + k];
    rig_compose(f0, off, fr);
  }
  // ---- the 8 rigid-group frames in the backbone frame, then global
  Rig grp[8];
  for (int g = 0; g < 8; ++g) {
    const float* d4 = default_frames + ((size_t)aa * 8 + g) * 16;
    float s = 0.f, c = 1.f;                                             // backbone group: no rotation
    if (g > 0) {
      s = x[tors_off + 2 * (g - 1)];
      c = x[tors_off + 2 * (g - 1) + 1];
      if (!(flags & DEC_RAW_TORSIONS)) {
        const float n = sqrtf(s * s + c * c);
        s /= n; c /= n;
      }
    }
    Rig f;                                                              // default o Rx(torsion): [[1,0,0],[0,c,-s],[0,s,c]]
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const float d0 = d4[r * 4], d1 = d4[r * 4 + 1], d2 = d4[r * 4 + 2];
      f.R[r * 3] = d0;
      f.R[r * 3 + 1] = d1 * c + d2 * s;
      f.R[r * 3 + 2] = -d1 * s + d2 * c;
      f.t[r] = d4[r * 4 + 3];
    }
    if (g >= 5) rig_compose(grp[g - 1], f, grp[g]);                     // chi2..chi4 chained onto chi1
    else grp[g] = f;
  }
  for (int a = 0; a < 14; ++a) {
    const int g = (int)group_idx[aa * 14 + a];
    Rig G;
    rig_compose(fr, grp[g], G);
    const float* lp = lit_positions + ((size_t)aa * 14 + a) * 3;
    const float m = atom_mask[aa * 14 + a];
    float* o = atom14 + ((size_t)i * 14 + a) * 3;
#pragma unroll
    for (int r = 0; r < 3; ++r) o[r] = (G.R[r * 3] * lp[0] + G.R[r * 3 + 1] * lp[1] + G.R[r * 3 + 2] * lp[2] + G.t[r]) * m;
  }
  if (aa_out) {
    long long best = aa;
    if (flags & DEC_DESIGN) {                                           // torch.argmax over the 20 logits (first maximum)
      best = 0;
      float bv = x[D - 20];
      for (int k = 1; k < 20; ++k)
        if (x[D - 20 + k] > bv) { bv = x[D - 20 + k]; best = k; }
    }
    aa_out[i] = best;
  }
}

}  // namespace mdgen
