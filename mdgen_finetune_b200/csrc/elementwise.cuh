// elementwise.cuh -- HBM-bound glue kernels: LayerNorm+adaLN modulate, token embedding,
// timestep embedding, adaLN table, RoPE tables.  One warp per 384-wide row, float4 lanes.
#pragma once
#include "common.cuh"

namespace mdgen {

// ---- LayerNorm (no affine, eps) + modulate ------------------------------------------
// out[m,:] = LN(x[m,:]) * (1 + scale[b]) + shift[b],  b = (m / rows_per_batch) % nbatch.
// latent_model.py:385,390,467,475,489 + layers.py:14-15,73.  If shift == nullptr: affine
// LayerNorm with weight/bias (ipa_norm, latent_model.py:360,382; eps 1e-5).
// LN_ROWS consecutive rows per warp: all their loads are issued before the first reduction (memory-level parallelism
// per warp), and the batch-index division plus the shift / scale loads are shared when the rows fall in one batch element.
constexpr int LN_ROWS = 2;
template <typename OT, int C>
__global__ void __launch_bounds__(256)
ln_modulate_kernel(const float* __restrict__ x, OT* __restrict__ out, int M, const float* __restrict__ shift,
                   const float* __restrict__ scale, int mod_stride, int rows_per_batch, int nbatch,
                   const float* __restrict__ w, const float* __restrict__ bvec, float eps) {
  static_assert(C % 128 == 0, "C must be a multiple of 128");
  constexpr int V = C / 128;  // float4 per lane
  pdl_trigger();       // common.cuh: programmatic dependent launch
  pdl_wait();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int row0 = warp * LN_ROWS;
  if (row0 >= M) return;
  float4 v[LN_ROWS][V];
#pragma unroll
  for (int r = 0; r < LN_ROWS; ++r) {
    const int row = min(row0 + r, M - 1);
    const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * C);
#pragma unroll
    for (int i = 0; i < V; ++i) v[r][i] = xr[lane + 32 * i];
  }
  // per-channel multiplier / offset of the first row's batch element (or the affine LayerNorm weights)
  const int q0 = row0 / rows_per_batch;
  const int b0 = q0 % nbatch;
  const int split = (q0 + 1) * rows_per_batch - row0;      // rows >= split belong to the next batch element
  float4 mul[V], add[V];
#pragma unroll
  for (int i = 0; i < V; ++i) {
    const int c0 = (lane + 32 * i) * 4;
    if (shift) {
      const float4 sc = *reinterpret_cast<const float4*>(scale + (size_t)b0 * mod_stride + c0);
      mul[i] = make_float4(1.f + sc.x, 1.f + sc.y, 1.f + sc.z, 1.f + sc.w);
      add[i] = *reinterpret_cast<const float4*>(shift + (size_t)b0 * mod_stride + c0);
    } else if (w) {
      mul[i] = *reinterpret_cast<const float4*>(w + c0);
      add[i] = *reinterpret_cast<const float4*>(bvec + c0);
    } else {
      mul[i] = make_float4(1.f, 1.f, 1.f, 1.f);
      add[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  float mean[LN_ROWS], rstd[LN_ROWS];
#pragma unroll
  for (int r = 0; r < LN_ROWS; ++r) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i) s += (v[r][i].x + v[r][i].y) + (v[r][i].z + v[r][i].w);
    mean[r] = s;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
#pragma unroll
    for (int r = 0; r < LN_ROWS; ++r) mean[r] += __shfl_xor_sync(0xffffffffu, mean[r], o);
#pragma unroll
  for (int r = 0; r < LN_ROWS; ++r) {
    mean[r] *= (1.0f / C);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const float a = v[r][i].x - mean[r], b = v[r][i].y - mean[r], c = v[r][i].z - mean[r], d = v[r][i].w - mean[r];
      q += (a * a + b * b) + (c * c + d * d);
    }
    rstd[r] = q;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
#pragma unroll
    for (int r = 0; r < LN_ROWS; ++r) rstd[r] += __shfl_xor_sync(0xffffffffu, rstd[r], o);
#pragma unroll
  for (int r = 0; r < LN_ROWS; ++r) {
    const int row = row0 + r;
    if (row >= M) break;
    const float rs = rsqrtf(rstd[r] * (1.0f / C) + eps), mu = mean[r];
    const bool next = shift && r >= split;                  // crossed into the next batch element: reload its vectors
    const int b1 = (q0 + 1) % nbatch;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const int c0 = (lane + 32 * i) * 4;
      float4 m4 = mul[i], a4 = add[i];
      if (next) {
        const float4 sc = *reinterpret_cast<const float4*>(scale + (size_t)b1 * mod_stride + c0);
        m4 = make_float4(1.f + sc.x, 1.f + sc.y, 1.f + sc.z, 1.f + sc.w);
        a4 = *reinterpret_cast<const float4*>(shift + (size_t)b1 * mod_stride + c0);
      }
      const float o0 = fmaf((v[r][i].x - mu) * rs, m4.x, a4.x), o1 = fmaf((v[r][i].y - mu) * rs, m4.y, a4.y);
      const float o2 = fmaf((v[r][i].z - mu) * rs, m4.z, a4.z), o3 = fmaf((v[r][i].w - mu) * rs, m4.w, a4.w);
      OT* o = out + (size_t)row * C + c0;
      if constexpr (sizeof(OT) == 4) {
        *reinterpret_cast<float4*>(o) = make_float4(o0, o1, o2, o3);
      } else {
        const __nv_bfloat162 p0 = __floats2bfloat162_rn(o0, o1), p1 = __floats2bfloat162_rn(o2, o3);
        uint2 pk;
        pk.x = *reinterpret_cast<const uint32_t*>(&p0);
        pk.y = *reinterpret_cast<const uint32_t*>(&p1);
        *reinterpret_cast<uint2*>(o) = pk;
      }
    }
  }
}

// ---- token embedding -------------------------------------------------------------------
// emb[m,:] = W_x x[m] + W_c x_cond[m] + b_x + b_c + mask_emb[cm[m]] + pos[l] + time[t] + ipa[b,l]
// latent_model.py:242-255.  thread = channel (its weight column lives in registers, read coalesced
// from the transposed weight WT [D + Dc, C]); one block = EMB_TOK tokens staged in shared memory.
constexpr int EMB_TOK = 64;
template <int C, int DDP>
__global__ void __launch_bounds__(C)
embed_kernel(const float* __restrict__ x, int D, const float* __restrict__ xc, int Dc,
             const int64_t* __restrict__ cm, const float* __restrict__ WT, const float* __restrict__ bsum,
             const float* __restrict__ mask_emb, const float* __restrict__ pos, const float* __restrict__ time_emb,
             const float* __restrict__ ipa, float* __restrict__ out, int N, int T, int L) {
  __shared__ __align__(16) float sm[EMB_TOK][DDP];
  __shared__ int4 meta[EMB_TOK];   // per token: mask_emb row, pos row, time row, ipa row (element offsets / C)
  const int c = threadIdx.x;
  const int m0 = blockIdx.x * EMB_TOK;
  const int DD = D + Dc;
  float w[DDP];
#pragma unroll
  for (int d = 0; d < DDP; ++d) w[d] = d < DD ? WT[(size_t)d * C + c] : 0.f;
  for (int i = threadIdx.x; i < EMB_TOK * DDP; i += C) {
    const int tk = i / DDP, d = i % DDP, m = m0 + tk;
    float v = 0.f;
    if (m < N && d < DD) v = d < D ? x[(size_t)m * D + d] : xc[(size_t)m * Dc + (d - D)];
    sm[tk][d] = v;
  }
  if (threadIdx.x < EMB_TOK) {
    const int m = min(m0 + (int)threadIdx.x, N - 1);
    const int l = m % L, t = (m / L) % T, b = m / (L * T);
    meta[threadIdx.x] = make_int4(cm ? (int)cm[m] : 0, l, t, b * L + l);
  }
  __syncthreads();
  const float base = bsum[c];
  const int ntk = min(EMB_TOK, N - m0);
#pragma unroll 4
  for (int tk = 0; tk < ntk; ++tk) {
    const int4 mt = meta[tk];
    float add = 0.f;
    if (cm) add += __ldg(mask_emb + (size_t)mt.x * C + c);
    if (pos) add += __ldg(pos + (size_t)mt.y * C + c);
    if (time_emb) add += __ldg(time_emb + (size_t)mt.z * C + c);
    if (ipa) add += __ldg(ipa + (size_t)mt.w * C + c);
    float a0 = base, a1 = 0.f;
#pragma unroll
    for (int d = 0; d < DDP; d += 8) {
      const float4 xv = *reinterpret_cast<const float4*>(&sm[tk][d]);
      const float4 yv = *reinterpret_cast<const float4*>(&sm[tk][d + 4]);
      a0 = fmaf(w[d], xv.x, a0);
      a1 = fmaf(w[d + 4], yv.x, a1);
      a0 = fmaf(w[d + 1], xv.y, a0);
      a1 = fmaf(w[d + 5], yv.y, a1);
      a0 = fmaf(w[d + 2], xv.z, a0);
      a1 = fmaf(w[d + 6], yv.z, a1);
      a0 = fmaf(w[d + 3], xv.w, a0);
      a1 = fmaf(w[d + 7], yv.w, a1);
    }
    out[(size_t)(m0 + tk) * C + c] = (a0 + a1) + add;
  }
}

// Per-step part of the embedding: out[m,:] = stat[m,:] + ipa[b,l] + W_x x[m].  `stat` holds everything that does not
// change between sampler steps (W_c x_cond + biases + mask / position / time embeddings, written once per
// conditioning by embed_kernel with D = 0).  thread = CH channels (c, c + C/CH, ...) with their weight columns in
// registers, so one broadcast LDS.128 of the staged inputs feeds 4 * CH FMAs: the one-channel kernel above was
// LSU-bound (ncu: 78 % LSU, 68 us for config 3).
constexpr int EMB_STEP_TOK = 16;    // tokens per block: 2000 blocks on config 3, ~13 resident warps per SM stay busy
constexpr int EMB_STEP_GRP = 8;     // tokens whose loads are in flight together
// Round 2: the tokens of a block are processed four (then eight) at a time with every load of the group issued before the first FMA
// (the one-token loop had ~8 loads in flight per thread at ~15 warps per SM): 2.29 -> 2.07 ms per rollout on config 3.
// Eight at a time: 2.08 -> 1.90 ms (168 registers, no spills).
// Fusing the first block's LayerNorm + modulate in here (block reductions for the row statistics) was measured SLOWER than
// the separate ln_modulate launch (3.32 vs 3.01 ms per rollout for the two together) and is not kept.
template <int C, int DP, int CH>
__global__ void __launch_bounds__(C / CH)
embed_step_kernel(const float* __restrict__ x, int D, const float* __restrict__ WT, const float* __restrict__ stat,
                  const float* __restrict__ ipa, float* __restrict__ out, int N, int T, int L) {
  constexpr int NT = C / CH, G = EMB_STEP_GRP;
  static_assert(NT >= EMB_STEP_TOK && EMB_STEP_TOK % G == 0, "one thread per staged token for the row table");
  __shared__ __align__(16) float sm[EMB_STEP_TOK][DP];
  __shared__ int ipa_row[EMB_STEP_TOK];
  pdl_trigger();       // common.cuh: programmatic dependent launch
  pdl_wait();
  const int c0 = threadIdx.x;
  const int m0 = blockIdx.x * EMB_STEP_TOK;
  float w[CH][DP];
#pragma unroll
  for (int k = 0; k < CH; ++k)
#pragma unroll
    for (int d = 0; d < DP; ++d) w[k][d] = d < D ? WT[(size_t)d * C + c0 + k * NT] : 0.f;
  for (int i = threadIdx.x; i < EMB_STEP_TOK * DP; i += NT) {
    const int tk = i / DP, d = i % DP, m = m0 + tk;
    sm[tk][d] = (m < N && d < D) ? x[(size_t)m * D + d] : 0.f;
  }
  if (threadIdx.x < EMB_STEP_TOK) {
    const int m = min(m0 + (int)threadIdx.x, N - 1);
    ipa_row[threadIdx.x] = (m / (L * T)) * L + m % L;
  }
  __syncthreads();
#pragma unroll 1
  for (int g = 0; g < EMB_STEP_TOK; g += G) {
    if (m0 + g >= N) break;
    float a[G][CH];
#pragma unroll
    for (int tk = 0; tk < G; ++tk) {
      const size_t row = (size_t)min(m0 + g + tk, N - 1) * C;
#pragma unroll
      for (int k = 0; k < CH; ++k) a[tk][k] = __ldg(stat + row + c0 + k * NT);
    }
    if (ipa) {
#pragma unroll
      for (int tk = 0; tk < G; ++tk) {
        const size_t ir = (size_t)ipa_row[g + tk] * C;
#pragma unroll
        for (int k = 0; k < CH; ++k) a[tk][k] += __ldg(ipa + ir + c0 + k * NT);
      }
    }
#pragma unroll
    for (int tk = 0; tk < G; ++tk) {
#pragma unroll
      for (int d = 0; d < DP; d += 4) {
        const float4 xv = *reinterpret_cast<const float4*>(&sm[g + tk][d]);
#pragma unroll
        for (int k = 0; k < CH; ++k) {
          a[tk][k] = fmaf(w[k][d], xv.x, a[tk][k]);
          a[tk][k] = fmaf(w[k][d + 1], xv.y, a[tk][k]);
          a[tk][k] = fmaf(w[k][d + 2], xv.z, a[tk][k]);
          a[tk][k] = fmaf(w[k][d + 3], xv.w, a[tk][k]);
        }
      }
    }
#pragma unroll
    for (int tk = 0; tk < G; ++tk) {
      const int m = m0 + g + tk;
      if (m < N) {
#pragma unroll
        for (int k = 0; k < CH; ++k) out[(size_t)m * C + c0 + k * NT] = a[tk][k];
      }
    }
  }
}

// ---- design mode: Dirichlet conditional-flow velocity of the 20 sequence channels -----------------------------
// LatentMDGenModel.forward_inference (latent_model.py:279-326) with DirichletConditionalFlow.c_factor (utils.py:45-57)
// on the device.  thread = token.  In place on `latent` [N, D]: the last 20 channels hold the logits on entry and
//   flow_i = dalpha * sum_j softmax(logits / temp)_j * (delta_ij - x_i) * c_j        on exit,
//   c_j = -interp(x_j; linspace(0, 1, 1000), ifunc) * [x_j^(alpha-1) > 0] * [x_j < 1] * beta / ((1 - x_j)^19 * x_j^(alpha-1))
// with numpy's types: the powers in float32 (float32 array ** python scalar), the quotients and the interpolation in
// float64, the result cast to float32.  `ifunc` = the row of d/dalpha I_b(alpha, 19) nearest to alpha, `beta` =
// B(alpha, 19): both come from the host tables (scipy at model construction, like the reference).
// flags[0] += tokens off the simplex (|sum - 1| > 1e-4 or a negative entry: the reference prints a warning),
// flags[1] += NaN c-factors (RuntimeError in the reference unless allow_nan_cfactor, which maps them like nan_to_num).
constexpr int DFLOW_K = 20, DFLOW_GRID = 1000;
__global__ void dirichlet_flow_kernel(const float* __restrict__ x, float* __restrict__ latent, int N, int D,
                                      const double* __restrict__ ifunc, float alpha_m1, double beta, float temp, float dalpha,
                                      int allow_nan, int* __restrict__ flags) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= N) return;
  const float* xd = x + (size_t)m * D + (D - DFLOW_K);
  float* lg = latent + (size_t)m * D + (D - DFLOW_K);
  float xv[DFLOW_K], p[DFLOW_K];
  float sum = 0.f, mx = -INFINITY;
  bool neg = false;
#pragma unroll
  for (int j = 0; j < DFLOW_K; ++j) {
    xv[j] = xd[j];
    sum += xv[j];
    neg |= xv[j] < 0.f;
    p[j] = lg[j] / temp;
    mx = fmaxf(mx, p[j]);
  }
  if (neg || !(fabsf(sum - 1.0f) <= 1e-4f)) atomicAdd(&flags[0], 1);
  float ps = 0.f;
#pragma unroll
  for (int j = 0; j < DFLOW_K; ++j) {
    p[j] = expf(p[j] - mx);
    ps += p[j];
  }
  float pc = 0.f;          // sum_j p_j c_j
  float c[DFLOW_K];
  int nans = 0;
#pragma unroll
  for (int j = 0; j < DFLOW_K; ++j) {
    p[j] /= ps;
    const float b = xv[j];
    const float om = 1.0f - b;
    float om19 = om;                                   // (1 - b)^19 in float32
#pragma unroll
    for (int e = 1; e < DFLOW_K - 1; ++e) om19 *= om;
    const double beta_div = b < 1.0f ? beta / (double)om19 : 0.0;
    const float bp = powf(b, alpha_m1);                // NaN for b < 0: the comparison below is false like numpy's
    const double full = bp > 0.f ? beta_div / (double)bp : 0.0;
    // np.interp on the uniform grid (clamped at both ends)
    const double bd = (double)b;
    double iv;
    if (!(bd > 0.0)) iv = ifunc[0];
    else if (bd >= 1.0) iv = ifunc[DFLOW_GRID - 1];
    else {
      int k = (int)(bd * (DFLOW_GRID - 1));
      k = k > DFLOW_GRID - 2 ? DFLOW_GRID - 2 : k;
      const double x0 = (double)k / (DFLOW_GRID - 1), x1 = (double)(k + 1) / (DFLOW_GRID - 1);
      iv = (ifunc[k + 1] - ifunc[k]) / (x1 - x0) * (bd - x0) + ifunc[k];
    }
    float cj = (float)(-iv * full);
    if (isnan(cj)) {
      ++nans;
      if (allow_nan) cj = 0.f;
    } else if (allow_nan && isinf(cj)) {
      cj = cj > 0.f ? 3.4028234663852886e38f : -3.4028234663852886e38f;
    }
    c[j] = cj;
    pc += p[j] * cj;
  }
  if (nans) atomicAdd(&flags[1], nans);
#pragma unroll
  for (int i = 0; i < DFLOW_K; ++i) lg[i] = (p[i] * c[i] - xv[i] * pc) * dalpha;
}

// WT[d, c] = d < D ? Wx[c, d] : Wc[c, d - D];  bsum[c] = bx[c] + bc[c]
__global__ void pack_embed_weights_kernel(const float* __restrict__ Wx, int D, const float* __restrict__ Wc, int Dc,
                                          const float* __restrict__ bx, const float* __restrict__ bc, int C,
                                          float* __restrict__ WT, float* __restrict__ bsum, float* __restrict__ bx_only) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (D + Dc) * C) {
    const int d = i / C, c = i % C;
    WT[i] = d < D ? Wx[(size_t)c * D + d] : Wc[(size_t)c * Dc + (d - D)];
  }
  if (i < C) {
    bsum[i] = bx[i] + bc[i];
    bx_only[i] = bx[i];
  }
}

// ---- small dense layer for a handful of rows: y[r, j] = act(b[j] + sum_k W[j,k] x[r,k]) ----
// One warp per output j (lanes along k, coalesced W reads), loops over rows r (<= MAXR per pass).
// Used for t_embedder (layers.py:52-55), the adaLN table (latent_model.py:380,461; layers.py:72),
// the 7-vector / x_d / design-head linears.  in_act: 0 none, 1 SiLU on the input; out_act: 0 none, 2 GELU.
__global__ void __launch_bounds__(256)
rowvec_linear_kernel(const float* __restrict__ x, int ldx, int R, const float* __restrict__ W, int K,
                     const float* __restrict__ bias, int J, float* __restrict__ y, int ldy, int in_act,
                     int out_act, int accumulate) {
  const int j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (j >= J) return;
  const float* wr = W + (size_t)j * K;
  for (int r = 0; r < R; ++r) {
    float acc = 0.f;
    for (int k = lane; k < K; k += 32) {
      float xv = x[(size_t)r * ldx + k];
      if (in_act == 1) xv = silu(xv);
      acc = fmaf(wr[k], xv, acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) {
      acc += bias ? bias[j] : 0.f;
      if (out_act == 2) acc = gelu_erf(acc);
      float* o = y + (size_t)r * ldy + j;
      *o = accumulate ? *o + acc : acc;
    }
  }
}

// sinusoidal timestep features, layers.py:30-50: [cos(t f_i) | sin(t f_i)], f_i = exp(-ln(1e4) i/128)
__global__ void timestep_features_kernel(const float* __restrict__ t, float mult, int R, float* __restrict__ out) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= R * 128) return;
  int r = idx / 128, i = idx % 128;
  float f = expf(-9.210340371976184f * (float)i / 128.0f);
  float a = t[r] * mult * f;
  out[(size_t)r * 256 + i] = cosf(a);
  out[(size_t)r * 256 + 128 + i] = sinf(a);
}

// RoPE tables (fair-esm): cos/sin[s, i] of s * inv_freq[i], s in [0, n), i in [0, half)
__global__ void rope_table_kernel(const float* __restrict__ inv_freq, int half, int n, float* __restrict__ cs,
                                  float* __restrict__ sn) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n * half) return;
  int s = idx / half, i = idx % half;
  float a = (float)s * inv_freq[i];
  cs[idx] = cosf(a);
  sn[idx] = sinf(a);
}

// aatype embedding + optional adds: s[r,:] (+)= table[idx[r % idx_mod], :]
__global__ void gather_rows_kernel(const float* __restrict__ table, const int64_t* __restrict__ idx, int idx_mod,
                                   int R, int C, float* __restrict__ out, int accumulate) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= R * C) return;
  int r = i / C, c = i % C;
  float v = table[(size_t)idx[r % idx_mod] * C + c];
  out[i] = accumulate ? out[i] + v : v;
}

// mean over the frame axis of a [B,T,L,W] tensor slice: out[b,l,w] = mean_t in[b,t,l,off+w]
// block = (b, l) x 64 channels; its 256 threads are 64 channels x 4 interleaved slices of the frame axis (the one-thread-
// per-output version walked T serially: 21 us for 100 frames on the design config, launched twice per evaluation)
constexpr int MOT_W = 64, MOT_S = 4;
template <typename IT>
__global__ void __launch_bounds__(MOT_W * MOT_S)
mean_over_T_kernel(const IT* __restrict__ in, int ld, int off, int B, int T, int L, int W, float* __restrict__ out) {
  __shared__ float part[MOT_S][MOT_W];
  const int bl = blockIdx.x, b = bl / L, l = bl % L;
  const int wl = threadIdx.x % MOT_W, sl = threadIdx.x / MOT_W, w = blockIdx.y * MOT_W + wl;
  float s = 0.f;
  if (w < W) {
    const IT* src = in + ((size_t)b * T * L + l) * ld + off + w;
#pragma unroll 4
    for (int t = sl; t < T; t += MOT_S) s += to_f(src[(size_t)t * L * ld]);
  }
  part[sl][wl] = s;
  __syncthreads();
  if (sl == 0 && w < W) out[(size_t)bl * W + w] = ((part[0][wl] + part[1][wl]) + (part[2][wl] + part[3][wl])) / (float)T;
}

// design head tail: latent[b,t,l,D-20+k] += logits[b,l,k]   (latent_model.py:268)
__global__ void add_logits_kernel(float* __restrict__ latent, int D, const float* __restrict__ logits, int B, int T,
                                  int L) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * T * L * 20) return;
  int k = i % 20, m = i / 20, l = m % L, b = m / (L * T);
  latent[(size_t)m * D + D - 20 + k] += logits[((size_t)b * L + l) * 20 + k];
}

// dst[s, j, :] = src[s, :] for j < nt   (chunk elements per source row-block)
__global__ void broadcast_rows_kernel(const float* __restrict__ src, float* __restrict__ dst, int streams, int nt,
                                      int chunk) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)streams * nt * chunk) return;
  const size_t e = i % chunk, s = i / ((size_t)nt * chunk);
  dst[i] = src[s * chunk + e];
}
__global__ void axpy_kernel(float* __restrict__ y, const float* __restrict__ x, float a, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] += a * x[i];
}
__global__ void silu_kernel(const float* __restrict__ x, float* __restrict__ y, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = silu(x[i]);
}
__global__ void fill_kernel(float* __restrict__ y, float v, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = v;
}
__global__ void f32_to_bf16_kernel(const float* __restrict__ x, bf16* __restrict__ y, size_t n, float scale) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = __float2bfloat16_rn(x[i] * scale);
}

__global__ void bf16_to_f32_kernel(const bf16* __restrict__ x, float* __restrict__ y, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = __bfloat162float(x[i]);
}

__global__ void bf16_rows_to_f32_kernel(const bf16* __restrict__ x, int ldx, float* __restrict__ y, int ldy, int M, int N) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)M * N) return;
  const int m = (int)(i / N), n = (int)(i % N);
  y[(size_t)m * ldy + n] = __bfloat162float(x[(size_t)m * ldx + n]);
}

}  // namespace mdgen
