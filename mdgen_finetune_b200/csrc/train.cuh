// train.cuh -- row N4: gradients of LatentMDGenModel.forward (latent_model.py:216-269) for RTB posterior fine-tuning
// (HGFNode.run_policy, samplers.py:120-146; loss gfn_diffusion.py:459-460) and Transport.training_losses
// (transport.py:137-222).  Included by model.cu after the forward orchestration (uses MdgenHandle, find_w, the SIMT
// GEMM / attention / IPA launchers).
//
// mdgen_backward RECOMPUTES the forward in fp32 on a small reverse-mode tape -- every op launches its forward kernels
// and records a closure that launches its adjoint (backward*.cuh) -- then runs the closures in reverse.  Activations and
// their gradients live in a per-handle slot pool: the k-th allocation of a pass reuses the k-th slot of the previous pass,
// so after the first call of a shape nothing is allocated.  Parameter gradients ACCUMULATE into caller buffers.
#pragma once
#include <functional>

#include "backward.cuh"
#include "backward_attn.cuh"
#include "backward_ipa.cuh"

namespace mdgen {
namespace train {

struct Var {          // contiguous [rows, cols] fp32 value + gradient (same layout)
  float* v = nullptr;
  float* g = nullptr;
  int rows = 0, cols = 0;
  size_t n() const { return (size_t)rows * cols; }
  Var view(int row0, int nrows) const { return Var{v + (size_t)row0 * cols, g ? g + (size_t)row0 * cols : nullptr, nrows, cols}; }
};

__global__ void gather_rows_fwd_kernel(const float* __restrict__ table, const int64_t* __restrict__ idx, int idx_mod, int M,
                                       int C, int nrows_table, float* __restrict__ out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)M * C) return;
  const int r = (int)(i / C), c = (int)(i % C);
  const long long k = idx[r % idx_mod];
  out[i] = (k >= 0 && k < nrows_table) ? table[(size_t)k * C + c] : 0.f;
}
__global__ void gather_rows_bwd_kernel(const float* __restrict__ dout, const int64_t* __restrict__ idx, int idx_mod, int M, int C,
                                       int nrows_table, float* __restrict__ dtable) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)M * C) return;
  const int r = (int)(i / C), c = (int)(i % C);
  const long long k = idx[r % idx_mod];
  if (k >= 0 && k < nrows_table) atomicAdd(dtable + (size_t)k * C + c, dout[i]);
}
// out[b, t, l, :] = x[b, t, l, :] + pos[l, :] + time[t, :] + s[b, l, :]   (each term optional)
__global__ void embed_sum_kernel(const float* __restrict__ x, const float* __restrict__ pos, const float* __restrict__ tim,
                                 const float* __restrict__ s, int B, int T, int L, int C, float* __restrict__ out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)B * T * L * C) return;
  const int c = (int)(i % C);
  const size_t tok = i / C;
  const int l = (int)(tok % L), t = (int)((tok / L) % T), b = (int)(tok / ((size_t)L * T));
  float v = x[i];
  if (pos) v += pos[(size_t)l * C + c];
  if (tim) v += tim[(size_t)t * C + c];
  if (s) v += s[((size_t)b * L + l) * C + c];
  out[i] = v;
}
// ds[b, l, :] += sum_t dout[b, t, l, :]
__global__ void sum_over_T_kernel(const float* __restrict__ dout, int B, int T, int L, int C, float* __restrict__ ds) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)B * L * C) return;
  const int c = (int)(i % C), l = (int)((i / C) % L), b = (int)(i / ((size_t)C * L));
  float a = 0.f;
  for (int t = 0; t < T; ++t) a += dout[(((size_t)b * T + t) * L + l) * C + c];
  ds[i] += a;
}
// out[b, l, w] = mean_t in[(b, t, l), off + w]   (row pitch ld)
__global__ void mean_T_fwd_kernel(const float* __restrict__ in, int ld, int off, int B, int T, int L, int W, float* __restrict__ out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)B * L * W) return;
  const int w = (int)(i % W), l = (int)((i / W) % L), b = (int)(i / ((size_t)W * L));
  float a = 0.f;
  for (int t = 0; t < T; ++t) a += in[(((size_t)b * T + t) * L + l) * ld + off + w];
  out[i] = a / (float)T;
}
// din[(b, t, l), off + w] += dout[b, l, w] / T
__global__ void mean_T_bwd_kernel(const float* __restrict__ dout, int ld, int off, int B, int T, int L, int W, float* __restrict__ din) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)B * T * L * W) return;
  const int w = (int)(i % W);
  const size_t tok = i / W;
  const int l = (int)(tok % L), b = (int)(tok / ((size_t)L * T));
  din[tok * ld + off + w] += dout[((size_t)b * L + l) * W + w] / (float)T;
}
// latent[(b, t, l), D - 20 + w] += logits[b, l, w]  (latent_model.py:268), out of place
__global__ void add_logits_fwd_kernel(const float* __restrict__ lat, int D, const float* __restrict__ logits, int B, int T, int L,
                                      float* __restrict__ out) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)B * T * L * D) return;
  const int dch = (int)(i % D);
  const size_t tok = i / D;
  float v = lat[i];
  if (dch >= D - 20) v += logits[((tok / ((size_t)L * T)) * L + tok % L) * 20 + (dch - (D - 20))];
  out[i] = v;
}

inline unsigned nblk(size_t n, int th = 256) { return (unsigned)((n + th - 1) / th); }

}  // namespace train
}  // namespace mdgen
