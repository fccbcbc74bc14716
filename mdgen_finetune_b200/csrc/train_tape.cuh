// train_tape.cuh -- the reverse-mode tape of row N4 (see train.cuh).  Included by model.cu after MdgenHandle.
#pragma once
#include "train.cuh"

namespace mdgen {
namespace train {

struct Tape {
  MdgenHandle* h;
  cudaStream_t st;
  std::map<std::string, float*> pgrad;          // parameter name -> caller's gradient buffer (accumulated); absent = frozen
  std::vector<std::function<int()>> bw;          // adjoints, run in reverse
  size_t slot = 0;
  int err = 0;

  Tape(MdgenHandle* h_, cudaStream_t s) : h(h_), st(s) {}

  float* grad_of(const std::string& name) const {
    auto it = pgrad.find(name);
    return it == pgrad.end() ? nullptr : it->second;
  }
  const float* param(const std::string& name) {
    const Tensor* t = find_w(h, name);
    if (!t) {
      set_error("mdgen_backward: missing weight " + name);
      err = MDGEN_ENOTFOUND;
      return nullptr;
    }
    return t->d;
  }
  // k-th buffer of this pass = k-th slot of the pool (grown on first use / when a shape needs more)
  float* buf(size_t nfloat, bool zero) {
    const size_t bytes = (nfloat ? nfloat : 1) * sizeof(float);
    if (slot >= h->train_slots.size()) h->train_slots.push_back(DevBuf());
    DevBuf& b = h->train_slots[slot++];
    if (b.bytes < bytes) {
      if (b.p) {
        cudaStreamSynchronize(st);
        cudaFree(b.p);
      }
      b.p = nullptr;
      b.bytes = 0;
      if (cudaMalloc(&b.p, bytes) != cudaSuccess) {
        set_error("mdgen_backward: out of device memory for the activation tape");
        err = MDGEN_ECUDA;
        return nullptr;
      }
      b.bytes = bytes;
    }
    if (zero) cudaMemsetAsync(b.p, 0, bytes, st);
    return (float*)b.p;
  }
  Var var(int rows, int cols) {
    Var v;
    v.rows = rows;
    v.cols = cols;
    v.v = buf(v.n(), false);
    v.g = buf(v.n(), true);
    return v;
  }
  // a tensor the caller supplies (no gradient unless `want_grad`)
  Var input(const float* ptr, int rows, int cols, bool want_grad) {
    Var v;
    v.rows = rows;
    v.cols = cols;
    v.v = const_cast<float*>(ptr);
    v.g = want_grad ? buf(v.n(), true) : nullptr;
    return v;
  }
  int check(cudaError_t e, const char* what) {
    if (e != cudaSuccess && !err) {
      set_error(std::string("mdgen_backward: ") + what + ": " + cudaGetErrorString(e));
      err = MDGEN_ECUDA;
    }
    return err;
  }

  // ---- Y[:, ycol : ycol + N] = X W^T + b on raw pointers (row pitches ldx / ldy); W [N, K] ----------------------------
  // wparts: the packed weight W is the row-wise concatenation of several reference parameters (fused q|k|v, fused IPA
  // projections): each part (name prefix, rows, alpha) receives its slice of the weight / bias gradient times alpha.
  struct WPart { std::string name; int rows; float alpha; };
  void linear_raw(const float* Xv, float* Xg, int ldx, int M, int K, const float* W, const float* b, int N, float* Yv, float* Yg,
                  int ldy, const std::vector<WPart>& parts) {
    if (err || M <= 0) return;
    EpiParams e;
    e.bias = b;
    e.out = Yv;
    e.ldo = ldy;
    if (launch_gemm_simt<float, float>(Xv, ldx, W, K, M, N, K, EPI_STORE, e, st) != 0) err = err ? err : MDGEN_ECUDA;
    std::vector<std::pair<float*, float*>> g;
    bool any = false;
    for (auto& p : parts) {
      g.emplace_back(grad_of(p.name + ".weight"), grad_of(p.name + ".bias"));
      any = any || g.back().first || g.back().second;
    }
    cudaStream_t s = st;
    bw.push_back([=]() -> int {
      if (Xg) MD_TRY(bwd::gemm_strided(s, Yg, ldy, 1, W, K, 1, Xg, ldx, M, K, N, true));      // dX += dY W
      if (any) {
        int r0 = 0;
        for (size_t i = 0; i < parts.size(); ++i) {
          const int rows = parts[i].rows;
          if (g[i].first)       // dW[n, k] += alpha sum_m dY[m, r0 + n] X[m, k]
            MD_TRY(bwd::gemm_strided(s, Yg + r0, 1, ldy, Xv, ldx, 1, g[i].first, K, rows, K, M, true, parts[i].alpha));
          if (g[i].second) MD_TRY(bwd::colsum(s, Yg + r0, ldy, M, rows, g[i].second, parts[i].alpha));
          r0 += rows;
        }
      }
      return 0;
    });
  }
  Var linear(const Var& X, const std::string& name) {          // a reference nn.Linear by name
    const Tensor* w = find_w(h, name + ".weight");
    const Tensor* b = find_w(h, name + ".bias");
    if (!w || !b) {
      set_error("mdgen_backward: missing weight " + name);
      err = MDGEN_ENOTFOUND;
      return Var();
    }
    const int N = (int)w->shape[0], K = (int)w->shape[1];
    Var Y = var(X.rows, N);
    linear_raw(X.v, X.g, K, X.rows, K, w->d, b->d, N, Y.v, Y.g, N, {{name, N, 1.f}});
    return Y;
  }
  Var linear_packed(const Var& X, const LinearW& lw, const std::vector<WPart>& parts) {
    Var Y = var(X.rows, lw.N);
    linear_raw(X.v, X.g, lw.K, X.rows, lw.K, lw.w, lw.b, lw.N, Y.v, Y.g, lw.N, parts);
    return Y;
  }

  // ---- element-wise ------------------------------------------------------------------------------------------------
  Var silu(const Var& X) {
    Var Y = var(X.rows, X.cols);
    if (err) return Y;
    const size_t n = X.n();
    silu_kernel<<<nblk(n), 256, 0, st>>>(X.v, Y.v, n);
    MD_LAUNCHED();
    cudaStream_t s = st;
    bw.push_back([=]() -> int {
      if (X.g) bwd::silu_bwd_kernel<<<nblk(n), 256, 0, s>>>(X.v, Y.g, X.g, n);
      MD_LAUNCHED();
      return 0;
    });
    return Y;
  }
  Var gelu(const Var& X) {
    Var Y = var(X.rows, X.cols);
    if (err) return Y;
    const size_t n = X.n();
    bwd::gelu_fwd_kernel<<<nblk(n), 256, 0, st>>>(X.v, Y.v, n);
    MD_LAUNCHED();
    cudaStream_t s = st;
    bw.push_back([=]() -> int {
      if (X.g) bwd::gelu_bwd_kernel<<<nblk(n), 256, 0, s>>>(X.v, Y.g, X.g, n);
      MD_LAUNCHED();
      return 0;
    });
    return Y;
  }
  Var add(const Var& A, const Var& B) {
    Var Y = var(A.rows, A.cols);
    if (err) return Y;
    const size_t n = A.n();
    bwd::add_kernel<<<nblk(n), 256, 0, st>>>(A.v, B.v, Y.v, n);
    MD_LAUNCHED();
    cudaStream_t s = st;
    bw.push_back([=]() -> int {
      if (A.g) bwd::acc_kernel<<<nblk(n), 256, 0, s>>>(A.g, Y.g, 1.f, n);
      if (B.g) bwd::acc_kernel<<<nblk(n), 256, 0, s>>>(B.g, Y.g, 1.f, n);
      MD_LAUNCHED();
      return 0;
    });
    return Y;
  }
  // LayerNorm(eps, no affine) * (1 + scale[b]) + shift[b] with shift / scale = column chunks of `ada` [nbatch, ada.cols]
  // (layers.py:14-15), or the affine LayerNorm of ipa_norm when gamma_name is given.
  Var layernorm(const Var& X, const Var* ada, int shift_chunk, int scale_chunk, int rows_per_batch, int nbatch, float eps,
                const std::string& affine_name = "") {
    Var Y = var(X.rows, X.cols);
    if (err) return Y;
    const int M = X.rows, C = X.cols;
    const float *shift = nullptr, *scale = nullptr, *gamma = nullptr, *beta = nullptr;
    float *dshift = nullptr, *dscale = nullptr, *dgamma = nullptr, *dbeta = nullptr;
    int ld = 0;
    if (ada) {
      ld = ada->cols;
      shift = ada->v + (size_t)shift_chunk * C;
      scale = ada->v + (size_t)scale_chunk * C;
      dshift = ada->g + (size_t)shift_chunk * C;
      dscale = ada->g + (size_t)scale_chunk * C;
    } else {
      gamma = param(affine_name + ".weight");
      beta = param(affine_name + ".bias");
      dgamma = grad_of(affine_name + ".weight");
      dbeta = grad_of(affine_name + ".bias");
      rows_per_batch = nbatch = 1;
    }
    if (err) return Y;
    ln_modulate_kernel<float, 384><<<((M + LN_ROWS - 1) / LN_ROWS * 32 + 255) / 256, 256, 0, st>>>(
        X.v, Y.v, M, shift, scale, ld, rows_per_batch, nbatch, gamma, beta, eps);
    MD_LAUNCHED();
    cudaStream_t s = st;
    bw.push_back([=]() -> int {
      bwd::ln_bwd_kernel<<<(M * 32 + 255) / 256, 256, 0, s>>>(X.v, Y.g, X.g, M, C, scale, dshift, dscale, ld, rows_per_batch, nbatch,
                                                            gamma, dgamma, dbeta, eps);
      MD_LAUNCHED();
      return 0;
    });
    return Y;
  }
  // Xo = X + gate[b] * Y   (gate = a column chunk of `ada`; ada == null: plain residual)
  Var gate_resid(const Var& X, const Var& Y, const Var* ada, int gate_chunk, int rows_per_batch, int nbatch) {
    Var O = var(X.rows, X.cols);
    if (err) return O;
    const int M = X.rows, C = X.cols;
    const float* gate = ada ? ada->v + (size_t)gate_chunk * C : nullptr;
    float* dgate = ada ? ada->g + (size_t)gate_chunk * C : nullptr;
    const int ld = ada ? ada->cols : 0;
    const size_t n = X.n();
    bwd::gate_resid_fwd_kernel<<<nblk(n), 256, 0, st>>>(X.v, Y.v, gate, ld, rows_per_batch, nbatch, O.v, M, C);
    MD_LAUNCHED();
    cudaStream_t s = st;
    bw.push_back([=]() -> int {
      bwd::gate_resid_bwd_kernel<<<nblk(n), 256, 0, s>>>(O.g, Y.v, gate, dgate, ld, rows_per_batch, nbatch, X.g, Y.g, M, C);
      MD_LAUNCHED();
      return 0;
    });
    return O;
  }
  Var gather(const std::string& table_name, const int64_t* idx, int idx_mod, int M, int C) {
    Var Y = var(M, C);
    const Tensor* tb = find_w(h, table_name);
    if (!tb) {
      set_error("mdgen_backward: missing weight " + table_name);
      err = MDGEN_ENOTFOUND;
    }
    if (err) return Y;
    const int nrows = (int)tb->shape[0];
    const size_t n = Y.n();
    gather_rows_fwd_kernel<<<nblk(n), 256, 0, st>>>(tb->d, idx, idx_mod, M, C, nrows, Y.v);
    MD_LAUNCHED();
    float* dt = grad_of(table_name);
    cudaStream_t s = st;
    bw.push_back([=]() -> int {
      if (dt) gather_rows_bwd_kernel<<<nblk(n), 256, 0, s>>>(Y.g, idx, idx_mod, M, C, nrows, dt);
      MD_LAUNCHED();
      return 0;
    });
    return Y;
  }

  int run_backward() {
    for (auto it = bw.rbegin(); it != bw.rend(); ++it) {
      const int r = (*it)();
      if (r != 0) return r;
    }
    return check(cudaGetLastError(), "backward launch");
  }
};

}  // namespace train
}  // namespace mdgen
