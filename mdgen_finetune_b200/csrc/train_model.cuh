// train_model.cuh -- LatentMDGenModel.forward (latent_model.py:216-269) written once on the tape of train_tape.cuh.
#pragma once
#include "train_tape.cuh"

namespace mdgen {
namespace train {

// MultiheadAttention core on packed qkv [M, 3C] (q pre-scaled by the packed weight): forward = the fp32 SIMT kernels of
// the check mode, adjoint = backward_attn.cuh.
static Var attention_op(Tape& T, const Var& qkv, const MhaW& m, const std::string& prefix, int nb, int S, int inner,
                        const float* key_valid, int mask_mod, bool frame_axis) {
  MdgenHandle* h = T.h;
  const int C = h->C(), H = h->cfg.mha_heads, hd = C / H;
  Var att = T.var(qkv.rows, C);
  if (T.err) return att;
  AttnParams<float> p;
  p.qkv = qkv.v;
  p.ld = 3 * C;
  p.out = att.v;
  p.ldo = C;
  p.nb = nb;
  p.S = S;
  p.inner = inner;
  p.H = H;
  p.bias_k = m.bias_k;
  p.bias_v = m.bias_v;
  p.key_valid = key_valid;
  p.mask_mod = mask_mod;
  const bool rope = !h->flag(MDGEN_F_NO_ROPE);
  p.cs = rope ? (frame_axis ? h->rope_t_cs : h->rope_l_cs) : nullptr;
  p.sn = rope ? (frame_axis ? h->rope_t_sn : h->rope_l_sn) : nullptr;
  if (launch_attention_simt<float>(p, hd, T.st) != 0) T.err = T.err ? T.err : MDGEN_ECUDA;
  float* stats = T.buf((size_t)qkv.rows * H * 3, false);
  float* dbk = T.grad_of(prefix + ".attn.bias_k");
  float* dbv = T.grad_of(prefix + ".attn.bias_v");
  cudaStream_t s = T.st;
  T.bw.push_back([=]() -> int { return bwd::attention_backward(p, att.v, att.g, C, qkv.g, stats, dbk, dbv, s); });
  return att;
}

// x += gate * MHA(modulate(LN(x)))   (latent_model.py:384-387, 466-486)
static Var attn_block(Tape& T, const Var& x, const MhaW& m, const std::string& prefix, const Var& ada, int chunk0,
                      int rows_per_batch, int nbatch, int nb, int S, int inner, const float* key_valid, int mask_mod,
                      bool frame_axis) {
  const float scaling = 1.0f / sqrtf((float)(T.h->C() / T.h->cfg.mha_heads));
  const int C = T.h->C();
  Var hn = T.layernorm(x, &ada, chunk0, chunk0 + 1, rows_per_batch, nbatch, 1e-6f);
  Var qkv = T.linear_packed(hn, m.qkv, {{prefix + ".attn.q_proj", C, scaling}, {prefix + ".attn.k_proj", C, 1.f},
                                        {prefix + ".attn.v_proj", C, 1.f}});
  Var att = attention_op(T, qkv, m, prefix, nb, S, inner, key_valid, mask_mod, frame_axis);
  Var y = T.linear(att, prefix + ".attn.out_proj");
  return T.gate_resid(x, y, &ada, chunk0 + 2, rows_per_batch, nbatch);
}
// x += gate * fc2(gelu(fc1(modulate(LN(x)))))   (latent_model.py:389-392, 488-491)
static Var mlp_block(Tape& T, const Var& x, const std::string& prefix, const Var& ada, int chunk0, int rows_per_batch, int nbatch) {
  Var hn = T.layernorm(x, &ada, chunk0, chunk0 + 1, rows_per_batch, nbatch, 1e-6f);
  Var u = T.linear(hn, prefix + ".fc1");
  Var g = T.gelu(u);
  Var y = T.linear(g, prefix + ".fc2");
  return T.gate_resid(x, y, &ada, chunk0 + 2, rows_per_batch, nbatch);
}

// InvariantPointAttention (ipa.py:92-255) on rows (stream, b, l): n [M, C] -> cat [M, H (c + 4 Pv)]
static Var ipa_core(Tape& T, const Var& n, const IpaLayerW& Lw, const std::string& prefix, int BL, int L) {
  MdgenHandle* h = T.h;
  const IpaDims d{h->cfg.ipa_heads, h->cfg.ipa_head_dim, h->cfg.ipa_qk, h->cfg.ipa_v};
  const int M = n.rows, ntBL = BL;        // one time row per pass
  Var P = T.linear_packed(n, Lw.proj, {{prefix + ".linear_q", d.H * d.c, 1.f}, {prefix + ".linear_kv", 2 * d.H * d.c, 1.f},
                                       {prefix + ".linear_q_points", d.H * d.Pq * 3, 1.f},
                                       {prefix + ".linear_kv_points", d.H * (d.Pq + d.Pv) * 3, 1.f}});
  Var Gq = T.var(M, d.H * d.Pq * 3), Gkv = T.var(M, d.H * (d.Pq + d.Pv) * 3);
  Var cat = T.var(M, d.cat_width());
  if (T.err) return cat;
  const int ldp = d.proj_width();
  const int npts = M * d.H * (2 * d.Pq + d.Pv);
  ipa_points_kernel<<<(npts + 255) / 256, 256, 0, T.st>>>(P.v, ldp, d, h->frames_rot, h->frames_trans, M, ntBL, BL, Gq.v, Gkv.v);
  MD_LAUNCHED();
  const size_t smem = (size_t)d.H * (L + 80) * 4;
  ipa_attention_kernel<<<M, d.H * 32, smem, T.st>>>(P.v, ldp, d, Gq.v, Gkv.v, h->frames_rot, h->frames_trans, ntBL, Lw.head_w,
                                                    h->mask_bl, BL, L, cat.v, 1e5f, 1e-8f);
  MD_LAUNCHED();
  float* dhw = T.grad_of(prefix + ".head_weights");
  cudaStream_t s = T.st;
  const float *rot = h->frames_rot, *trans = h->frames_trans, *mask = h->mask_bl, *hw = Lw.head_w;
  T.bw.push_back([=]() -> int {
    const size_t sm = (size_t)d.H * (3 * L + 80) * 4;
    bwd::ipa_attention_bwd_kernel<<<M, d.H * 32, sm, s>>>(P.v, ldp, d, Gq.v, Gkv.v, rot, trans, ntBL, hw, mask, BL, L, cat.g, 1e5f,
                                                         1e-8f, P.g, Gq.g, Gkv.g, dhw);
    MD_LAUNCHED();
    bwd::ipa_points_bwd_kernel<<<(npts + 255) / 256, 256, 0, s>>>(Gq.g, Gkv.g, d, rot, M, ntBL, BL, P.g, ldp);
    MD_LAUNCHED();
    MD_CUDA(cudaGetLastError());
    return 0;
  });
  return cat;
}

// IPALayer.forward (latent_model.py:378-394) on s [M, C], rows (stream, b, l); c = SiLU(t embedding) [B, C]
static Var ipa_layer(Tape& T, const Var& s, const Var& c, int i, int B, int L) {
  MdgenHandle* h = T.h;
  const IpaLayerW& Lw = h->ipa[i];
  const std::string p = "ipa_layers." + std::to_string(i);
  const int M = s.rows, BL = B * L;
  Var ada = T.linear(c, p + ".adaLN_modulation.1");                       // [B, 6C]
  Var n = T.layernorm(s, nullptr, 0, 0, 1, 1, 1e-5f, p + ".ipa_norm");
  Var cat = ipa_core(T, n, Lw, p + ".ipa", BL, L);
  Var y = T.linear(cat, p + ".ipa.linear_out");
  Var s1 = T.add(s, y);
  Var s2 = attn_block(T, s1, Lw.mha, p + ".mha_l", ada, 0, L, B, M / L, L, 1, h->mask_bl, BL, false);
  return mlp_block(T, s2, p, ada, 3, L, B);
}

struct ForwardOut {
  Var latent;   // [N, D]
  Var x_in;     // the input x as a Var (gradient when requested)
};

// The whole velocity evaluation on the tape.  x [N, D], t [B] device pointers; conditioning from the handle.
static int build_forward(Tape& T, const float* x, const float* t, bool want_dx, ForwardOut* out) {
  MdgenHandle* h = T.h;
  cudaStream_t st = T.st;
  const int C = h->C(), B = h->B, Tn = h->T, L = h->L, D = h->D(), Dc = h->Dc();
  const int N = B * Tn * L, BL = B * L;
  const bool design = h->flag(MDGEN_F_DESIGN), pre = h->flag(MDGEN_F_PREPEND_IPA);
  Var xin = T.input(x, N, D, want_dx);
  out->x_in = xin;
  // ---- timestep embedding (layers.py:30-55) and c = SiLU(t_emb), the input of every adaLN linear
  float* tfeat = T.buf((size_t)B * 256, false);
  if (T.err) return T.err;
  timestep_features_kernel<<<(B * 128 + 127) / 128, 128, 0, st>>>(t, h->cfg.time_multiplier, B, tfeat);
  MD_LAUNCHED();
  Var tf = T.input(tfeat, B, 256, false);
  Var temb = T.linear(T.silu(T.linear(tf, "t_embedder.mlp.0")), "t_embedder.mlp.2");
  Var c = T.silu(temb);
  // ---- prepended IPA stack (run_ipa, latent_model.py:179-214)
  Var ipa_out;
  if (pre) {
    const int streams = h->streams, M = streams * BL;
    Var s = T.var(M, C);
    if (T.err) return T.err;
    if (h->two_stream()) {      // stream 0 = latent_to_emb_r(r7) with start frames, stream 1 = latent_to_emb_f(f7) with end frames
      const Tensor *wr = find_w(h, "latent_to_emb_r.weight"), *br = find_w(h, "latent_to_emb_r.bias");
      const Tensor *wf = find_w(h, "latent_to_emb_f.weight"), *bf_ = find_w(h, "latent_to_emb_f.bias");
      MD_CHECK(wr && br && wf && bf_, "missing latent_to_emb_f/r");
      T.linear_raw(h->rel7, nullptr, 7, BL, 7, wr->d, br->d, C, s.v, s.g, C, {{"latent_to_emb_r", C, 1.f}});
      T.linear_raw(h->rel7 + (size_t)BL * 7, nullptr, 7, BL, 7, wf->d, bf_->d, C, s.v + (size_t)BL * C, s.g + (size_t)BL * C, C,
                   {{"latent_to_emb_f", C, 1.f}});
    } else {
      MD_CUDA(cudaMemsetAsync(s.v, 0, s.n() * 4, st));
    }
    if (h->has_aatype && !h->flag(MDGEN_F_NO_AA_EMB)) s = T.add(s, T.gather("aatype_to_emb.weight", h->aatype, BL, M, C));
    if (design) {               // x_d = x[..., -20:].mean(1) -> x_d_to_emb, added to every stream (latent_model.py:237-240, 193-207)
      Var xd = T.var(BL, 20);
      if (T.err) return T.err;
      mean_T_fwd_kernel<<<nblk((size_t)BL * 20), 256, 0, st>>>(x, D, D - 20, B, Tn, L, 20, xd.v);
      MD_LAUNCHED();
      T.bw.push_back([=]() -> int {
        if (xin.g) mean_T_bwd_kernel<<<nblk((size_t)N * 20), 256, 0, st>>>(xd.g, D, D - 20, B, Tn, L, 20, xin.g);
        MD_LAUNCHED();
        return 0;
      });
      Var e = T.linear(xd, "x_d_to_emb");                                 // [BL, C]
      Var rep = T.var(M, C);
      if (T.err) return T.err;
      for (int k = 0; k < streams; ++k) MD_CUDA(cudaMemcpyAsync(rep.v + (size_t)k * BL * C, e.v, e.n() * 4, cudaMemcpyDeviceToDevice, st));
      T.bw.push_back([=]() -> int {
        for (int k = 0; k < streams; ++k) bwd::acc_kernel<<<nblk(e.n()), 256, 0, st>>>(e.g, rep.g + (size_t)k * BL * C, 1.f, e.n());
        MD_LAUNCHED();
        return 0;
      });
      s = T.add(s, rep);
    }
    for (size_t i = 0; i < h->ipa.size(); ++i) s = ipa_layer(T, s, c, (int)i, B, L);
    ipa_out = streams == 2 ? T.add(s.view(0, BL), s.view(BL, BL)) : s;
  }
  // ---- token embedding (latent_model.py:242-255)
  Var tok = T.linear(xin, "latent_to_emb");
  if (h->has_cond) {
    Var xc = T.input(h->x_cond, N, Dc, false);
    tok = T.add(tok, T.add(T.linear(xc, "cond_to_emb"), T.gather("mask_to_emb.weight", h->cond_mask, N, N, C)));
  }
  Var xres = T.var(N, C);
  if (T.err) return T.err;
  {
    const Tensor* pe = h->flag(MDGEN_F_ABS_POS) ? find_w(h, "pos_embed") : nullptr;
    const Tensor* te = h->flag(MDGEN_F_ABS_TIME) ? find_w(h, "time_embed") : nullptr;
    embed_sum_kernel<<<nblk(xres.n()), 256, 0, st>>>(tok.v, pe ? pe->d : nullptr, te ? te->d : nullptr, pre ? ipa_out.v : nullptr, B, Tn,
                                                    L, C, xres.v);
    MD_LAUNCHED();
    T.bw.push_back([=]() -> int {
      bwd::acc_kernel<<<nblk(xres.n()), 256, 0, st>>>(tok.g, xres.g, 1.f, xres.n());
      if (pre) sum_over_T_kernel<<<nblk((size_t)BL * C), 256, 0, st>>>(xres.g, B, Tn, L, C, ipa_out.g);
      MD_LAUNCHED();
      return 0;
    });
  }
  // ---- factorised spatio-temporal blocks (latent_model.py:456-493)
  Var xr = xres;
  for (size_t i = 0; i < h->layers.size(); ++i) {
    const MainLayerW& Lw = h->layers[i];
    const std::string p = "layers." + std::to_string(i);
    Var ada = T.linear(c, p + ".adaLN_modulation.1");                     // [B, 9C]
    xr = attn_block(T, xr, Lw.mha_l, p + ".mha_l", ada, 0, Tn * L, B, B * Tn, L, 1, h->mask, N, false);
    xr = attn_block(T, xr, Lw.mha_t, p + ".mha_t", ada, 3, Tn * L, B, B * L, Tn, L, h->mask, N, true);
    xr = mlp_block(T, xr, p, ada, 6, Tn * L, B);
  }
  // ---- final layer (layers.py:71-75)
  Var adaF = T.linear(c, "emb_to_latent.adaLN_modulation.1");             // [B, 2C]: shift | scale
  Var hn = T.layernorm(xr, &adaF, 0, 1, Tn * L, B, 1e-6f);
  Var latent = T.linear(hn, "emb_to_latent.linear");
  if (design) {                 // latent[..., -20:] += emb_to_logits(gelu(fc3(mean_T fc2(gelu(fc1(x))))))   (latent_model.py:262-268)
    Var g1 = T.gelu(T.linear(xr, "fc1"));
    Var x2 = T.linear(g1, "fc2");
    Var mean = T.var(BL, C);
    if (T.err) return T.err;
    mean_T_fwd_kernel<<<nblk(mean.n()), 256, 0, st>>>(x2.v, C, 0, B, Tn, L, C, mean.v);
    MD_LAUNCHED();
    T.bw.push_back([=]() -> int {
      mean_T_bwd_kernel<<<nblk(x2.n()), 256, 0, st>>>(mean.g, C, 0, B, Tn, L, C, x2.g);
      MD_LAUNCHED();
      return 0;
    });
    Var logits = T.linear(T.gelu(T.linear(mean, "fc3")), "emb_to_logits");   // [BL, 20]
    Var lat2 = T.var(N, D);
    if (T.err) return T.err;
    add_logits_fwd_kernel<<<nblk(lat2.n()), 256, 0, st>>>(latent.v, D, logits.v, B, Tn, L, lat2.v);
    MD_LAUNCHED();
    T.bw.push_back([=]() -> int {
      bwd::acc_kernel<<<nblk(lat2.n()), 256, 0, st>>>(latent.g, lat2.g, 1.f, lat2.n());
      // d logits[b, l, w] = sum_t d latent[(b, t, l), D - 20 + w] = T * mean_T
      mean_T_fwd_kernel<<<nblk(logits.n()), 256, 0, st>>>(lat2.g, D, D - 20, B, Tn, L, 20, logits.g);
      bwd::acc_kernel<<<nblk(logits.n()), 256, 0, st>>>(logits.g, logits.g, (float)(Tn - 1), logits.n());
      MD_LAUNCHED();
      return 0;
    });
    latent = lat2;
  }
  out->latent = latent;
  MD_CUDA(cudaGetLastError());
  return T.err;
}

}  // namespace train
}  // namespace mdgen
