// model.cu -- handle, weight packing, per-forward orchestration and the C ABI of libmdgen_b200.so.
// Mirrors LatentMDGenModel.forward (src/mdgen/model/latent_model.py:216-269) and the fixed-grid
// Euler sampler (transport.py:407-450, integrators.py:74-114); see include/mdgen_b200.h.
#include <cuda.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/mdgen_b200.h"
#include "attention_simt.cuh"
#include "common.cuh"
#include "elementwise.cuh"
#include "gemm_simt.cuh"
#include "gemm_tc.cuh"
#include "ipa.cuh"

namespace mdgen {

static thread_local std::string g_err;
thread_local long long g_launches = 0;
void set_error(const std::string& msg) { g_err = msg; }

// ---- driver entry point for TMA descriptors (no link-time dependency on libcuda) ----------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// 2-D bf16 row-major [rows, cols] (cols contiguous, row pitch ld elements), box [box_rows, 64], 128B swizzle
static int make_tmap_2d(CUtensorMap* m, const bf16* ptr, int rows, int cols, int ld, int box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  MD_CHECK(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  MD_CHECK(((uintptr_t)ptr & 15) == 0 && (ld * 2) % 16 == 0, "TMA operand must be 16-byte aligned");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)tc::BLOCK_K, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)ptr, dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  MD_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed");
  return 0;
}

static int g_num_sms = 0;
static int num_sms() {
  if (!g_num_sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
  }
  return g_num_sms;
}

template <int BN, int EPI, typename OT>
static int launch_tc_inst(const CUtensorMap& ta, const CUtensorMap& tw, int M, int N, int K, const EpiParams& ep,
                          cudaStream_t st) {
  static bool configured = false;
  auto kern = tc::gemm_tc_kernel<BN, EPI, OT>;
  if (!configured) {
    MD_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::smem_bytes(BN)));
    configured = true;
  }
  int tiles = ((M + tc::BLOCK_M - 1) / tc::BLOCK_M) * ((N + BN - 1) / BN);
  int grid = tiles < num_sms() ? tiles : num_sms();
  kern<<<grid, tc::NUM_THREADS, tc::smem_bytes(BN), st>>>(ta, tw, M, N, K, ep);
  MD_LAUNCHED();
  MD_CUDA(cudaGetLastError());
  return 0;
}

template <int BN>
static int launch_tc_bn(const CUtensorMap& ta, const CUtensorMap& tw, int M, int N, int K, int epi, bool out_bf16,
                        const EpiParams& ep, cudaStream_t st) {
  switch (epi) {
    case EPI_STORE:
      return out_bf16 ? launch_tc_inst<BN, EPI_STORE, bf16>(ta, tw, M, N, K, ep, st)
                      : launch_tc_inst<BN, EPI_STORE, float>(ta, tw, M, N, K, ep, st);
    case EPI_GELU:
      return out_bf16 ? launch_tc_inst<BN, EPI_GELU, bf16>(ta, tw, M, N, K, ep, st)
                      : launch_tc_inst<BN, EPI_GELU, float>(ta, tw, M, N, K, ep, st);
    case EPI_GATE_RESID: return launch_tc_inst<BN, EPI_GATE_RESID, float>(ta, tw, M, N, K, ep, st);
    case EPI_RESID: return launch_tc_inst<BN, EPI_RESID, float>(ta, tw, M, N, K, ep, st);
    case EPI_EULER: return launch_tc_inst<BN, EPI_EULER, float>(ta, tw, M, N, K, ep, st);
  }
  set_error("launch_gemm_tc: bad epilogue");
  return -1;
}

// C = A[M,K] . W[N,K]^T on tcgen05.  block_n = 0 picks a tile width from the problem shape.
int launch_gemm_tc(const bf16* A, int lda, const bf16* W, int ldw, int M, int N, int K, int epi, bool out_bf16,
                   const EpiParams& ep, int block_n, cudaStream_t st) {
  if (M <= 0 || N <= 0) return 0;
  MD_CHECK(K % 8 == 0, "gemm_tc: K must be a multiple of 8");
  if (block_n == 0) {
    if (N <= 32) block_n = 32;
    else if (N <= 64) block_n = 64;
    else {
      // fill the machine: prefer 128-wide tiles unless that leaves SMs idle
      long t128 = (long)((M + 127) / 128) * ((N + 127) / 128);
      block_n = t128 >= num_sms() ? 128 : 64;
    }
  }
  CUtensorMap ta, tw;
  MD_TRY(make_tmap_2d(&ta, A, M, K, lda, tc::BLOCK_M));
  MD_TRY(make_tmap_2d(&tw, W, N, K, ldw, block_n));
  switch (block_n) {
    case 32: return launch_tc_bn<32>(ta, tw, M, N, K, epi, out_bf16, ep, st);
    case 64: return launch_tc_bn<64>(ta, tw, M, N, K, epi, out_bf16, ep, st);
    case 128: return launch_tc_bn<128>(ta, tw, M, N, K, epi, out_bf16, ep, st);
  }
  set_error("launch_gemm_tc: block_n must be 32, 64 or 128");
  return -1;
}

// =============================================================================================
// handle
// =============================================================================================
struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
};

struct Tensor {
  float* d = nullptr;  // device fp32
  std::vector<int64_t> shape;
  size_t numel() const {
    size_t n = 1;
    for (auto s : shape) n *= (size_t)s;
    return n;
  }
};

struct LinearW {      // one GEMM-able projection
  const float* w = nullptr;  // [N, K] fp32
  const float* b = nullptr;  // [N]
  const bf16* wb = nullptr;  // [N, K] bf16 copy (bf16 mode)
  int N = 0, K = 0;
};

struct MhaW {
  LinearW qkv, out;
  const float* bias_k = nullptr;
  const float* bias_v = nullptr;
};

struct IpaLayerW {
  const float* norm_w = nullptr;
  const float* norm_b = nullptr;
  const float* head_w = nullptr;
  LinearW proj, lin_out;
  MhaW mha;
  LinearW fc1, fc2;
  int ada_off = 0;  // column offset into the adaLN table
};

struct MainLayerW {
  MhaW mha_l, mha_t;
  LinearW fc1, fc2;
  int ada_off = 0;
};

}  // namespace mdgen

using namespace mdgen;

struct MdgenHandle {
  MdgenConfig cfg;
  int device = 0;
  std::map<std::string, Tensor> weights;  // raw fp32 copies, reference names
  std::vector<void*> packed;              // buffers owned by finalize()
  bool finalized = false;

  // packed views
  std::vector<IpaLayerW> ipa;
  std::vector<MainLayerW> layers;
  LinearW final_lin, design_fc1, design_fc2, design_fc3, design_logits;
  float* ada_w = nullptr;  // [ada_cols, C]
  float* ada_b = nullptr;
  int ada_cols = 0, ada_final_off = 0;

  // conditioning / workspace (valid after set_conditioning)
  int B = 0, T = 0, L = 0;
  bool has_cond = false;
  std::vector<DevBuf> ws_owned;
  size_t ws_bytes = 0;
  // step-invariant
  float* mask = nullptr;      // [N]
  float* mask_bl = nullptr;   // [B*L]  (mask[:, 0])
  float* x_cond = nullptr;    // [N, Dc] or null
  int64_t* cond_mask = nullptr;
  int64_t* aatype = nullptr;
  float* frames_rot = nullptr;    // [streams*B*L, 9]
  float* frames_trans = nullptr;  // [streams*B*L, 3]
  float* rel7 = nullptr;          // [2, B*L, 7]  (stream 0 = r7, stream 1 = f7)
  int streams = 1;
  float *rope_l_cs = nullptr, *rope_l_sn = nullptr, *rope_t_cs = nullptr, *rope_t_sn = nullptr;
  // per-step
  float *tbuf = nullptr, *tfeat = nullptr, *th1 = nullptr, *temb = nullptr, *ada = nullptr;
  float* xres = nullptr;   // [N, C] residual stream
  void* hbuf = nullptr;    // [N, C]  AT
  void* qkv = nullptr;     // [N, 3C] AT
  void* att = nullptr;     // [N, C]  AT
  void* mlp = nullptr;     // [N, 4C] AT
  float* vel = nullptr;    // [N, D] scratch velocity
  float *ipa_s = nullptr, *ipa_n = nullptr, *ipa_p = nullptr, *ipa_gq = nullptr, *ipa_gkv = nullptr,
        *ipa_cat = nullptr, *ipa_out = nullptr;
  float *xd = nullptr, *dh_mean = nullptr, *dh_a = nullptr, *dh_b = nullptr, *dh_logits = nullptr;
  int launches_last_forward = 0;

  int C() const { return cfg.embed_dim; }
  int D() const { return cfg.latent_dim; }
  int Dc() const { return (cfg.flags & MDGEN_F_DESIGN) ? cfg.latent_dim - 20 : cfg.latent_dim; }
  bool flag(int f) const { return (cfg.flags & f) != 0; }
  bool bf16_mode() const { return cfg.precision == MDGEN_PREC_BF16; }
  bool two_stream() const { return flag(MDGEN_F_TPS) || flag(MDGEN_F_INPAINT); }
};

namespace mdgen {

static int dev_alloc(MdgenHandle* h, void** out, size_t bytes, bool packed = false) {
  void* p = nullptr;
  if (bytes == 0) bytes = 16;
  MD_CUDA(cudaMalloc(&p, bytes));
  if (packed) h->packed.push_back(p);
  else {
    DevBuf b;
    b.p = p;
    b.bytes = bytes;
    h->ws_owned.push_back(b);
    h->ws_bytes += bytes;
  }
  *out = p;
  return 0;
}

static const Tensor* find_w(MdgenHandle* h, const std::string& name) {
  auto it = h->weights.find(name);
  return it == h->weights.end() ? nullptr : &it->second;
}

#define NEED_W(var, name)                                                   \
  const Tensor* var = find_w(h, name);                                      \
  if (!var) {                                                               \
    set_error(std::string("missing weight: ") + (name));                    \
    return MDGEN_ENOTFOUND;                                                 \
  }

static int to_bf16(MdgenHandle* h, const float* src, size_t n, const bf16** out, cudaStream_t st) {
  void* p;
  MD_TRY(dev_alloc(h, &p, n * sizeof(bf16), true));
  f32_to_bf16_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(src, (bf16*)p, n, 1.0f);
  MD_CUDA(cudaGetLastError());
  *out = (const bf16*)p;
  return 0;
}

static int pack_linear(MdgenHandle* h, const std::string& name, LinearW* lw, bool want_bf16, cudaStream_t st) {
  NEED_W(w, name + ".weight");
  NEED_W(b, name + ".bias");
  MD_CHECK(w->shape.size() == 2, "linear weight must be 2-D");
  lw->w = w->d;
  lw->b = b->d;
  lw->N = (int)w->shape[0];
  lw->K = (int)w->shape[1];
  if (want_bf16) MD_TRY(to_bf16(h, w->d, w->numel(), &lw->wb, st));
  return 0;
}

__global__ void scale_copy_kernel(const float* __restrict__ src, float* __restrict__ dst, size_t n, float s) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[i] * s;
}
static int copy_scaled(const float* src, float* dst, size_t n, float s, cudaStream_t st) {
  scale_copy_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(src, dst, n, s);
  MD_CUDA(cudaGetLastError());
  return 0;
}

// fused [q*d^-0.5 | k | v] projection (mha.py:260-263) + out_proj + bias_k/bias_v
static int pack_mha(MdgenHandle* h, const std::string& p, MhaW* m, cudaStream_t st) {
  const int C = h->C();
  const float scaling = 1.0f / sqrtf((float)(C / h->cfg.mha_heads));
  NEED_W(qw, p + ".attn.q_proj.weight");
  NEED_W(qb, p + ".attn.q_proj.bias");
  NEED_W(kw, p + ".attn.k_proj.weight");
  NEED_W(kb, p + ".attn.k_proj.bias");
  NEED_W(vw, p + ".attn.v_proj.weight");
  NEED_W(vb, p + ".attn.v_proj.bias");
  NEED_W(bk, p + ".attn.bias_k");
  NEED_W(bv, p + ".attn.bias_v");
  float *w, *b;
  MD_TRY(dev_alloc(h, (void**)&w, (size_t)3 * C * C * 4, true));
  MD_TRY(dev_alloc(h, (void**)&b, (size_t)3 * C * 4, true));
  MD_TRY(copy_scaled(qw->d, w, (size_t)C * C, scaling, st));
  MD_TRY(copy_scaled(kw->d, w + (size_t)C * C, (size_t)C * C, 1.f, st));
  MD_TRY(copy_scaled(vw->d, w + (size_t)2 * C * C, (size_t)C * C, 1.f, st));
  MD_TRY(copy_scaled(qb->d, b, C, scaling, st));
  MD_TRY(copy_scaled(kb->d, b + C, C, 1.f, st));
  MD_TRY(copy_scaled(vb->d, b + 2 * C, C, 1.f, st));
  m->qkv.w = w;
  m->qkv.b = b;
  m->qkv.N = 3 * C;
  m->qkv.K = C;
  if (h->bf16_mode()) MD_TRY(to_bf16(h, w, (size_t)3 * C * C, &m->qkv.wb, st));
  MD_TRY(pack_linear(h, p + ".attn.out_proj", &m->out, h->bf16_mode(), st));
  m->bias_k = bk->d;
  m->bias_v = bv->d;
  return 0;
}

static int finalize_weights(MdgenHandle* h, cudaStream_t st) {
  for (void* p : h->packed) cudaFree(p);
  h->packed.clear();
  h->ipa.clear();
  h->layers.clear();
  const int C = h->C(), nl = h->cfg.num_layers;
  const bool bfm = h->bf16_mode();
  const bool pre = h->flag(MDGEN_F_PREPEND_IPA);
  // adaLN table: [ipa layers 6C each | main layers 9C each | final 2C]
  h->ada_cols = (pre ? nl * 6 * C : 0) + nl * 9 * C + 2 * C;
  MD_TRY(dev_alloc(h, (void**)&h->ada_w, (size_t)h->ada_cols * C * 4, true));
  MD_TRY(dev_alloc(h, (void**)&h->ada_b, (size_t)h->ada_cols * 4, true));
  int off = 0;
  auto put_ada = [&](const std::string& name, int rows) -> int {
    NEED_W(w, name + ".weight");
    NEED_W(b, name + ".bias");
    MD_CHECK((int)w->shape[0] == rows && (int)w->shape[1] == C, "adaLN weight shape");
    MD_TRY(copy_scaled(w->d, h->ada_w + (size_t)off * C, (size_t)rows * C, 1.f, st));
    MD_TRY(copy_scaled(b->d, h->ada_b + off, rows, 1.f, st));
    off += rows;
    return 0;
  };
  if (pre) {
    const IpaDims d{h->cfg.ipa_heads, h->cfg.ipa_head_dim, h->cfg.ipa_qk, h->cfg.ipa_v};
    MD_CHECK(d.c == 32 && d.Pq == 8 && d.Pv == 8 && d.H <= 4, "IPA kernels need c=32, 8 q/v points, <=4 heads");
    for (int i = 0; i < nl; ++i) {
      IpaLayerW L;
      const std::string p = "ipa_layers." + std::to_string(i);
      L.ada_off = off;
      MD_TRY(put_ada(p + ".adaLN_modulation.1", 6 * C));
      NEED_W(nw, p + ".ipa_norm.weight");
      NEED_W(nb, p + ".ipa_norm.bias");
      NEED_W(hw, p + ".ipa.head_weights");
      L.norm_w = nw->d;
      L.norm_b = nb->d;
      L.head_w = hw->d;
      // fused projection [linear_q | linear_kv | linear_q_points | linear_kv_points]  (ipa.py:113-138)
      const char* names[4] = {".ipa.linear_q", ".ipa.linear_kv", ".ipa.linear_q_points", ".ipa.linear_kv_points"};
      const int pw = d.proj_width();
      float *w, *b;
      MD_TRY(dev_alloc(h, (void**)&w, (size_t)pw * C * 4, true));
      MD_TRY(dev_alloc(h, (void**)&b, (size_t)pw * 4, true));
      int r = 0;
      for (int k = 0; k < 4; ++k) {
        NEED_W(ww, p + names[k] + ".weight");
        NEED_W(bb, p + names[k] + ".bias");
        MD_TRY(copy_scaled(ww->d, w + (size_t)r * C, ww->numel(), 1.f, st));
        MD_TRY(copy_scaled(bb->d, b + r, bb->numel(), 1.f, st));
        r += (int)ww->shape[0];
      }
      MD_CHECK(r == pw, "IPA projection widths");
      L.proj.w = w;
      L.proj.b = b;
      L.proj.N = pw;
      L.proj.K = C;
      MD_TRY(pack_linear(h, p + ".ipa.linear_out", &L.lin_out, false, st));
      MD_TRY(pack_mha(h, p + ".mha_l", &L.mha, st));
      MD_TRY(pack_linear(h, p + ".fc1", &L.fc1, bfm, st));
      MD_TRY(pack_linear(h, p + ".fc2", &L.fc2, bfm, st));
      h->ipa.push_back(L);
    }
  }
  for (int i = 0; i < nl; ++i) {
    MainLayerW L;
    const std::string p = "layers." + std::to_string(i);
    L.ada_off = off;
    MD_TRY(put_ada(p + ".adaLN_modulation.1", 9 * C));
    MD_TRY(pack_mha(h, p + ".mha_l", &L.mha_l, st));
    MD_TRY(pack_mha(h, p + ".mha_t", &L.mha_t, st));
    MD_TRY(pack_linear(h, p + ".fc1", &L.fc1, bfm, st));
    MD_TRY(pack_linear(h, p + ".fc2", &L.fc2, bfm, st));
    h->layers.push_back(L);
  }
  h->ada_final_off = off;
  MD_TRY(put_ada("emb_to_latent.adaLN_modulation.1", 2 * C));
  MD_TRY(pack_linear(h, "emb_to_latent.linear", &h->final_lin, bfm, st));
  if (h->flag(MDGEN_F_DESIGN)) {
    MD_TRY(pack_linear(h, "fc1", &h->design_fc1, bfm, st));
    MD_TRY(pack_linear(h, "fc2", &h->design_fc2, false, st));
    MD_TRY(pack_linear(h, "fc3", &h->design_fc3, false, st));
    MD_TRY(pack_linear(h, "emb_to_logits", &h->design_logits, false, st));
  }
  h->finalized = true;
  return 0;
}

}  // namespace mdgen
