/* mdgen_b200.h -- C ABI of the B200-native MDGen denoiser (libmdgen_b200.so).
 *
 * The reference (lucascimeca/mdgen_finetune) has no FFI layer: its boundary for this path is
 * the Python API of LatentMDGenModel / Sampler.  Each entry point below
 * names the reference interface it replaces.  All pointers are plain device (or, where stated,
 * host) pointers; no torch types cross this boundary.  Every call returns 0 on success and a
 * negative MDGEN_E* code otherwise; mdgen_last_error() returns the message of the last failure
 * on the calling thread.  A handle is bound to the CUDA device that was current at
 * mdgen_create(); calls are asynchronous on the stream passed in.  mdgen_euler_rollout replays ONE CUDA graph of the
 * whole solve (captured on first use per shape / step count, MDGEN_NO_GRAPH=1 disables it); under a caller's own
 * stream capture it emits the plain launch sequence instead (no allocation, copy from host or synchronisation in the
 * step; one un-captured call of the same shape must have run first) -- tests/test_graph.py covers both.
 */
#ifndef MDGEN_B200_H
#define MDGEN_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct MdgenHandle MdgenHandle;
typedef void* mdgen_stream_t; /* cudaStream_t */

enum {
  MDGEN_OK = 0,
  MDGEN_EINVAL = -1,   /* bad argument / shape (reference: assert / ValueError)            */
  MDGEN_ECUDA = -2,    /* CUDA runtime error                                               */
  MDGEN_ESTATE = -3,   /* call order violated (weights not finalised, no conditioning set) */
  MDGEN_ENOTFOUND = -4 /* unknown weight name                                              */
};

/* task flags: the argparse flags LatentMDGenModel reads (src/mdgen/parsing.py) */
enum {
  MDGEN_F_SIM = 1 << 0,         /* args.sim_condition                     */
  MDGEN_F_TPS = 1 << 1,         /* args.tps_condition                     */
  MDGEN_F_INPAINT = 1 << 2,     /* args.inpainting                        */
  MDGEN_F_DESIGN = 1 << 3,      /* args.design                            */
  MDGEN_F_NO_AA_EMB = 1 << 4,   /* args.no_aa_emb                         */
  MDGEN_F_ABS_POS = 1 << 5,     /* args.abs_pos_emb                       */
  MDGEN_F_ABS_TIME = 1 << 6,    /* args.abs_time_emb                      */
  MDGEN_F_NO_ROPE = 1 << 7,     /* args.no_rope                           */
  MDGEN_F_PREPEND_IPA = 1 << 8  /* args.prepend_ipa                       */
};

enum { MDGEN_PREC_BF16 = 0, MDGEN_PREC_FP32_CHECK = 1 };

/* Replaces the constructor arguments LatentMDGenModel(args, latent_dim)
 * (src/mdgen/model/latent_model.py:43-126). */
typedef struct MdgenConfig {
  int32_t embed_dim;      /* args.embed_dim   (384) */
  int32_t num_layers;     /* args.num_layers  (5)   */
  int32_t mha_heads;      /* args.mha_heads   (16)  */
  int32_t ipa_heads;      /* args.ipa_heads   (4)   */
  int32_t ipa_head_dim;   /* args.ipa_head_dim(32)  */
  int32_t ipa_qk;         /* args.ipa_qk      (8)   */
  int32_t ipa_v;          /* args.ipa_v       (8)   */
  int32_t latent_dim;     /* latent_dim (21 / 28 / 48) */
  int32_t crop;           /* args.crop (rows of pos_embed) */
  int32_t num_frames;     /* args.num_frames (rows of time_embed) */
  int32_t flags;          /* MDGEN_F_* */
  int32_t precision;      /* MDGEN_PREC_* : bf16 tensor-core path or fp32 check mode */
  float time_multiplier;  /* args.time_multiplier (100) */
} MdgenConfig;

/* lifetime ---------------------------------------------------------------------------- */
int mdgen_create(const MdgenConfig* cfg, MdgenHandle** out);
int mdgen_destroy(MdgenHandle* h);
const char* mdgen_last_error(void);
int mdgen_abi_version(void);

/* weights: one call per state_dict entry (names and shapes of SURVEY 8b; replaces
 * nn.Module.load_state_dict, latent_model.py:50-126).  `data` is fp32, device memory if
 * is_device != 0 else host memory; it is copied (the caller keeps ownership). */
int mdgen_load_weight(MdgenHandle* h, const char* name, const float* data, const int64_t* shape,
                      int32_t ndim, int32_t is_device, mdgen_stream_t stream);
/* Packs the loaded weights for the kernels (bf16 copies, fused QKV, folded q scaling).
 * Must be called after the last mdgen_load_weight and again after any weight update. */
int mdgen_finalize_weights(MdgenHandle* h, mdgen_stream_t stream);

/* Step-invariant conditioning of one batch: everything forward() receives besides (x, t)
 * (latent_model.py:216-220).  Replaces the per-call recomputation of
 * pos_embed + cond_to_emb(x_cond) + mask_to_emb(x_cond_mask) (latent_model.py:243-250)
 * and the frame inputs of run_ipa (latent_model.py:179-214).
 *   mask        [B,T,L] f32 (1 = valid)
 *   start_rot   [B,L,3,3] f32, start_trans [B,L,3] f32    (Rigid start_frames)
 *   end_rot/end_trans: same for end_frames, NULL unless TPS / inpainting
 *   rel_f7/rel_r7 [B,L,7] f32: to_tensor_7 of start^-1 o end and end^-1 o start
 *       (latent_model.py:198-199); NULL => computed on device with the closed-form
 *       quaternion (w >= 0) instead of the reference's sign-ambiguous eigh
 *   x_cond      [B,T,L,cond_dim] f32 or NULL;  x_cond_mask [B,T,L] i64 or NULL
 *   aatype      [B,L] i64 or NULL
 * All device pointers; borrowed only for the duration of the call. */
int mdgen_set_conditioning(MdgenHandle* h, int32_t B, int32_t T, int32_t L, const float* mask,
                           const float* start_rot, const float* start_trans,
                           const float* end_rot, const float* end_trans, const float* rel_f7,
                           const float* rel_r7, const float* x_cond, const int64_t* x_cond_mask,
                           const int64_t* aatype, mdgen_stream_t stream);

/* One velocity evaluation: LatentMDGenModel.forward / forward_inference
 * (latent_model.py:216-269, 272-278).  x [B,T,L,D] f32, t [B] f32, out [B,T,L,D] f32. */
int mdgen_forward(MdgenHandle* h, const float* x, const float* t, float* out,
                  mdgen_stream_t stream);

/* Fixed-grid Euler sampler: Sampler.sample_ode(sampling_method="euler", num_steps)
 * (transport.py:407-450, integrators.py:74-114, torchdiffeq fixed-grid Euler).
 * grid = linspace(0,1,num_steps); num_steps-1 velocity evaluations.
 *   x0 [B,T,L,D] f32; out_final [B,T,L,D] f32; traj: NULL or [num_steps,B,T,L,D] f32
 * (the reference returns the whole stack, callers take [-1], wrapper.py:478). */
int mdgen_euler_rollout(MdgenHandle* h, const float* x0, int32_t num_steps, float* traj,
                        float* out_final, mdgen_stream_t stream);

/* Gradients of one velocity evaluation (row N4): what autograd does when the reference trains through the module --
 * `policy.unet(x, t_, **condition)` with gradients in HGFNode.run_policy (src/models/samplers.py:120-146), the RTB loss
 * (src/rtb_utils/gfn_diffusion.py:459-460) and Transport.training_losses (src/mdgen/transport/transport.py:173-188).
 * Given dout = dLoss/d(forward output) [B,T,L,D] it RECOMPUTES LatentMDGenModel.forward (latent_model.py:216-269, the
 * raw forward: design logits added, no Dirichlet flow) in fp32 for the conditioning last set, and back-propagates:
 *   dx     [B,T,L,D] f32 or NULL: receives dLoss/dx (overwritten)
 *   names / grads [nparams]: reference state_dict names and device fp32 buffers of the parameter's shape; the parameter
 *       gradients are ACCUMULATED (+=) into them (zero them first for a fresh gradient).  Parameters not listed are
 *       treated as frozen (no gradient computed).  q/k/v projections, bias_k / bias_v, IPA weights incl. head_weights,
 *       LayerNorm affine, embeddings (aatype_to_emb.weight / mask_to_emb.weight as tables), adaLN and timestep MLP.
 *   out    [B,T,L,D] f32 or NULL: the recomputed fp32 forward output
 * All arithmetic is fp32 regardless of the handle's precision (a bf16 handle: the gradient is that of the fp32 function,
 * within the bf16 tolerance of the bf16 forward).  The first call of a shape allocates the activation tape. */
int mdgen_backward(MdgenHandle* h, const float* x, const float* t, const float* dout, float* dx, int32_t nparams,
                   const char* const* names, float* const* grads, float* out, mdgen_stream_t stream);

/* Design mode: turns the logits in the last 20 channels of `latent` [B,T,L,D] (what mdgen_forward wrote) into the
 * Dirichlet conditional-flow velocity, in place: the tail of LatentMDGenModel.forward_inference
 * (latent_model.py:296-326) with DirichletConditionalFlow.c_factor (utils.py:45-57) evaluated on the device instead
 * of through .cpu().numpy().  x [B,T,L,D] f32 (device) = the state the velocity was evaluated at; ifunc [1000] f64
 * (device) = row of beta_cdfs_derivative nearest to alpha (utils.py:53); beta = B(alpha, 19); temp =
 * args.dirichlet_flow_temp; dalpha = alpha_max - 1 (transport.py:52-57).  flags_out [2] i32 (host): tokens of x off
 * the simplex (the reference prints a warning), NaN c-factors (the reference raises unless allow_nan_cfactor; with
 * allow_nan != 0 they are replaced like torch.nan_to_num).  Synchronises the stream to read the flags. */
int mdgen_dirichlet_flow(MdgenHandle* h, const float* x, float* latent, const double* ifunc, double alpha,
                         double beta, float temp, float dalpha, int32_t allow_nan, int32_t* flags_out,
                         mdgen_stream_t stream);

/* Design-mode sampler: the tables mdgen_euler_rollout needs to apply the Dirichlet flow at every step of the next
 * rollout (nt = num_steps - 1).  HOST pointers: rows [nt, 1000] f64 (the beta_cdfs_derivative row nearest to alpha_i,
 * utils.py:53), alphas [nt], betas [nt] (B(alpha_i, 19)); temp / dalpha / allow_nan as in mdgen_dirichlet_flow.
 * The upload is asynchronous on `stream` (ordered after a rollout still in flight there) from a handle-owned staging
 * copy, so the host buffers are free on return; unchanged tables are not uploaded again.
 * After the rollout mdgen_dirichlet_flags returns the counters accumulated over all steps (off-simplex tokens,
 * NaN c-factors): the reference warns / raises at the step where they occur, the fused rollout at its end. */
int mdgen_set_dirichlet_tables(MdgenHandle* h, int32_t nt, const double* rows, const double* alphas,
                               const double* betas, float temp, float dalpha, int32_t allow_nan,
                               mdgen_stream_t stream);
int mdgen_dirichlet_flags(MdgenHandle* h, int32_t* flags_out, mdgen_stream_t stream);

/* NewMDGenWrapper.prep_batch (wrapper.py:283-365), the step before the sampler, on the device and stateless:
 * trans [B,T,L,3], rots [B,T,L,3,3], torsions [B,T,L,14] f32 -> latents / x_cond [B,T,L,D] f32 with
 * D = 7 * (1 + two_ref) + 14 and x_cond_mask [B,T,L] i64.  flags: 1 second reference frame (tps / inpainting /
 * dynamic_mpnn), 2 no_offsets, 4 no_torsion, 8 no_design_torsion, 16 design_key_frames, 32 sim_condition,
 * 64 tps_condition, 128 inpainting-style residue conditioning; cond_interval as args.cond_interval (0 = off). */
int mdgen_prep_batch(int32_t B, int32_t T, int32_t L, const float* trans, const float* rots,
                     const float* torsions, int32_t flags, int32_t cond_interval, float* latents,
                     float* x_cond, int64_t* cond_mask, mdgen_stream_t stream);

/* The decode step of NewMDGenWrapper.inference (wrapper.py:487-508), the step after the sampler: samples [B,T,L,D] f32 ->
 * atom14 [B,T,L,14,3] f32 (frames_torsions_to_atom14, geometry.py:61-79,236-334) and aa_out [B,T,L] i64 (NULL to skip).
 * tors_off = first torsion channel (7, or 14 with tps / inpainting); rot0 [B,L,3,3] / trans0 [B,L,3] = rigids[:, 0]
 * (NULL with flag 1 = no_offsets); seqres [B,L] i64.  The caller supplies the four mdgen.residue_constants tables:
 * default_frames [21,8,4,4] f32, group_idx [21,14] i64, atom_mask [21,14] f32, lit_positions [21,14,3] f32.
 * flags: 1 no_offsets, 2 --oracle (torsions not re-normalised), 4 design (aa_out = argmax of the last 20 channels). */
int mdgen_decode_atom14(int32_t B, int32_t T, int32_t L, int32_t D, const float* samples, int32_t tors_off,
                        const float* rot0, const float* trans0, const int64_t* seqres,
                        const float* default_frames, const int64_t* group_idx, const float* atom_mask,
                        const float* lit_positions, int32_t flags, float* atom14, int64_t* aa_out,
                        mdgen_stream_t stream);

/* introspection ------------------------------------------------------------------------ */
/* Number of kernels one mdgen_forward launches for the current conditioning shape. */
int mdgen_launches_per_forward(MdgenHandle* h);
/* Total kernels launched by the library on the calling thread since load (graph replays count their kernel nodes). */
long long mdgen_launch_count(void);
/* Launches per kernel variant since load: 0 gemm_ln (fused projection + residual + LayerNorm), 1 gemm_tc2 (CTA pair),
 * 2 gemm_tc, 3 attention_tc, 4 attention_seq4 (residue axis, 5 keys), 5 whole-rollout graph replays, 6 qkv_attn4 (q/k/v
 * projection fused with the residue-axis attention at L = 4), 7 gemm_mlp (fc1 + GELU + fc2 + residual + LayerNorm in one
 * kernel), 8 gemm_ln4 (projection + residual + LayerNorm for fewer than 8192 rows, row split over a cluster of 4 CTAs);
 * -1 if out of range.
 * Launches captured into a rollout graph are counted once, at capture. */
long long mdgen_variant_launches(int32_t which);
/* Bytes of device workspace currently held by the handle. */
int64_t mdgen_workspace_bytes(MdgenHandle* h);

/* Per-kernel-family device timing (CUDA events on the launching stream) for bench.py's roofline:
 * between begin and end every launch of the library is bracketed by two events.  ncat <= 11. */
int mdgen_profile_begin(void);
int mdgen_profile_end(float* ms_by_cat, int32_t* launches_by_cat, int32_t ncat);
const char* mdgen_profile_category_name(int32_t i);

/* Kernel-developer hook: when device_buf != NULL, the attention kernel's CTA `cta` writes per-role
 * clock64() stamps into device_buf[0..95] (tools/attn_probe.py).  NULL (default) disables it. */
int mdgen_debug_set_buffer(long long* device_buf, int32_t cta);

/* unit-test entry points (tests/ only): run one kernel family on caller buffers --------- */
/* C[M,N] = A[M,K] . W[N,K]^T + bias, A/W fp32 on device (converted to bf16 inside when
 * use_tc != 0 -> tcgen05 path; fp32 SIMT path otherwise).  tcgen05 epilogue variants: use_tc = 1 fp32 store,
 * 2 bf16 store, 3 bf16 erf-GELU (both converted back to fp32 into C), 4 C[m,n] += bias[n] * (acc + bias[n])
 * (the gated-residual epilogue with gate = bias), 5 the fused projection + gated residual + LayerNorm/modulate kernel
 * (N = 384): bias = [linear bias (N) | gate (2,N) | shift (2,N) | scale (2,N)], batch element (m / 200) % 2,
 * C = [2M, N]: rows [0,M) residual stream in/out, rows [M,2M) the normalised rows. */
int mdgen_test_gemm(int32_t M, int32_t N, int32_t K, const float* A, const float* W,
                    const float* bias, float* C, int32_t use_tc, int32_t block_n,
                    mdgen_stream_t stream);
/* Attention core as reached from MultiheadAttention (mha.py:258-397) on packed qkv:
 * qkv [rows, 3C] f32 (q already scaled), rows = nb * S * inner, token (b, s, i) at row
 * (b*S + s)*inner + i  (inner = 1 ... residue axis: L tokens contiguous; frame axis: stride L).
 * bias_k/bias_v [C]; key_valid [nb*S*inner] f32 in the same row order; out [rows, C] f32. */
int mdgen_test_attention(int32_t nb, int32_t S, int32_t inner, int32_t heads, int32_t head_dim,
                         const float* qkv, const float* bias_k, const float* bias_v,
                         const float* key_valid, float* out, int32_t use_tc, int32_t rope,
                         mdgen_stream_t stream);

/* Allocation-free GEMM.  NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * C[M,N] (fp32) = bf16(A[M,K]) . bf16(W[N,K])^T + bias[N] (bias may be NULL) on the tcgen05 GEMM -- the arithmetic of
 * mdgen_test_gemm(use_tc = 1) -- with the bf16 operand copies held in a file-scope `static __device__` array of
 * MDGEN_NOALLOC_SCRATCH_BYTES (8 MiB; every GPU that loads the library holds it until the process exits) instead of
 * cudaMalloc'ed scratch.  All pointers are device pointers on the CURRENT device; K % 8 == 0.
 * Limit: M*K <= MDGEN_NOALLOC_MAX_OPERAND_ELEMS and N*K <= MDGEN_NOALLOC_MAX_OPERAND_ELEMS (2^21 elements each, e.g.
 * 5461 x 384 or 1365 x 1536); larger problems return MDGEN_EINVAL with a message naming the limit.
 * The call performs no cudaMalloc / cudaFree / allocating library call, does not synchronise, and is asynchronous on
 * `stream` (capturable).  ALL calls share the one array, which keeps the previous call's contents (nothing in it has to
 * start at zero: every element read is rewritten first, on the same stream).  Calls on one stream are ordered; calls
 * that overlap on the GPU (different streams, or replays of graphs that contain it, running concurrently) RACE on the
 * scratch -- serialise them. */
#define MDGEN_NOALLOC_MAX_OPERAND_ELEMS (1 << 21)
#define MDGEN_NOALLOC_SCRATCH_BYTES (2 * MDGEN_NOALLOC_MAX_OPERAND_ELEMS * 2)
int mdgen_gemm_noalloc(int32_t M, int32_t N, int32_t K, const float* A, const float* W, const float* bias,
                       float* C, mdgen_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* MDGEN_B200_H */
