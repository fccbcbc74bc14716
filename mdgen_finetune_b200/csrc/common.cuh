// common.cuh -- shared helpers for the mdgen_b200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <string>
#include <vector>

namespace mdgen {

typedef __nv_bfloat16 bf16;

// ---- error plumbing (host) -----------------------------------------------------------
void set_error(const std::string& msg);
#define MD_CUDA(call)                                                                       \
  do {                                                                                      \
    cudaError_t e__ = (call);                                                               \
    if (e__ != cudaSuccess) {                                                               \
      mdgen::set_error(std::string(#call) + ": " + cudaGetErrorString(e__) + " @" +         \
                       __FILE__ + ":" + std::to_string(__LINE__));                          \
      return -2;                                                                            \
    }                                                                                       \
  } while (0)
#define MD_CHECK(cond, msg)                                                                 \
  do {                                                                                      \
    if (!(cond)) {                                                                          \
      mdgen::set_error(std::string(msg) + " (" #cond ") @" + __FILE__ + ":" +               \
                       std::to_string(__LINE__));                                           \
      return -1;                                                                            \
    }                                                                                       \
  } while (0)
#define MD_TRY(call)                                                                        \
  do {                                                                                      \
    int r__ = (call);                                                                       \
    if (r__ != 0) return r__;                                                               \
  } while (0)

// launch counter: every kernel launch of the library goes through MD_LAUNCHED()
extern thread_local long long g_launches;
#define MD_LAUNCHED() (++mdgen::g_launches)

// ---- per-DEVICE (not per-process) launcher state ---------------------------------------------------------------
// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) and the SM count belong to a device: a process that drives several
// GPUs (nn.DataParallel replicas, samplers.py:75-79; one handle per device) configures each kernel once per device.
int cur_device();
int num_sms();                                     // SM count of the current device (cached per device)
struct DevOnce {
  std::atomic<unsigned long long> mask{0};         // bit d set: configured on device d (d < 64)
  bool need(int dev) const { return dev < 0 || dev >= 64 || !((mask.load(std::memory_order_acquire) >> dev) & 1ull); }
  void done(int dev) { if (dev >= 0 && dev < 64) mask.fetch_or(1ull << dev, std::memory_order_release); }
};
#define MD_SMEM_ONCE(kern, bytes)                                                                           \
  do {                                                                                                      \
    static mdgen::DevOnce once__;                                                                           \
    const int dev__ = mdgen::cur_device();                                                                  \
    if (once__.need(dev__)) {                                                                               \
      MD_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)));       \
      once__.done(dev__);                                                                                   \
    }                                                                                                       \
  } while (0)

// Kernel-variant switches (A/B measurements, debugging): the environment is read ONCE PER C-ABI
// CALL (read_switches() at the entry points), never per launch.
struct Switches {
  bool no_gemm_ln = false, no_gemm_2cta = false, no_gemm_ws = false, no_l2_persist = false, no_attn_tc = false,
       no_attn_bulk = false, gelu_2cta = false, no_graph = false, no_qkv_attn = false, no_pair16 = false, pair16_store = false, no_pdl = false, no_mlp_fused = false, no_gemm_ln4 = false;
  int attn_var = 0, attn_poly = 0, asb_seq = 4, attn_sb = 48, gemm_ln_min_m = 8192, mlp_min_m = 8192, ln4_min_m = 1024;
};
extern thread_local Switches g_sw;
void read_switches();

// launches per kernel variant (mdgen_variant_launches): tests assert that the full-size cases reach the timed kernels
enum Variant : int { V_GEMM_LN = 0, V_GEMM_TC2, V_GEMM_TC, V_ATTN_TC, V_ATTN_SEQ4, V_GRAPH_REPLAY, V_QKV_ATTN, V_GEMM_MLP, V_GEMM_LN4, V_COUNT };
extern std::atomic<long long> g_variant[V_COUNT];
#define MD_VARIANT(v) (mdgen::g_variant[v].fetch_add(1, std::memory_order_relaxed))

// optional per-kernel-family timing with CUDA events on the launching stream (bench.py roofline)
enum ProfCat : int { PC_GEMM_TC = 0, PC_GEMM_SIMT, PC_ATTN_TC, PC_ATTN_SIMT, PC_ATTN_SHORT, PC_LN, PC_EMBED, PC_IPA,
                     PC_ADALN, PC_MISC, PC_QKV_ATTN, PC_COUNT };
struct ProfRec { int cat; cudaEvent_t a, b; };
struct Profiler {
  bool on = false;
  std::vector<ProfRec> recs;
};
extern thread_local Profiler g_prof;
struct ProfScope {
  cudaStream_t st;
  bool active;
  ProfRec r;
  ProfScope(int cat, cudaStream_t s) : st(s), active(g_prof.on) {
    if (!active) return;
    r.cat = cat;
    cudaEventCreate(&r.a);
    cudaEventCreate(&r.b);
    cudaEventRecord(r.a, st);
  }
  ~ProfScope() {
    if (!active) return;
    cudaEventRecord(r.b, st);
    g_prof.recs.push_back(r);
  }
};

// ---- programmatic dependent launch ---------------------------------------------------
// The per-evaluation kernels are launched with cudaLaunchAttributeProgrammaticStreamSerialization (also inside the captured
// rollout graph, where the attribute becomes a programmatic edge): kernel N+1 may become resident -- barrier init, TMEM
// allocation, tensor-map prefetch -- while kernel N is still running, and blocks in pdl_wait() until N has COMPLETED and
// its memory is visible.  Rules every such kernel follows:
//   * pdl_wait() comes before the first access to global memory (reads AND writes: write-after-read on reused scratch);
//   * pdl_trigger() comes after the kernel's own TMEM allocation (a dependent that became resident on the same SM and
//     allocated first would otherwise starve it while waiting for it: deadlock);
//   * every thread executes pdl_wait(), so "N+1 complete" implies "N complete" and the ordering is transitive.
// A kernel launched WITHOUT the attribute (plain <<<>>>) is an ordinary full dependency on both sides.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = g_sw.no_pdl ? 0 : 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// ---- device helpers ------------------------------------------------------------------
__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// exact erf GELU (layers.py:78-85)
__device__ __forceinline__ float gelu_erf(float x) {
  return x * 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
}
__device__ __forceinline__ float silu(float x) { return x / (1.0f + expf(-x)); }

// epilogue kinds shared by the SIMT and the tcgen05 GEMM
enum Epi : int {
  EPI_STORE = 0,       // out = acc + bias
  EPI_GELU = 1,        // out = gelu(acc + bias)
  EPI_GATE_RESID = 2,  // resid[m,n] += gate[batch(m)][n] * (acc + bias)     (fp32 resid)
  EPI_RESID = 3,       // resid[m,n] += acc + bias
  EPI_EULER = 4        // v = acc + bias; out[m,n] = v; state[m,n] += dt * v  (fp32)
};

struct EpiParams {
  const float* bias = nullptr;   // [N] or null
  void* out = nullptr;           // [M, ldo] of OT (STORE/GELU/EULER: velocity out, fp32)
  int ldo = 0;
  float* resid = nullptr;        // [M, ldr] fp32 (GATE_RESID / RESID) or Euler state
  int ldr = 0;
  const float* gate = nullptr;   // gate[b * gate_stride + n]
  int gate_stride = 0;
  int rows_per_batch = 1;        // batch(m) = (m / rows_per_batch) % nbatch
  int nbatch = 1;
  float dt = 0.f;
  long long* dbg = nullptr;      // kernel-developer clock stamps of CTA dbg_cta (mdgen_debug_set_buffer), normally null
  int dbg_cta = 0;
  int staged = 0;                // set by launch_gemm_tc: pointers / pitches allow the coalesced 128-bit epilogue
};

}  // namespace mdgen
