// attention_tc.cuh -- tcgen05 attention core for the long axes (frame axis T of the 4AA configs, 257 / 251
// keys of ATLAS).  Semantics = attention_simt.cuh (mha.py:258-397): bias-KV token as key S, RoPE on q/k,
// key-padding -> -inf, fp32 softmax statistics, P.V.
//
// head_dim = 24 makes this kernel exp- and issue-bound, not tensor-bound: one ex2, one FFMA, half a bf16 pack and
// one 4-byte TMEM read per 96 MMA FLOPs.  ncu on the previous versions showed no saturated pipe (XU 54 %, issue
// 52 %) but softmax warps waiting on each other through per-block barriers, so the design minimises instructions
// and synchronisation per key and never lets the softmax wait for operands:
//   * PERSISTENT CTAs (one per SM) walk a static list of work units; a unit = one (sequence, head, group of up
//     to four 128-query tiles).  Four PRODUCER warps stream the unit's K (RoPE'd) and V through a ring of
//     96-key operand blocks in shared memory (no-swizzle UMMA layout [d group of 8][96 keys][8]; V is consumed
//     MN-major, so no transpose is ever materialised) and stage the Q tiles of the NEXT unit into the second Q
//     buffer of each warpgroup -- staging is completely hidden behind the softmax of the current unit.
//     d is zero-padded 24 -> 32; padding column 24 of V is 1, which makes O[:, 24] the softmax row sum for free.
//     The padding groups are written once per CTA, never restaged.
//   * single pass over key blocks of 48: S = Q.K^T (128x48 fp32 in TMEM) is DOUBLE-BUFFERED per warpgroup
//     (2 x 48 columns) and P = exp2(S log2e - m), packed bf16, overwrites the first 24 columns of the S buffer it
//     came from (a thread only ever touches its own TMEM lane, so no cross-warp hazard).  O += P.V takes P as the
//     TMEM A operand.  The tensor pipe executes one thread's MMAs in order, so the issuer simply queues
//     P.V(j) ; Q.K^T(j+2) behind each "P(j) ready": no S-empty / P-consumed barriers exist, the softmax does ONE
//     wait (S ready) and ONE arrive (P ready) per block and S(j+1) is computed while it works on S(j).
//     Online softmax at block granularity with lazy rescaling: O is rescaled only when the running max grows by
//     more than 2^8.
//   * FOUR softmax warpgroups (warps 4-19; thread = query row) work on four query tiles at once with their own
//     S0 / S1 / O TMEM columns (4 x (48 + 48 + 32) = all 512), i.e. four warps per SM sub-partition, so that one
//     warp's MUFU phase overlaps the others' TMEM-read / max / store phases.  Lane 0 of warp w (w < 4) issues the
//     tcgen05.mma of warpgroup w (blocking mbarrier waits, no polling) and hands ring slots / Q buffers back to
//     the producers with tcgen05.commit.
//   * registers follow the roles (setmaxnreg): 40 for the MMA issuers, 56 for the producers, 96 for the softmax.
#pragma once
#include "attention_simt.cuh"
#include "gemm_tc.cuh"

namespace mdgen {
namespace atc {

using tc::elect_one;
using tc::mbar_arrive;
using tc::mbar_init;
using tc::mbar_wait;
using tc::smem_u32;
using tc::tcgen05_commit;
using tc::tcgen05_fence_after;
using tc::tcgen05_fence_before;
using tc::tmem_ld_32x32b_x32;
using tc::tmem_ld_wait;
using tc::umma_bf16;

constexpr int HD = 24, DP = 32;          // head dim and its MMA-padded width
constexpr int QT = 128;                  // query tile rows
constexpr int KB = 128;                  // rows of a Q operand block
constexpr int RB = 96;                   // keys (rows) per K/V ring block = 2 softmax blocks of 48 or 3 of 32
constexpr int MAXBUF = 3;                // S buffers per warpgroup: 2 x 48 or 3 x 32 columns (both 96)
// Kernel variants <SB, NBUF> (keys per MMA / softmax block, S buffers in flight per warpgroup):
//   <48, 2>  v7: Q.K^T(j+2) is queued behind "P(j) ready" -- one block of slack for the round trip softmax -> issuer ->
//            tensor pipe -> commit -> softmax, which the clock stamps (tools/attn_probe.py) show exposed: 400-1000 of every
//            ~2200 cycles per block the softmax warps wait for S while the MUFU, saturated in between, idles.
//   <32, 3>  v8: the same 96 S columns as THREE buffers of 32 keys: Q.K^T(j+3) behind P(j), two blocks of slack.
constexpr int MAX_KEYS = 2048;           // S + 1 <= MAX_KEYS (K/V are streamed; the limit only bounds what is tested)
constexpr int NWG = 4;                   // softmax warpgroups = query tiles in flight per CTA
constexpr int NPW = 4;                   // producer warps (thread = one key row / one query row of a block)
constexpr int NSLOT = 12;                // K/V ring blocks in flight (a 1001-key unit uses 11)
constexpr int MMA_WARP0 = 0, SOFT_WARP0 = 4, PROD_WARP0 = 4 + 4 * NWG;
constexpr int THREADS = 32 * (PROD_WARP0 + NPW);
constexpr int REGS_MMA = 40, REGS_SOFT = 96, REGS_PROD = 56;
static_assert(128 * REGS_MMA + NWG * 128 * REGS_SOFT + NPW * 32 * REGS_PROD <= THREADS * 80, "register pool (launch bound: 80 / thread)");
constexpr int BLK_BYTES = 4 * KB * 16;   // Q operand block: [4 d-groups][128 rows][8 bf16] = 8 KB
constexpr int GRP_BYTES = KB * 16;       // 2048: stride between d-groups of a Q block (K-major LBO)
constexpr int RGRP_BYTES = RB * 16;      // 1536: stride between d-groups of a K/V ring block (K-major LBO / MN-major SBO)
constexpr int RBLK_BYTES = 4 * RGRP_BYTES;   // 6 KB
constexpr int OFF_K = 0, OFF_V = OFF_K + NSLOT * RBLK_BYTES, OFF_Q = OFF_V + NSLOT * RBLK_BYTES,
              OFF_BAR = OFF_Q + 2 * NWG * BLK_BYTES, SMEM_BYTES = OFF_BAR + 1024;
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
constexpr uint32_t TMEM_COLS = 512;
// per-warpgroup TMEM columns: S buffers fp32 [96 in all] (P packed bf16 overwrites the first half of its S), O fp32 [32]
constexpr uint32_t TM_WG = 128, TM_S = 0, TM_O = 96;
constexpr float RESCALE_TAU = 8.0f;      // log2 units: P <= 2^8 between rescales

struct Bars {
  uint64_t s_full[NWG][MAXBUF];    // Q.K^T of a block has landed in S buffer b (and every MMA queued before it has retired)
  uint64_t p_full[NWG][MAXBUF];    // the warpgroup has written P over S buffer b
  uint64_t q_full[2][NWG], q_empty[2][NWG];      // per warpgroup: two Q buffers
  uint64_t kv_full[NSLOT], kv_empty[NSLOT];      // K/V ring
  uint64_t o_free[NWG];       // the warpgroup has consumed the unit's last two s_full phases: the next unit may signal again
  uint32_t tmem_base;
};
static_assert(sizeof(Bars) <= 1024, "barrier block");

// no-swizzle descriptor: start>>4 | LBO>>4 << 16 | SBO>>4 << 32 | version 1 << 46 | layout 0
__device__ __forceinline__ uint64_t desc_nosw(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) |
         ((uint64_t)1 << 46);
}
// bf16 x bf16 -> f32, A K-major, B K-major or MN-major (bit 16)
__host__ __device__ constexpr uint32_t idesc(int M, int N, bool b_mn_major) {
  return tc::make_idesc_bf16(M, N) | (b_mn_major ? (1u << 16) : 0u);
}
// D[tmem] (+)= A[tmem] . B[smem desc]
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t id, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(id), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x32b_x16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t* r) {
  tmem_st_32x32b_x16(taddr, r);
  tmem_st_32x32b_x16(taddr + 16, r + 16);
}
__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
// non-blocking phase test (the MMA issuer polls several warpgroups)
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float max3(float a, float b, float c) {   // one FMNMX3
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// optional per-role cycle stamps of one CTA (tools/attn_probe.py); dbg == nullptr in production
constexpr int DBG_SLOTS = 256;
__device__ __forceinline__ void dbg_stamp(long long* dbg, int slot) {
  if (dbg) dbg[slot] = clock64();
}

// rotate one 24-wide row held as 3 x (8 bf16) groups: pairs (i, i+12), fair-esm rotate_half
__device__ __forceinline__ void rope_row(uint4* g, const float* __restrict__ cs, const float* __restrict__ sn) {
  float f[HD];
  const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(g);
#pragma unroll
  for (int i = 0; i < HD / 2; ++i) {
    const float2 v = __bfloat1622float2(h2[i]);
    f[2 * i] = v.x;
    f[2 * i + 1] = v.y;
  }
  const float4* c4 = reinterpret_cast<const float4*>(cs);
  const float4* s4 = reinterpret_cast<const float4*>(sn);
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const float4 c = __ldg(c4 + j), s = __ldg(s4 + j);
    const float cc[4] = {c.x, c.y, c.z, c.w}, ss[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int i = 4 * j + e;
      const float a = f[i], b = f[i + HD / 2];
      f[i] = a * cc[e] - b * ss[e];
      f[i + HD / 2] = b * cc[e] + a * ss[e];
    }
  }
  uint32_t* o = reinterpret_cast<uint32_t*>(g);
#pragma unroll
  for (int i = 0; i < HD / 2; ++i) o[i] = pack_bf16(f[2 * i], f[2 * i + 1]);
}

// same rotation with the cos (cst[0..2]) / sin (cst[3..5]) rows already in registers
__device__ __forceinline__ void rope_row_regs(uint4* g, const float4* cst) {
  float f[HD];
  const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(g);
#pragma unroll
  for (int i = 0; i < HD / 2; ++i) {
    const float2 v = __bfloat1622float2(h2[i]);
    f[2 * i] = v.x;
    f[2 * i + 1] = v.y;
  }
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const float cc[4] = {cst[j].x, cst[j].y, cst[j].z, cst[j].w};
    const float ss[4] = {cst[3 + j].x, cst[3 + j].y, cst[3 + j].z, cst[3 + j].w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int i = 4 * j + e;
      const float a = f[i], b = f[i + HD / 2];
      f[i] = a * cc[e] - b * ss[e];
      f[i + HD / 2] = b * cc[e] + a * ss[e];
    }
  }
  uint32_t* o = reinterpret_cast<uint32_t*>(g);
#pragma unroll
  for (int i = 0; i < HD / 2; ++i) o[i] = pack_bf16(f[2 * i], f[2 * i + 1]);
}

// one Q row: global -> RoPE -> [d group][row][8] operand block
__device__ __forceinline__ void stage_q_row(const AttnParams<bf16>& p, const bf16* qkv_h, uint8_t* qb, int r, int q,
                                            uint32_t seq_row0) {
  uint4 g[3] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
  if (q < p.S) {
    const uint4* src = reinterpret_cast<const uint4*>(qkv_h + (size_t)(seq_row0 + (uint32_t)q * p.inner) * p.ld);
    g[0] = __ldg(src);
    g[1] = __ldg(src + 1);
    g[2] = __ldg(src + 2);
    if (p.cs) rope_row(g, p.cs + q * (HD / 2), p.sn + q * (HD / 2));
  }
#pragma unroll
  for (int j = 0; j < 3; ++j) *reinterpret_cast<uint4*>(qb + j * GRP_BYTES + r * 16) = g[j];
}

__device__ __forceinline__ void tmem_st_32x32b_x8(uint32_t taddr, const uint32_t* r) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
               "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}

// exp2 on the FMA / ALU pipes (the MUFU unit does 16 ex2 / clk / SM and is this kernel's busiest pipe): round-to-nearest
// split x = n + f with the 1.5 * 2^23 trick, degree-3 minimax polynomial for 2^f on [-0.5, 0.5] (relative error
// 7.5e-5, far below the bf16 rounding of P), n added into the exponent field.  x <= 8 by construction (lazy
// rescaling); x = -inf (masked key) gives 2^-126 instead of 0, which vanishes against the row sum.
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -126.0f);
  const float t = x + 12582912.0f;
  const float f = x - (t - 12582912.0f);
  const float p = fmaf(fmaf(fmaf(0.0551716611f, f, 0.2426111251f), f, 0.6932609677f), f, 0.9999280572f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
// element j of a block takes the polynomial path when POLY > 0 and j % POLY == POLY - 1
template <int POLY> __device__ __forceinline__ float ex2_sel(float x, int j) {
  return (POLY > 0 && (j % (POLY > 0 ? POLY : 1)) == POLY - 1) ? ex2_poly(x) : ex2(x);
}

// One 48-key block of scores of one query row (x[0..47], fp32 bit patterns; padded keys arrive as ~ -32768 through
// the mask column of K, see the producers): running-max update with
// lazy rescale of O (in TMEM; `pv_bar` / `pv_parity` = the barrier phase that tells every P.V issued so far has
// retired), then P = exp2(x log2e - m) packed to bf16 (pk[0..23]).
template <bool DBG, int POLY, int SB>
__device__ __forceinline__ void soft_block(uint32_t (&x)[SB], uint32_t (&pk)[SB / 2], bool have_o,
                                           float& m_used, uint32_t tO, uint64_t* pv_bar, uint32_t pv_parity, int var) {
  const float LOG2E = 1.4426950408889634f;
  float bm[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
  for (int j = 0; j < SB; j += 2) bm[(j >> 1) & 3] = max3(bm[(j >> 1) & 3], __uint_as_float(x[j]), __uint_as_float(x[j + 1]));
  const float mb = fmaxf(fmaxf(bm[0], bm[1]), fmaxf(bm[2], bm[3])) * LOG2E;
  const bool grow = mb > m_used + RESCALE_TAU;           // false when both are -inf
  if (__any_sync(0xffffffffu, grow)) {                   // rare after the first blocks of a tile
    const float sc = grow ? ex2(m_used - mb) : 1.0f;     // m_used = -inf -> 0 (nothing accumulated yet)
    if (have_o) {
      mbar_wait(pv_bar, pv_parity);                      // every P.V issued so far has landed in O
      tcgen05_fence_after();
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        uint32_t o[16];
        tmem_ld_32x32b_x16(tO + hh * 16, o);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * sc);
        tmem_st_32x32b_x16(tO + hh * 16, o);
      }
    }
    if (grow) m_used = mb;
  }
  const float mneg = (m_used == -INFINITY) ? 0.f : -m_used;
  if (DBG && (var & 1)) {                                // timing experiment: no MUFU
#pragma unroll
    for (int j = 0; j < SB / 2; ++j)
      pk[j] = pack_bf16(fmaf(__uint_as_float(x[2 * j]), LOG2E, mneg), fmaf(__uint_as_float(x[2 * j + 1]), LOG2E, mneg));
    return;
  }
#pragma unroll
  for (int j = 0; j < SB / 2; ++j)
    pk[j] = pack_bf16(ex2_sel<POLY>(fmaf(__uint_as_float(x[2 * j]), LOG2E, mneg), 2 * j),
                      ex2_sel<POLY>(fmaf(__uint_as_float(x[2 * j + 1]), LOG2E, mneg), 2 * j + 1));
}

// work unit -> (sequence, head, query range)
struct Unit {
  int h, ntiles, q_begin;
  uint32_t seq_row0;
};
__device__ __forceinline__ Unit decode_unit(const AttnParams<bf16>& p, int unit, int chunks_per_seq) {
  Unit u;
  const int chunk = unit % chunks_per_seq, nbi = unit / (chunks_per_seq * p.H);
  u.h = (unit / chunks_per_seq) % p.H;
  u.q_begin = chunk * NWG * QT;
  u.ntiles = (min(p.S, u.q_begin + NWG * QT) - u.q_begin + QT - 1) / QT;
  u.seq_row0 = (uint32_t)(nbi / p.inner) * p.S * p.inner + (nbi % p.inner);   // row of key/query 0
  return u;
}
template <int N> __device__ __forceinline__ void setmaxnreg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void setmaxnreg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }


template <bool DBG, int POLY, int SB, int NBUF>
__global__ void __launch_bounds__(THREADS, 1)
attention_tc_kernel(AttnParams<bf16> p, int chunks_per_seq, int nunits, long long* dbg_buf, int dbg_cta, int dbg_var) {
  static_assert(SB * NBUF == 96 && RB % SB == 0 && NBUF <= MAXBUF && SB % 16 == 0, "S buffers fill 96 TMEM columns");
  constexpr int BPR = RB / SB;         // softmax blocks per K/V ring block
  const int var = DBG ? dbg_var : 0;   // timing experiments (tools/attn_probe.py); results are wrong when non-zero
  // No integer round trip on this pointer: the compiler must keep seeing a shared-space address, otherwise every
  // staging access becomes a generic LD.E / ST.E (long-scoreboard latency instead of LDS / STS).
  extern __shared__ __align__(128) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 127u) != 0) __trap();     // dynamic shared memory starts at the window base (no static shared)
  Bars* bars = reinterpret_cast<Bars*>(smem + OFF_BAR);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, tid = threadIdx.x;
  long long* dbg = (DBG && dbg_buf && (int)blockIdx.x == dbg_cta) ? dbg_buf : nullptr;
  if (tid == 0) dbg_stamp(dbg, 0);
  const int C = p.H * HD, S = p.S, nkeys = S + 1;
  const int nkb = (nkeys + RB - 1) / RB;                     // ring blocks per unit
  const int nsb = (nkeys + SB - 1) / SB;                     // MMA / softmax blocks of SB keys (>= NBUF: S >= 64)

  if (tid == 0) {
    for (int i = 0; i < NWG; ++i) {
      for (int b = 0; b < MAXBUF; ++b) {
        mbar_init(&bars->s_full[i][b], 1);
        mbar_init(&bars->p_full[i][b], 4);
      }
      for (int b = 0; b < 2; ++b) {
        mbar_init(&bars->q_full[b][i], NPW);
        mbar_init(&bars->q_empty[b][i], 1);
      }
    }
    for (int i = 0; i < NWG; ++i) mbar_init(&bars->o_free[i], 4);
    for (int i = 0; i < NSLOT; ++i) {
      mbar_init(&bars->kv_full[i], NPW);
      mbar_init(&bars->kv_empty[i], NWG);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&bars->tmem_base)), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // padding d-group (d = 24..31) of the V and Q operand blocks, written once: [1, 0, ..] for both.  V's ones column
  // makes O[:, 24] the row sum; Q's multiplies the MASK COLUMN of K (d = 24: 0 for an attendable key, -32768 for a
  // padded one, written per key by the producers), so key padding costs the softmax warps nothing.
  for (int i = tid; i < NSLOT * RB; i += THREADS) {
    const int blk = i / RB, r = i - blk * RB;
    *reinterpret_cast<uint4*>(smem + OFF_V + blk * RBLK_BYTES + 3 * RGRP_BYTES + r * 16) = make_uint4(0x3F80u, 0, 0, 0);
  }
  for (int i = tid; i < 2 * NWG * KB; i += THREADS)
    *reinterpret_cast<uint4*>(smem + OFF_Q + (i >> 7) * BLK_BYTES + 3 * GRP_BYTES + (i & 127) * 16) = make_uint4(0x3F80u, 0, 0, 0);
  fence_async_smem();
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = bars->tmem_base;
  pdl_trigger();       // common.cuh: after the TMEM allocation
  pdl_wait();          // qkv (the previous kernel's output) is complete from here on
  if (tid == 0) dbg_stamp(dbg, 1);
  const uint32_t sK = smem_u32(smem + OFF_K), sV = smem_u32(smem + OFF_V), sQ = smem_u32(smem + OFF_Q);

  if (warp < SOFT_WARP0) {
    // =============================== MMA issuers ===============================
    // Warp w walks the issue loop of warpgroup w warp-uniformly; one elected lane issues the tcgen05.mma / commits
    // (tc::elect_one: from a `lane == 0` branch every MMA costs ~10 dependent instructions).  Queue order per unit:
    //   QK(0) QK(1) | PV(0) QK(2) | PV(1) QK(3) | ... | PV(n-2) | PV(n-1)
    // each group behind "P(j) ready" and followed by a commit on s_full[j & 1], which therefore means both
    // "S(j+2) has landed" and "P.V(j) and everything before it has retired".
    setmaxnreg_dec<REGS_MMA>();
    const int w = warp;
    {
      uint32_t np0 = 0, np1 = 0, np2 = 0; // P-ready phases consumed per S buffer
      uint32_t gblk = 0, quse = 0;        // ring blocks consumed by earlier units, units this warpgroup was active in
      const uint32_t id_qk = idesc(QT, SB, false), id_pv = idesc(QT, DP, true);
      if (w == 0 && lane == 0) dbg_stamp(dbg, 14);
      for (int unit = blockIdx.x; unit < nunits; unit += gridDim.x, gblk += nkb) {
        const Unit u = decode_unit(p, unit, chunks_per_seq);
        if (w >= u.ntiles) {              // idle warpgroup: only hands the ring slots back
          for (int kb = 0; kb < nkb; ++kb) {
            const uint32_t g = gblk + kb, slot = g % NSLOT;
            mbar_wait(&bars->kv_full[slot], (g / NSLOT) & 1);
            if (elect_one()) mbar_arrive(&bars->kv_empty[slot]);
            __syncwarp();
          }
          continue;
        }
        const uint32_t qbuf = quse & 1;
        // an mbarrier phase must be consumed before the barrier completes again: the softmax warps have seen the
        // previous unit's closing s_full phases
        if (quse > 0) mbar_wait(&bars->o_free[w], (quse - 1) & 1);
        mbar_wait(&bars->q_full[qbuf][w], (quse >> 1) & 1);
        tcgen05_fence_after();
        const uint32_t qa = sQ + (qbuf * NWG + w) * BLK_BYTES;
        const uint32_t tW = tmem + w * TM_WG;
        auto issue_qk = [&](int j) {
          const int sub = j % BPR;                                 // position of the block inside its ring block
          const uint32_t g = gblk + j / BPR, slot = g % NSLOT;
          if (sub == 0) {                                          // first use of the ring block
            mbar_wait(&bars->kv_full[slot], (g / NSLOT) & 1);
            tcgen05_fence_after();
          }
          const uint32_t ka = sK + slot * RBLK_BYTES + sub * (SB * 16);
          if (elect_one()) {
            if (!(var & 8) || j < NBUF) {
#pragma unroll
              for (int jj = 0; jj < DP / 16; ++jj)
                umma_bf16(tW + TM_S + (j % NBUF) * SB, desc_nosw(qa + jj * 2 * GRP_BYTES, GRP_BYTES, 128),
                          desc_nosw(ka + jj * 2 * RGRP_BYTES, RGRP_BYTES, 128), id_qk, jj);
            }
            if (j == nsb - 1) tcgen05_commit(&bars->q_empty[qbuf][w]);   // every Q.K^T of the unit is queued: Q buffer free when they retire
          }
          __syncwarp();
        };
#pragma unroll
        for (int j0 = 0; j0 < NBUF; ++j0) {                        // prime every S buffer (nsb >= NBUF)
          issue_qk(j0);
          if (elect_one()) tcgen05_commit(&bars->s_full[w][j0]);
          __syncwarp();
        }
        for (int j = 0; j < nsb; ++j) {
          const int b = j % NBUF, sub = j % BPR;
          const uint32_t par = (b == 0 ? np0 : b == 1 ? np1 : np2) & 1;
          mbar_wait(&bars->p_full[w][b], par);
          if (b == 0) ++np0; else if (b == 1) ++np1; else ++np2;
          tcgen05_fence_after();
          const uint32_t g = gblk + j / BPR, slot = g % NSLOT;
          const uint32_t va = sV + slot * RBLK_BYTES + sub * (SB * 16);
          if (elect_one()) {
            if (!(var & 4)) {
#pragma unroll
              for (int kk = 0; kk < SB / 16; ++kk)      // 16 keys per MMA: 8 packed P columns, two 8-key groups of V
                umma_bf16_ts(tW + TM_O, tW + TM_S + b * SB + kk * 8, desc_nosw(va + kk * 256, 128, RGRP_BYTES), id_pv,
                             (j == 0 && kk == 0) ? 0u : 1u);
            }
            if (sub == BPR - 1 || j == nsb - 1) tcgen05_commit(&bars->kv_empty[slot]);   // this warpgroup is done with the ring block
          }
          __syncwarp();
          if (j + NBUF < nsb) issue_qk(j + NBUF);
          if (elect_one()) tcgen05_commit(&bars->s_full[w][b]);
          __syncwarp();
        }
        ++quse;
      }
      if (w == 0 && lane == 0) dbg_stamp(dbg, 15);
    }
  } else if (warp < PROD_WARP0) {
    // =============================== softmax warps ===============================
    setmaxnreg_inc<REGS_SOFT>();
    const int w = (warp - SOFT_WARP0) >> 2, quarter = warp & 3;
    const int r = quarter * 32 + lane;                       // query row in tile = TMEM lane
    const uint32_t tbase = tmem + w * TM_WG + ((uint32_t)(quarter * 32) << 16);
    const uint32_t tO = tbase + TM_O;
    uint64_t* const sfull = &bars->s_full[w][0];
    uint64_t* const pfull = &bars->p_full[w][0];
    const bool stamp = DBG && (r == 0 && w < 2);
    uint32_t sc0 = 0, sc1 = 0, sc2 = 0;                      // s_full phases consumed per S buffer
    uint32_t it = 0;
    for (int unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
      const Unit u = decode_unit(p, unit, chunks_per_seq);
      if (w < u.ntiles) {
        const int q = u.q_begin + w * QT + r;
        float m_used = -INFINITY;
        int b = 0;
        for (int j = 0; j < nsb; ++j, ++it) {
          uint32_t x[SB], pk[SB / 2];
          const uint32_t spar = (b == 0 ? sc0 : b == 1 ? sc1 : sc2) & 1;
          mbar_wait(sfull + b, spar);
          if (b == 0) ++sc0; else if (b == 1) ++sc1; else ++sc2;
          tcgen05_fence_after();
          if (stamp && it < 16) dbg_stamp(dbg, 16 + w * 40 + 2 * it);
          const uint32_t tS = tbase + TM_S + b * SB;
          tmem_ld_32x32b_x32(tS, x);
          if constexpr (SB == 48) tmem_ld_32x32b_x16(tS + 32, x + 32);
          tmem_ld_wait();
          // "P.V(j-1) has retired" = the next, not yet consumed phase of the S buffer block j-1 used (its commit follows
          // [P.V(j-1), Q.K^T(j-1+NBUF)])
          const int bp = b == 0 ? NBUF - 1 : b - 1;
          soft_block<DBG, POLY, SB>(x, pk, j > 0, m_used, tO, sfull + bp, (bp == 0 ? sc0 : bp == 1 ? sc1 : sc2) & 1, var);
          if (!(var & 16)) {
            tmem_st_32x32b_x16(tS, pk);
            if constexpr (SB == 48) tmem_st_32x32b_x8(tS + 16, pk + 16);
          }
          tmem_st_wait();
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(pfull + b);
          if (stamp && it < 16) dbg_stamp(dbg, 17 + w * 40 + 2 * it);
          b = (b + 1 == NBUF) ? 0 : b + 1;
        }
        // ---- epilogue: every S buffer signals once more (the last NBUF P.V retired); O / rowsum -> bf16 -> global.
        // The next unit's P.V cannot start before this warpgroup has stored P of its first block, i.e. after O
        // has been read here.
        mbar_wait(sfull + 0, sc0 & 1);
        mbar_wait(sfull + 1, sc1 & 1);
        ++sc0; ++sc1;
        if constexpr (NBUF == 3) {
          mbar_wait(sfull + 2, sc2 & 1);
          ++sc2;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->o_free[w]);
        tcgen05_fence_after();
        uint32_t o[32];
        tmem_ld_32x32b_x32(tO, o);
        tmem_ld_wait();
        tcgen05_fence_before();
        if (stamp && unit == (int)blockIdx.x) dbg_stamp(dbg, 48 + w * 40);
        if (q < S) {
          const float inv = 1.0f / __uint_as_float(o[HD]);
          uint4* dst = reinterpret_cast<uint4*>(p.out + (size_t)(u.seq_row0 + (uint32_t)q * p.inner) * p.ldo + u.h * HD);
#pragma unroll
          for (int g2 = 0; g2 < 3; ++g2)
            dst[g2] = make_uint4(pack_bf16(__uint_as_float(o[g2 * 8]) * inv, __uint_as_float(o[g2 * 8 + 1]) * inv),
                                 pack_bf16(__uint_as_float(o[g2 * 8 + 2]) * inv, __uint_as_float(o[g2 * 8 + 3]) * inv),
                                 pack_bf16(__uint_as_float(o[g2 * 8 + 4]) * inv, __uint_as_float(o[g2 * 8 + 5]) * inv),
                                 pack_bf16(__uint_as_float(o[g2 * 8 + 6]) * inv, __uint_as_float(o[g2 * 8 + 7]) * inv));
        }
      }
    }
  } else {
    // =============================== producers ===============================
    // Thread r of the 128 stages row r of every operand block: Q tiles of the unit first (into the buffer the
    // warpgroup is not reading), then the unit's K / V ring blocks (96 rows: the last producer warp only arrives).
    setmaxnreg_dec<REGS_PROD>();
    const int r = tid - PROD_WARP0 * 32;
    uint32_t gblk = 0;
    int nbatch = 0;                       // DBG: producer batches (Q tiles and K/V blocks) stamped so far
    uint32_t quse[NWG] = {0, 0, 0, 0};
    for (int unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
      const Unit u = decode_unit(p, unit, chunks_per_seq);
      const bf16* qkv_h = p.qkv + u.h * HD;
#pragma unroll
      for (int w = 0; w < NWG; ++w) {
        if (w < u.ntiles) {
          const uint32_t qbuf = quse[w] & 1;
          mbar_wait(&bars->q_empty[qbuf][w], ((quse[w] >> 1) & 1) ^ 1);
          if (DBG && r == 0 && nbatch < 48) dbg_stamp(dbg, 96 + 2 * nbatch);
          stage_q_row(p, qkv_h, smem + OFF_Q + (qbuf * NWG + w) * BLK_BYTES, r, u.q_begin + w * QT + r, u.seq_row0);
          fence_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars->q_full[qbuf][w]);
          if (DBG && r == 0 && nbatch < 48) dbg_stamp(dbg, 97 + 2 * nbatch);
          if (DBG) ++nbatch;
          ++quse[w];
        }
      }
      for (int kb = 0; kb < nkb; ++kb, ++gblk) {
        const uint32_t slot = gblk % NSLOT;
        mbar_wait(&bars->kv_empty[slot], ((gblk / NSLOT) & 1) ^ 1);
        if (DBG && r == 0 && nbatch < 48) dbg_stamp(dbg, 96 + 2 * nbatch);
        if (r < RB) {
          const int key = kb * RB + r;
          const uint32_t row = u.seq_row0 + (uint32_t)(key < S ? key : 0) * p.inner;
          const bf16* src = qkv_h + (size_t)row * p.ld;
          // V: cp.async 16-byte pieces straight into the MN-major operand layout (rows >= S zero-filled)
          const uint32_t vrow = sV + slot * RBLK_BYTES + r * 16;
#pragma unroll
          for (int c = 0; c < 3; ++c) cp_async16(vrow + c * RGRP_BYTES, src + 2 * C + c * 8, key < S ? 16u : 0u);
          cp_async_commit();
          // K: global -> registers -> RoPE -> K-major operand layout; key S = learned bias token (mha.py:265-268)
          uint4 g[3] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
          float valid = (key == S) ? 1.f : 0.f;
          if (key < S) {
            const uint4* ks = reinterpret_cast<const uint4*>(src + C);
#pragma unroll
            for (int j = 0; j < 3; ++j) g[j] = __ldg(ks + j);
            valid = __ldg(p.key_valid + row % (uint32_t)p.mask_mod);
          } else if (key == S) {
            uint32_t* gw = reinterpret_cast<uint32_t*>(g);
#pragma unroll
            for (int e = 0; e < HD / 2; ++e) gw[e] = pack_bf16(p.bias_k[u.h * HD + 2 * e], p.bias_k[u.h * HD + 2 * e + 1]);
          }
          if (p.cs && key <= S) rope_row(g, p.cs + key * (HD / 2), p.sn + key * (HD / 2));
          uint8_t* kr = smem + OFF_K + slot * RBLK_BYTES + r * 16;
#pragma unroll
          for (int j = 0; j < 3; ++j) *reinterpret_cast<uint4*>(kr + j * RGRP_BYTES) = g[j];
          // mask column (d = 24): -32768 for padded keys and for the rows behind the bias token -- far below any
          // score, yet small enough that fp32 rounding of (s - max) log2e stays ~2^-8 while only padded keys have been seen
          *reinterpret_cast<uint4*>(kr + 3 * RGRP_BYTES) = make_uint4(valid != 0.f ? 0u : 0xC700u, 0, 0, 0);
          cp_async_wait<0>();
          if (key == S) {                   // bias_v row, after this thread's zero-filling cp.async has landed
            uint8_t* vr = smem + OFF_V + slot * RBLK_BYTES + r * 16;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              uint4 v;
              uint32_t* vw = reinterpret_cast<uint32_t*>(&v);
#pragma unroll
              for (int e = 0; e < 4; ++e) vw[e] = pack_bf16(p.bias_v[u.h * HD + c * 8 + 2 * e], p.bias_v[u.h * HD + c * 8 + 2 * e + 1]);
              *reinterpret_cast<uint4*>(vr + c * RGRP_BYTES) = v;
            }
          }
          fence_async_smem();
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->kv_full[slot]);
        if (DBG && r == 0 && nbatch < 48) dbg_stamp(dbg, 97 + 2 * nbatch);
        if (DBG) ++nbatch;
      }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
  }
}

}  // namespace atc

inline bool attention_tc_supported(const AttnParams<bf16>& p, int head_dim) {
  return head_dim == atc::HD && p.S + 1 <= atc::MAX_KEYS && p.S >= 64 && (p.ld % 8) == 0 && (p.ldo % 8) == 0;
}

extern thread_local long long* g_dbg_buf;   // device buffer for dbg_stamp (mdgen_debug_set_buffer), normally null
extern thread_local int g_dbg_cta;

constexpr int ATTN_POLY_DEFAULT = 0;   // MDGEN_ATTN_POLY=3: every third exp2 on the FMA pipe -- measured SLOWER (216 -> 234 us on config 3:
                                       // the kernel is issue-bound, not MUFU-bound), kept for A/B runs

template <bool DBG, int POLY, int SB, int NBUF>
inline cudaError_t launch_attention_tc_variant(const AttnParams<bf16>& p, int grid, int chunks, int nunits, int var, cudaStream_t st) {
  static DevOnce once;
  const int dev = cur_device();
  if (once.need(dev)) {
    cudaError_t e = cudaFuncSetAttribute(atc::attention_tc_kernel<DBG, POLY, SB, NBUF>, cudaFuncAttributeMaxDynamicSharedMemorySize, atc::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    once.done(dev);
  }
  MD_VARIANT(V_ATTN_TC);
  return launch_pdl(atc::attention_tc_kernel<DBG, POLY, SB, NBUF>, grid, atc::THREADS, atc::SMEM_BYTES, st, p, chunks, nunits,
                    DBG ? g_dbg_buf : nullptr, g_dbg_cta, var);
}

inline int launch_attention_tc(const AttnParams<bf16>& p, int head_dim, cudaStream_t st) {
  MD_CHECK(attention_tc_supported(p, head_dim), "attention_tc: unsupported shape");
  // work units of up to NWG query tiles; unit id = (sequence, head, chunk) with the chunk fastest, so the CTAs
  // that share a head's K / V run side by side (second reader hits L2)
  const int qtiles = (p.S + atc::QT - 1) / atc::QT;
  const int chunks = (qtiles + atc::NWG - 1) / atc::NWG;
  const long nunits = (long)p.nb * p.H * chunks;
  const int grid = (int)(nunits < num_sms() ? nunits : num_sms());
  ProfScope ps(PC_ATTN_TC, st);
  const int var = g_sw.attn_var, poly = g_sw.attn_poly;
  const bool v8 = g_sw.attn_sb == 32;       // <32, 3>: three S buffers of 32 keys; <48, 2>: v7
  if (g_dbg_buf || var) {
    if (v8) MD_CUDA((launch_attention_tc_variant<true, 0, 32, 3>(p, grid, chunks, (int)nunits, var, st)));
    else MD_CUDA((launch_attention_tc_variant<true, 0, 48, 2>(p, grid, chunks, (int)nunits, var, st)));
  } else if (v8) {
    MD_CHECK(poly == 0, "MDGEN_ATTN_POLY is only built for the 48-key variant (MDGEN_ATTN_SB=48)");
    MD_CUDA((launch_attention_tc_variant<false, 0, 32, 3>(p, grid, chunks, (int)nunits, 0, st)));
  } else {
    switch (poly) {
      case 0: MD_CUDA((launch_attention_tc_variant<false, 0, 48, 2>(p, grid, chunks, (int)nunits, 0, st))); break;
      case 3: MD_CUDA((launch_attention_tc_variant<false, 3, 48, 2>(p, grid, chunks, (int)nunits, 0, st))); break;
      default: MD_CHECK(false, "MDGEN_ATTN_POLY must be 0 or 3");
    }
  }
  MD_LAUNCHED();
  MD_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace mdgen
