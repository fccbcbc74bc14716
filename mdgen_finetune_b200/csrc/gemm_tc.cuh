// gemm_tc.cuh -- bf16 x bf16 -> fp32 GEMM on the sm_100a 5th-gen tensor cores.
//   C[M,N] = A[M,K] . W[N,K]^T (+ fused epilogue), A and W both K-major bf16 in HBM.
// Persistent, warp-specialised: warp 0 = TMA producer (cp.async.bulk.tensor, SWIZZLE_128B),
// warp 1 = single-thread tcgen05.mma issuer (accumulators in TMEM, 2 accumulator stages so the
// epilogue of tile i overlaps the main loop of tile i+1), warps 2..9 = epilogue
// (tcgen05.ld -> bias / erf-GELU / adaLN gate + residual / Euler -> global).
// Covers q/k/v (fused N = 3C), out_proj, fc1, fc2 of every block (latent_model.py:385-392,466-491;
// mha.py:260-262,397) and the final projection (layers.py:74).
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace mdgen {
namespace tc {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // 64 bf16 = 128 B = one SWIZZLE_128B row
constexpr int UMMA_K = 16;
constexpr int NUM_EPI_WARPS = 8;          // two warps per TMEM lane quarter, each takes half of the columns
constexpr int NUM_THREADS = 64 + 32 * NUM_EPI_WARPS;
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;

__host__ __device__ constexpr int stage_bytes(int bn) { return A_STAGE_BYTES + bn * BLOCK_K * 2; }
__host__ __device__ constexpr int num_stages(int bn) { return bn >= 128 ? 5 : (bn >= 64 ? 6 : 8); }
// W-stationary variant (K <= WS_MAX_KB * 64): the CTA keeps its whole [BN, K] weight slice in shared memory and
// streams only A tiles.  At 128x128 tiles a streamed-operand GEMM needs (128 + 128) * 2 B per 2 * 128 * 128 FLOP =
// 128 B/clk/SM at tensor peak, ~3x what the L2 -> SM fabric delivers (~6300 B/clk chip-wide); W-stationary halves it.
constexpr int WS_MAX_KB = 6, WS_STAGES = 5;
__host__ __device__ constexpr int ws_smem_bytes(int bn) {
  return WS_MAX_KB * bn * BLOCK_K * 2 + WS_STAGES * A_STAGE_BYTES + 8 * (32 * 144 + 3 * 128 * 4) + 1024 + 256;
}
// per-epilogue-warp staging tile: 32 rows x 128 B (+16 B pad per row: conflict-free 128-bit rows both ways)
constexpr int STG_PITCH = 144, STG_BYTES = 32 * STG_PITCH;
// per-epilogue-warp column vectors (bias, gate of the two batch elements a 32-row slab can touch) for up to 128 columns.
// With the shared-memory carve-out at its maximum the L1 keeps almost nothing: reading bias / gate straight from
// global inside the epilogue loops cost an L2 round trip (~700 clk) per use and dominated the tile time.
constexpr int CV_FLOATS = 128, CV_BYTES_PER_WARP = 3 * CV_FLOATS * 4;
constexpr int EPI_SMEM_BYTES = 8 * (STG_BYTES + CV_BYTES_PER_WARP);
__host__ __device__ constexpr int tmem_cols(int bn) { return 2 * bn <= 32 ? 32 : 2 * bn <= 64 ? 64 : 2 * bn <= 128 ? 128 : 2 * bn <= 256 ? 256 : 512; }
__host__ __device__ constexpr int smem_bytes(int bn) { return num_stages(bn) * stage_bytes(bn) + EPI_SMEM_BYTES + 1024 + 256; }

static_assert(smem_bytes(128) <= 227 * 1024 && smem_bytes(64) <= 227 * 1024 && smem_bytes(32) <= 227 * 1024 &&
                  ws_smem_bytes(128) <= 227 * 1024, "shared memory budget");

// ---- PTX wrappers ----------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "DONE:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
// one lane of a converged warp.  The MMA issuer runs its loop warp-uniformly and issues under elect_one(): descriptors stay in
// uniform registers and one tcgen05.mma costs ~3 instructions.  Issued from a `lane == 0` branch instead, the compiler
// wraps every tcgen05.mma in a uniformising loop (R2UR / ELECT / BRA.U.ANY, ~10 dependent instructions), and with two
// GELU warps competing for the same scheduler the ISSUE of the MMAs became slower than their execution (tools/mlp_trace.py; first seen in gemm_mlp.cuh).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "elect.sync _|P, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, P;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem desc] . B[smem desc]^T, kind::f16 (bf16 inputs, fp32 accumulate)
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// bulk tensor store shared -> global (rows / columns outside the tensor are clipped by the map), bulk-group completion
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map), "r"(smem_u32(src)), "r"(c0),
               "r"(c1)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor):
// start>>4 [0,14) | LBO>>4 [16,30) = 1 | SBO>>4 [32,46) = 1024>>4 (8 rows x 128 B) | version=1 [46,48)
// | layout_type = SWIZZLE_128B (2) [61,64).
__device__ __forceinline__ uint64_t make_smem_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// cute::UMMA::InstrDescriptor for bf16 x bf16 -> f32, both operands K-major
__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// erf-GELU (layers.py:78-85) = x * Phi(x) with one MUFU op: Phi(-|x|) = 2^-q(|x|), q a degree-5 polynomial fitted
// (weighted minimax of the GELU error, tools/fit_gelu.py) to -log2(erfc(|x|/sqrt 2)/2) on [0, 7]; beyond 7 the
// polynomial keeps growing so 2^-q underflows to 0 like the exact tail.  Max |error| 6.4e-7 in fp32 over [-30, 30]
// (the exact erff costs ~3x the issue slots and the rcp + ex2 formulation two MUFU ops per element, which made the
// fc1 epilogue slower than its tile's MMAs).
__device__ __forceinline__ float gelu_erf_fast(float x) {
  const float a = fabsf(x);
  float q = fmaf(a, 0.0004733167006634176f, -0.007084609474986792f);
  q = fmaf(a, q, 0.05182751268148422f);
  q = fmaf(a, q, 0.45999228954315186f);
  q = fmaf(a, q, 1.1507879495620728f);
  q = fmaf(a, q, 1.000037670135498f);
  float h;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(h) : "f"(-q));
  return fmaf(-a, h, fmaxf(x, 0.f));  // x > 0: x - x Phi(-x);  x < 0: x Phi(x) -- one FMNMX + one FFMA, no select
}

// ---- epilogue on 32 consecutive accumulator columns of one row --------------------------
template <int EPI, typename OT>
__device__ __forceinline__ void epilogue_chunk(const EpiParams& p, int m, int n0, int N, const uint32_t* r) {
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) + ((p.bias && n0 + j < N) ? __ldg(p.bias + n0 + j) : 0.f);
  const bool full = (n0 + 32 <= N);
  if constexpr (EPI == EPI_STORE || EPI == EPI_GELU) {
    if constexpr (EPI == EPI_GELU) {
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = gelu_erf_fast(v[j]);
    }
    OT* o = reinterpret_cast<OT*>(p.out) + (size_t)m * p.ldo + n0;
    if (full && (p.ldo % 8 == 0)) {
      if constexpr (sizeof(OT) == 2) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          __nv_bfloat162 a = __floats2bfloat162_rn(v[j], v[j + 1]), b = __floats2bfloat162_rn(v[j + 2], v[j + 3]);
          __nv_bfloat162 c = __floats2bfloat162_rn(v[j + 4], v[j + 5]), d = __floats2bfloat162_rn(v[j + 6], v[j + 7]);
          uint4 pk = make_uint4(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b),
                                *reinterpret_cast<uint32_t*>(&c), *reinterpret_cast<uint32_t*>(&d));
          *reinterpret_cast<uint4*>(o + j) = pk;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (n0 + j < N) o[j] = from_f<OT>(v[j]);
    }
  } else if constexpr (EPI == EPI_GATE_RESID || EPI == EPI_RESID) {
    float* x = p.resid + (size_t)m * p.ldr + n0;
    const float* g = nullptr;
    if constexpr (EPI == EPI_GATE_RESID) g = p.gate + (size_t)((m / p.rows_per_batch) % p.nbatch) * p.gate_stride + n0;
    const bool vec = full && (p.ldr % 4 == 0) && (((uintptr_t)p.resid & 15) == 0) &&
                     (EPI != EPI_GATE_RESID || ((p.gate_stride % 4 == 0) && (((uintptr_t)p.gate & 15) == 0)));
    if (vec) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) {
        float4 xv = *reinterpret_cast<float4*>(x + j);
        float4 gv = make_float4(1.f, 1.f, 1.f, 1.f);
        if constexpr (EPI == EPI_GATE_RESID) gv = __ldg(reinterpret_cast<const float4*>(g + j));
        xv.x = fmaf(gv.x, v[j], xv.x);
        xv.y = fmaf(gv.y, v[j + 1], xv.y);
        xv.z = fmaf(gv.z, v[j + 2], xv.z);
        xv.w = fmaf(gv.w, v[j + 3], xv.w);
        *reinterpret_cast<float4*>(x + j) = xv;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (n0 + j < N) x[j] += (g ? g[j] : 1.f) * v[j];
    }
  } else {  // EPI_EULER: velocity out (fp32) and optional in-place state update
    float* o = reinterpret_cast<float*>(p.out) + (size_t)m * p.ldo + n0;
    float* s = p.resid ? p.resid + (size_t)m * p.ldr + n0 : nullptr;
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (n0 + j < N) {
        o[j] = v[j];
        if (s) s[j] = fmaf(p.dt, v[j], s[j]);
      }
  }
}

// ---- staged epilogues: registers (thread = row) -> per-warp shared tile -> coalesced 128-bit global accesses ----
// The row-per-thread layout that tcgen05.ld delivers makes every direct global access touch 32 different
// cache lines per instruction; transposing through shared memory turns them into full 128-byte lines.

// dst[0..128) = src[n0 .. n0+128) (columns >= N and a null src read as `fill`); one float4 per lane, warp-private
__device__ __forceinline__ void stage_colvec(float* dst, const float* src, int n0, int N, int lane, float fill) {
  float4 v = make_float4(fill, fill, fill, fill);
  const int n = n0 + lane * 4;
  if (src) {
    if (n + 4 <= N) {
      v = __ldg(reinterpret_cast<const float4*>(src + n));
    } else {
      if (n < N) v.x = src[n];
      if (n + 1 < N) v.y = src[n + 1];
      if (n + 2 < N) v.z = src[n + 2];
    }
  }
  *reinterpret_cast<float4*>(dst + lane * 4) = v;
}

// bf16 output of W32 = 32 * PER consecutive columns of 32 rows (EPI_STORE / EPI_GELU); sbias = staged bias of column n0
template <int EPI, int PER>
__device__ __forceinline__ void epilogue_staged_bf16(const EpiParams& p, uint8_t* stg, const float* sbias, int lane, int mrow0,
                                                     int n0, int M, int N, uint32_t (&r)[PER][32]) {
  constexpr int PIECES = PER * 4;               // 16-byte pieces per row
  constexpr int ROWS_PER = 32 / PIECES;         // rows covered by one warp-wide access
  constexpr int ITERS = 32 / ROWS_PER;
  // Loads are batched ahead of their uses by hand (bias quads, then the staged rows): interleaved load/use pairs on one
  // recycled register quad left every shared-memory latency exposed (short-scoreboard stalls on each FADD / STG).
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    float4 b4[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) b4[k] = *reinterpret_cast<const float4*>(sbias + i * 32 + k * 4);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float v0 = __uint_as_float(r[i][4 * k]) + b4[k].x, v1 = __uint_as_float(r[i][4 * k + 1]) + b4[k].y;
      float v2 = __uint_as_float(r[i][4 * k + 2]) + b4[k].z, v3 = __uint_as_float(r[i][4 * k + 3]) + b4[k].w;
      if constexpr (EPI == EPI_GELU) {
        v0 = gelu_erf_fast(v0);
        v1 = gelu_erf_fast(v1);
        v2 = gelu_erf_fast(v2);
        v3 = gelu_erf_fast(v3);
      }
      const __nv_bfloat162 lo = __floats2bfloat162_rn(v0, v1), hi = __floats2bfloat162_rn(v2, v3);
      r[i][2 * k] = *reinterpret_cast<const uint32_t*>(&lo);       // packed in place: r[i][0..15] = 32 bf16
      r[i][2 * k + 1] = *reinterpret_cast<const uint32_t*>(&hi);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k)
      *reinterpret_cast<uint4*>(stg + lane * STG_PITCH + (i * 4 + k) * 16) =
          make_uint4(r[i][4 * k], r[i][4 * k + 1], r[i][4 * k + 2], r[i][4 * k + 3]);
  }
  __syncwarp();
  bf16* out = reinterpret_cast<bf16*>(p.out);
  const int piece = lane % PIECES, rsub = lane / PIECES;
  const int n = n0 + piece * 8;
  uint4 v[ITERS];
#pragma unroll
  for (int it = 0; it < ITERS; ++it) v[it] = *reinterpret_cast<const uint4*>(stg + (it * ROWS_PER + rsub) * STG_PITCH + piece * 16);
  bf16* o = out + (size_t)(mrow0 + rsub) * p.ldo + n;
  const size_t step = (size_t)ROWS_PER * p.ldo;
  if (mrow0 + 32 <= M && n0 + PER * 32 <= N) {       // interior slab: unpredicated 128-bit stores
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
      *reinterpret_cast<uint4*>(o) = v[it];
      o += step;
    }
  } else {
#pragma unroll
    for (int it = 0; it < ITERS; ++it) {
      if (mrow0 + it * ROWS_PER + rsub < M) {
        if (n + 8 <= N) {
          *reinterpret_cast<uint4*>(o) = v[it];
        } else {
          const bf16* e = reinterpret_cast<const bf16*>(&v[it]);
          for (int k = 0; k < 8; ++k)
            if (n + k < N) o[k] = e[k];
        }
      }
      o += step;
    }
  }
  __syncwarp();
}

// bf16 output of 64 columns x 32 rows through the bulk-copy engine: each thread (= row) packs its 128 bytes into a
// SWIZZLE_128B tile (16-byte chunk j of row r at chunk j ^ (r & 7): conflict-free), one lane issues the tensor store.
// No LDS / STG, no L1: the LSU path moved only ~25 B/clk/SM in these epilogues (gemm_ln.cuh measurements).
constexpr int TMA_TILE_BYTES = 4096;
template <int EPI>
__device__ __forceinline__ void epilogue_tma_bf16(const CUtensorMap* mapC, uint8_t* tile, const float* sbias, int lane, int mrow0,
                                                  int n0, uint32_t (&r)[2][32]) {
  if (lane == 0) bulk_wait_read<0>();              // the previous store from this tile has been read out
  __syncwarp();
  uint8_t* row = tile + lane * 128;
  const uint32_t swz = (uint32_t)(lane & 7);
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    float4 b4[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) b4[k] = *reinterpret_cast<const float4*>(sbias + i * 32 + k * 4);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float v0 = __uint_as_float(r[i][4 * k]) + b4[k].x, v1 = __uint_as_float(r[i][4 * k + 1]) + b4[k].y;
      float v2 = __uint_as_float(r[i][4 * k + 2]) + b4[k].z, v3 = __uint_as_float(r[i][4 * k + 3]) + b4[k].w;
      if constexpr (EPI == EPI_GELU) {
        v0 = gelu_erf_fast(v0);
        v1 = gelu_erf_fast(v1);
        v2 = gelu_erf_fast(v2);
        v3 = gelu_erf_fast(v3);
      }
      const __nv_bfloat162 lo = __floats2bfloat162_rn(v0, v1), hi = __floats2bfloat162_rn(v2, v3);
      r[i][2 * k] = *reinterpret_cast<const uint32_t*>(&lo);
      r[i][2 * k + 1] = *reinterpret_cast<const uint32_t*>(&hi);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k)
      *reinterpret_cast<uint4*>(row + (((uint32_t)(i * 4 + k) ^ swz) << 4)) =
          make_uint4(r[i][4 * k], r[i][4 * k + 1], r[i][4 * k + 2], r[i][4 * k + 3]);
  }
  fence_async_smem();
  __syncwarp();
  if (lane == 0) {
    tma_store_2d(mapC, tile, n0, mrow0);
    bulk_commit();
  }
}

// fp32 residual read-modify-write of 32 columns x 32 rows: x += gate * (acc + bias)   (EPI_GATE_RESID / EPI_RESID).
// Everything the inner loops need is precomputed per tile into registers (ResidLane): kernel parameters live in
// the constant bank and re-reading them inside the loops (LDC, long-scoreboard latency) dominated this epilogue.
struct ResidLane {
  float* x;            // this lane's first residual element: row rsub, columns n .. n+3 of the chunk
  size_t row_step;     // 4 rows, in floats
  int split;           // first row (0..32) that belongs to the next batch element
  int rows;            // valid rows of the 32 (M tail)
  int n;               // first of the lane's 4 columns in chunk 0 of the warp
};
template <int EPI>
__device__ __forceinline__ ResidLane resid_lane(const EpiParams& p, int lane, int mrow0, int n0, int M, int N) {
  ResidLane L;
  const int c4 = lane & 7, rsub = lane >> 3, n = n0 + c4 * 4;
  L.x = p.resid + (size_t)(mrow0 + rsub) * p.ldr + n;
  L.row_step = (size_t)4 * p.ldr;
  L.rows = M - mrow0;
  L.n = n;
  L.split = 32;
  if constexpr (EPI == EPI_GATE_RESID) {
    const int q = mrow0 / p.rows_per_batch;
    L.split = (q + 1) * p.rows_per_batch - mrow0;        // >= 1; a 32-row slab crosses at most one boundary
  }
  return L;
}
// stage bias and the two gate rows of the slab's 128 columns starting at n0 into the warp's column-vector buffer
template <int EPI>
__device__ __forceinline__ void stage_resid_colvecs(const EpiParams& p, float* cv, int lane, int mrow0, int n0, int N) {
  stage_colvec(cv, p.bias, n0, N, lane, 0.f);
  if constexpr (EPI == EPI_GATE_RESID) {
    const int q = mrow0 / p.rows_per_batch;
    stage_colvec(cv + CV_FLOATS, p.gate + (size_t)(q % p.nbatch) * p.gate_stride, n0, N, lane, 1.f);
    stage_colvec(cv + 2 * CV_FLOATS, p.gate + (size_t)((q + 1) % p.nbatch) * p.gate_stride, n0, N, lane, 1.f);
  }
}
// coff = column offset of the chunk inside the warp's span (0, 32, ..)
__device__ __forceinline__ void residual_prefetch(const ResidLane& L, int lane, int coff, int N, float4 (&xv)[8]) {
  const int rsub = lane >> 3;
  const float* x = L.x + coff;
  const bool vec = L.n + coff + 4 <= N;
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    xv[it] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (it * 4 + rsub < L.rows && vec) xv[it] = *reinterpret_cast<const float4*>(x);
    x += L.row_step;
  }
}
template <int EPI>
__device__ __forceinline__ void epilogue_staged_resid(const ResidLane& L, const float* cv, uint8_t* stg, int lane, int coff,
                                                      int N, const uint32_t (&r)[32], const float4 (&xv)[8]) {
#pragma unroll
  for (int j = 0; j < 32; j += 4)
    *reinterpret_cast<uint4*>(stg + lane * STG_PITCH + j * 4) = make_uint4(r[j], r[j + 1], r[j + 2], r[j + 3]);
  __syncwarp();
  const int c4 = lane & 7, rsub = lane >> 3;
  const int n = L.n + coff;
  const bool vec = n + 4 <= N;
  // in the transposed phase a lane owns 4 fixed columns: its bias and gate values are loop invariants (from shared memory)
  const float4 bz = *reinterpret_cast<const float4*>(cv + coff + c4 * 4);
  float4 g0 = make_float4(1.f, 1.f, 1.f, 1.f), g1 = g0;
  if constexpr (EPI == EPI_GATE_RESID) {
    g0 = *reinterpret_cast<const float4*>(cv + CV_FLOATS + coff + c4 * 4);
    g1 = *reinterpret_cast<const float4*>(cv + 2 * CV_FLOATS + coff + c4 * 4);
  }
  float* x = L.x + coff;
  float4 v[8];
#pragma unroll
  for (int it = 0; it < 8; ++it) v[it] = *reinterpret_cast<const float4*>(stg + (it * 4 + rsub) * STG_PITCH + c4 * 16);
  if (L.rows >= 32 && vec) {            // interior slab: unpredicated read-modify-write
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const float4 g = it * 4 + rsub < L.split ? g0 : g1;
      float4 o = xv[it];
      o.x = fmaf(g.x, v[it].x + bz.x, o.x);
      o.y = fmaf(g.y, v[it].y + bz.y, o.y);
      o.z = fmaf(g.z, v[it].z + bz.z, o.z);
      o.w = fmaf(g.w, v[it].w + bz.w, o.w);
      *reinterpret_cast<float4*>(x) = o;
      x += L.row_step;
    }
  } else {
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int row = it * 4 + rsub;
      if (row < L.rows) {
        const float4 g = row < L.split ? g0 : g1;
        const float vv[4] = {v[it].x, v[it].y, v[it].z, v[it].w}, gg[4] = {g.x, g.y, g.z, g.w}, bb[4] = {bz.x, bz.y, bz.z, bz.w};
        if (vec) {
          float4 o = xv[it];
          o.x = fmaf(gg[0], vv[0] + bb[0], o.x);
          o.y = fmaf(gg[1], vv[1] + bb[1], o.y);
          o.z = fmaf(gg[2], vv[2] + bb[2], o.z);
          o.w = fmaf(gg[3], vv[3] + bb[3], o.w);
          *reinterpret_cast<float4*>(x) = o;
        } else {
          for (int k = 0; k < 4; ++k)
            if (n + k < N) x[k] += gg[k] * (vv[k] + bb[k]);
        }
      }
      x += L.row_step;
    }
  }
  __syncwarp();
}

// The bf16 epilogue on ONE 32-column chunk, stored through a 32 x 32 SWIZZLE_64B tile (16-byte chunk j of row r at chunk
// j ^ ((r >> 1) & 3)): the unit of work of the pair kernel's 16-epilogue-warp variant (gemm_tc2.cuh).  (16 epilogue warps
// on THIS kernel were measured too -- GELU projection 53.5 vs 52.8 ms per rollout in the GEMM family, no gain: its
// 128 x 128 mainloop is shared-memory-bandwidth bound -- and removed.)
constexpr int TMA_TILE32_BYTES = 2048, EPI16_WARPS = 16;
// PENDING = bulk stores of this warp that may still be reading OTHER tiles (1 when the warp alternates between two tiles)
template <int EPI, int PENDING = 0>
__device__ __forceinline__ void epilogue_tma_bf16_32(const CUtensorMap* mapC, uint8_t* tile, const float* sbias, int lane, int mrow0,
                                                     int n0, uint32_t (&r)[32]) {
  if (lane == 0) bulk_wait_read<PENDING>();        // the previous store from this tile has been read out
  __syncwarp();
  uint8_t* row = tile + lane * 64;
  const uint32_t swz = (uint32_t)((lane >> 1) & 3);
  float4 b4[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) b4[k] = *reinterpret_cast<const float4*>(sbias + k * 4);
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    float v0 = __uint_as_float(r[4 * k]) + b4[k].x, v1 = __uint_as_float(r[4 * k + 1]) + b4[k].y;
    float v2 = __uint_as_float(r[4 * k + 2]) + b4[k].z, v3 = __uint_as_float(r[4 * k + 3]) + b4[k].w;
    if constexpr (EPI == EPI_GELU) {
      v0 = gelu_erf_fast(v0);
      v1 = gelu_erf_fast(v1);
      v2 = gelu_erf_fast(v2);
      v3 = gelu_erf_fast(v3);
    }
    const __nv_bfloat162 lo = __floats2bfloat162_rn(v0, v1), hi = __floats2bfloat162_rn(v2, v3);
    r[2 * k] = *reinterpret_cast<const uint32_t*>(&lo);
    r[2 * k + 1] = *reinterpret_cast<const uint32_t*>(&hi);
  }
#pragma unroll
  for (int k = 0; k < 4; ++k)
    *reinterpret_cast<uint4*>(row + (((uint32_t)k ^ swz) << 4)) = make_uint4(r[4 * k], r[4 * k + 1], r[4 * k + 2], r[4 * k + 3]);
  fence_async_smem();
  __syncwarp();
  if (lane == 0) {
    tma_store_2d(mapC, tile, n0, mrow0);
    bulk_commit();
  }
}

// ---- the kernel ----------------------------------------------------------------------------
template <int BN, int EPI, typename OT, int WS, int STG>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapW,
               const __grid_constant__ CUtensorMap tmapC, int M, int N, int K, EpiParams ep) {
  constexpr int STAGES = WS ? WS_STAGES : num_stages(BN);
  constexpr int STAGE_BYTES = WS ? A_STAGE_BYTES : stage_bytes(BN);
  constexpr int W_KB_BYTES = BN * BLOCK_K * 2;                 // one 64-wide K block of the weight slice
  constexpr int W_BYTES = WS ? WS_MAX_KB * W_KB_BYTES : 0;     // resident weight slice (WS only), in front of the ring
  constexpr uint32_t TMEM_COLS = tmem_cols(BN);
  // No integer round trip on this pointer: the compiler must keep seeing a shared-space address, otherwise every
  // staging access becomes a generic LD.E / ST.E (long-scoreboard latency instead of LDS / STS).
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) __trap();     // dynamic shared memory starts at the window base (no static shared)
  uint8_t* ring = smem + W_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ring + STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* w_bar = tmem_empty + 2;
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(w_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_m = (M + BLOCK_M - 1) / BLOCK_M, num_n = (N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int num_kb = (K + BLOCK_K - 1) / BLOCK_K;
  // tile schedule: streamed-W = round robin over all tiles; W-stationary = fixed n tile, strided m tiles
  const int ws_nb = blockIdx.x % num_n, ws_ci = blockIdx.x / num_n;
  const int ws_cn = ((int)gridDim.x - ws_nb + num_n - 1) / num_n;
  auto get_tile = [&](int i, int& mb, int& nb) -> bool {
    if constexpr (WS) {
      mb = ws_ci + i * ws_cn;
      nb = ws_nb;
      return mb < num_m;
    } else {
      const int tile = blockIdx.x + i * (int)gridDim.x;
      mb = tile / num_n;
      nb = tile % num_n;
      return tile < num_tiles;
    }
  };

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapW) : "memory");
    if constexpr (STG == 2) asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapC) : "memory");
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], NUM_EPI_WARPS);
    }
    mbar_init(w_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_slot)), "r"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;
  pdl_trigger();       // common.cuh: after the TMEM allocation
  pdl_wait();          // the previous kernel's outputs (A, residual) are complete from here on

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int mb, nb;
      if constexpr (WS) {
        if (get_tile(0, mb, nb)) {
          mbar_expect_tx(w_bar, num_kb * W_KB_BYTES);
          for (int kb = 0; kb < num_kb; ++kb) tma_load_2d(&tmapW, w_bar, smem + kb * W_KB_BYTES, kb * BLOCK_K, nb * BN);
        }
      }
      for (int i = 0; get_tile(i, mb, nb); ++i) {
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = ring + stage * STAGE_BYTES;
          mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
          tma_load_2d(&tmapA, &full_bar[stage], sa, kb * BLOCK_K, mb * BLOCK_M);
          if constexpr (!WS) tma_load_2d(&tmapW, &full_bar[stage], sa + A_STAGE_BYTES, kb * BLOCK_K, nb * BN);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the warp walks the loop, one elected lane issues (elect_one) =====
    {
      constexpr uint32_t idesc = make_idesc_bf16(BLOCK_M, BN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      int mb, nb;
      if constexpr (WS) {
        if (get_tile(0, mb, nb)) mbar_wait(w_bar, 0);          // the resident weight slice has landed
      }
      for (int i = 0; get_tile(i, mb, nb); ++i) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tcgen05_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tcgen05_fence_after();
          const uint32_t sa = smem_u32(ring + stage * STAGE_BYTES);
          const uint64_t adesc = make_smem_desc_sw128(sa);
          const uint64_t bdesc = make_smem_desc_sw128(WS ? smem_u32(smem + kb * W_KB_BYTES) : sa + A_STAGE_BYTES);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
              // advance 16 bf16 = 32 B inside the 128 B swizzle row: +2 in the (addr >> 4) field
              umma_bf16(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) != 0 ? 1u : 0u);
            }
            tcgen05_commit(&empty_bar[stage]);  // frees the smem slot when these MMAs retire
            if (kb == num_kb - 1) tcgen05_commit(&tmem_full[acc]);  // accumulator ready for the epilogue
          }
          __syncwarp();
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else {
    // ===== epilogue warps 2..9: TMEM lane quarter = warp % 4, column half = (warp - 2) / 4 =====
    const int quarter = warp & 3, half = (warp - 2) >> 2;
    constexpr int CHUNKS = BN / 32;                      // 32-column chunks per tile
    constexpr int PER = (CHUNKS + 1) / 2;                // chunks per warp
    // STG = coalesced shared-memory-staged epilogue (host guarantees the alignment it needs); else row-per-thread
    constexpr bool kResid = STG && (EPI == EPI_GATE_RESID || EPI == EPI_RESID);
    constexpr bool kBf16 = STG && (EPI == EPI_STORE || EPI == EPI_GELU) && sizeof(OT) == 2;
    static_assert(!STG || kResid || kBf16, "no staged epilogue for this combination");
    // STG == 2: bf16 output leaves through TMA stores from 1024-byte aligned SWIZZLE_128B tiles (64 columns per warp)
    constexpr bool kTma = STG == 2;
    static_assert(!kTma || (kBf16 && PER == 2), "the TMA-store epilogue is the 128-column bf16 one");
    uint8_t* epi = ring + STAGES * STAGE_BYTES + (kTma ? 1024 : 256);
    uint8_t* stg = epi + (warp - 2) * (kTma ? TMA_TILE_BYTES : STG_BYTES);
    float* cv = reinterpret_cast<float*>(epi + NUM_EPI_WARPS * (kTma ? TMA_TILE_BYTES : STG_BYTES) + (warp - 2) * CV_BYTES_PER_WARP);
    const bool active = half * PER < CHUNKS;             // BN = 32: only the first column half has work
    int acc = 0;
    uint32_t acc_phase = 0;
    int mb, nb;
    for (int i = 0; get_tile(i, mb, nb); ++i) {
      const int mrow0 = mb * BLOCK_M + quarter * 32;
      const int ncol0 = nb * BN + half * PER * 32;
      float4 xv[kResid ? PER : 1][8];
      ResidLane rl;
      if constexpr (kResid) {
        if (active) {
          rl = resid_lane<EPI>(ep, lane, mrow0, ncol0, M, N);
          stage_resid_colvecs<EPI>(ep, cv, lane, mrow0, ncol0, N);
          __syncwarp();
#pragma unroll
          for (int i = 0; i < PER; ++i) residual_prefetch(rl, lane, i * 32, N, xv[i]);
        }
      }
      if constexpr (kBf16) {
        if (active) stage_colvec(cv, ep.bias, ncol0, N, lane, 0.f);
        __syncwarp();
      }
      mbar_wait(&tmem_full[acc], acc_phase);
      tcgen05_fence_after();
      const int m = mrow0 + lane;
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * BN);
      if constexpr (kResid) {
        // one 32-column chunk at a time (keeps a single chunk of accumulators live next to the prefetched residual)
        {
#pragma unroll
          for (int i = 0; i < PER; ++i) {
            uint32_t rr[32];
            const int c = half * PER + i;
            if (c < CHUNKS) tmem_ld_32x32b_x32(taddr + c * 32, rr);
            tmem_ld_wait();
            if (i == PER - 1) {                     // accumulator fully read: hand the TMEM stage back
              tcgen05_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(&tmem_empty[acc]);
            }
            if (active && ncol0 + i * 32 < N) epilogue_staged_resid<EPI>(rl, cv, stg, lane, i * 32, N, rr, xv[i]);
          }
        }
      } else {
        uint32_t r[PER][32];
#pragma unroll
        for (int i = 0; i < PER; ++i) {
          const int c = half * PER + i;
          if (c < CHUNKS) tmem_ld_32x32b_x32(taddr + c * 32, r[i]);
        }
        tmem_ld_wait();
        // accumulator is in registers: hand the TMEM stage back before the math / stores
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        if (active && ncol0 < N) {
          if constexpr (kTma) {
            epilogue_tma_bf16<EPI>(&tmapC, stg, cv, lane, mrow0, ncol0, r);
          } else if constexpr (kBf16) {
            epilogue_staged_bf16<EPI, PER>(ep, stg, cv, lane, mrow0, ncol0, M, N, r);
          } else {
#pragma unroll
            for (int i2 = 0; i2 < PER; ++i2) {
              const int n0 = ncol0 + i2 * 32;
              if (half * PER + i2 < CHUNKS && m < M && n0 < N) epilogue_chunk<EPI, OT>(ep, m, n0, N, r[i2]);
            }
          }
        }
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if constexpr (kTma) {
      if (lane == 0) bulk_wait_read<0>();       // shared memory must outlive the last bulk store's read
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

}  // namespace tc
}  // namespace mdgen
