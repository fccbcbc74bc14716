// gemm_qkv_attn.cuh -- residue-axis attention at L = 4 fused into the q/k/v projection (latent_model.py:466-472,
// mha.py:260-286, 356-397): ONE kernel computes  att = MHA_core(h W_qkv^T + b)  for sequences of 4 tokens + the learned
// bias-KV token, and only att [N, C] bf16 ever reaches HBM.  Replaces, on the 4AA configs, the qkv GEMM's [N, 3C] store,
// attention_seq4_bulk_kernel's load of it (2 x 74 MB per layer on config 3) and one launch.
//
// Mainloop = the CTA-pair kernel of gemm_tc2.cuh (cluster of 2, tcgen05 cta_group::2, M = 256 rows per tile, TMA
// SWIZZLE_128B operand ring).  What differs:
//   * the packed weight is HEAD-MAJOR: tile t holds heads 3t .. 3t+2 as [q_h | k_h | v_h] = 72 rows each (q rows pre-scaled
//     by d^-0.5 like every packed qkv weight), padded to N = 224 (an MMA N for M = 256); 6 tiles cover the 16 heads;
//   * 12 epilogue warps per CTA: TMEM lane quarter = warp % 4, head of the tile = (warp - 2) / 4.  A thread owns one token
//     row and one head: tcgen05.ld brings its q, k, v (72 fp32); the 4 rows of a sequence are 4 consecutive lanes, which
//     exchange their (rotated) k and v rows in fp32 through a per-warp shared-memory tile read back with 128-bit loads
//     (a first version exchanged packed bf16x2 through 96 shuffles and spent twice the instructions on pack / unpack);
//     RoPE from a shared-memory table (position = row % 4; the bias token sits at position 4), key padding, the 4 x 5 fp32
//     softmax and P.V run in registers; 48 bytes per (row, head) are stored.
#pragma once
#include "gemm_tc2.cuh"

namespace mdgen {
namespace tqa {

using namespace tc;
using tc2::cluster_ctarank;
using tc2::cluster_sync_all;
using tc2::map_to_rank;
using tc2::mbar_arrive_cluster;
using tc2::tcgen05_commit_pair;
using tc2::tma_load_2d_pair;
using tc2::umma_bf16_pair;

constexpr int HD = 24, HEADS_PER_TILE = 3, HEAD_COLS = 3 * HD;       // 72 accumulator columns per head
constexpr int TILE_M = 256, TILE_N = 224, HALF_N = TILE_N / 2;         // 216 live columns + 8 zero rows of padding
constexpr int ACC_STRIDE = 256;                                        // TMEM columns per accumulator stage
constexpr int EPI_WARPS = 4 * HEADS_PER_TILE, THREADS = 64 + 32 * EPI_WARPS;
constexpr int STAGES = 4;                                              // (the kernel is epilogue-bound: the ring gives room to the exchange area)
constexpr int STAGE_BYTES = A_STAGE_BYTES + HALF_N * BLOCK_K * 2;      // 16 KB + 14 KB
static_assert(STAGE_BYTES % 1024 == 0, "SWIZZLE_128B operand tiles start on 1024-byte boundaries");
constexpr int SEQ = 4;                                                 // tokens per sequence (L)
// shared memory behind the operand ring: barriers | tmem slot | rope cos / sin [5][12] | bias_k rotated to position 4 [H][24]
// | bias_v [H][24] | the packed projection bias [6][224] | k / v exchange tiles
constexpr int OFF_BAR = STAGES * STAGE_BYTES;
constexpr int OFF_ROPE = OFF_BAR + 256;
constexpr int OFF_BK = OFF_ROPE + 2 * (SEQ + 1) * (HD / 2) * 4;
constexpr int OFF_BV = OFF_BK + 16 * HD * 4;
constexpr int MAX_N_TILES = 6;                                         // 16 heads
constexpr int OFF_BIAS = OFF_BV + 16 * HD * 4;                         // the whole packed projection bias [6 tiles][224], staged once
// k / v exchange: per epilogue warp 32 rows x [k 24 | v 24] fp32 = 12 chunks of 16 B per row, pitch 192 B.  The 8 sequences
// of a warp read the same chunk of rows 4 apart -- 4 x 192 B = a multiple of 16 banks whatever the pitch -- so sequence g
// keeps chunk c at position (c + g) mod 12: the eight 16-byte reads of one load instruction fall on distinct banks
// (the kernel is shared-memory-bandwidth bound: operand ring + 600 KB of exchange traffic per tile).
constexpr int XCH_CHUNKS = 12, XCH_PITCH = XCH_CHUNKS * 16, XCH_WARP_BYTES = 32 * XCH_PITCH;
constexpr int OFF_XCH = (OFF_BIAS + MAX_N_TILES * TILE_N * 4 + 15) / 16 * 16;
constexpr int SMEM_BYTES = OFF_XCH + EPI_WARPS * XCH_WARP_BYTES + 64;
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");

__host__ __device__ constexpr int num_n_tiles(int heads) { return (heads + HEADS_PER_TILE - 1) / HEADS_PER_TILE; }

struct Params {
  bf16* att;                 // [M, ldo] output
  int ldo;
  const float* bias;         // [num_n_tiles * 224] projection bias in the packed (head-major, padded) order
  const float* bias_k;       // [H * 24] learned bias key / value (mha.py:111-116)
  const float* bias_v;
  const float* key_valid;    // [mask_mod]: 1 = attendable token, indexed by row % mask_mod
  int mask_mod;
  const float* cs;           // RoPE tables [5][12] (positions 0..4) or null
  const float* sn;
  int H;
  int dbg_skip;              // timing experiment (MDGEN_ATTN_VAR bit 32): 1 = epilogue releases TMEM and stores nothing
};

__device__ __forceinline__ void tmem_ld_32x32b_x8(uint32_t taddr, uint32_t* r) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
// v[0..23] = acc + bias (bias through 128-bit shared-memory loads), then rotate_half pairing (i, i + 12) with the cos / sin
// rows of the token's position held in registers
__device__ __forceinline__ void bias_rope24(const uint32_t* acc, const float4* bias4, const float4* c4, const float4* s4,
                                            bool rope, float* v) {
#pragma unroll
  for (int g = 0; g < HD / 4; ++g) {
    const float4 b = bias4[g];
    v[4 * g] = __uint_as_float(acc[4 * g]) + b.x;
    v[4 * g + 1] = __uint_as_float(acc[4 * g + 1]) + b.y;
    v[4 * g + 2] = __uint_as_float(acc[4 * g + 2]) + b.z;
    v[4 * g + 3] = __uint_as_float(acc[4 * g + 3]) + b.w;
  }
  if (rope) {
#pragma unroll
    for (int g = 0; g < 3; ++g) {
      const float cc[4] = {c4[g].x, c4[g].y, c4[g].z, c4[g].w}, ss[4] = {s4[g].x, s4[g].y, s4[g].z, s4[g].w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int i = 4 * g + e;
        const float a = v[i], b = v[i + HD / 2];
        v[i] = a * cc[e] - b * ss[e];
        v[i + HD / 2] = b * cc[e] + a * ss[e];
      }
    }
  }
}

// One token row x one head: acc = [q | k | v] accumulator columns (fp32 bit patterns), bias = the 72 projection biases.
// Returns the attention output of the row in out[24].  All 32 lanes of the warp must call it (shuffles).
// Written for instruction-level parallelism (three epilogue warps per scheduler hide little latency): k and v are packed
// before q is formed (register pressure), every table comes in through 128-bit loads, the 4 scores accumulate in 8
// independent chains and the shuffles of one element pair are issued together.
// 24 accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld24(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  tmem_ld_32x32b_x8(taddr + 16, r + 16);
}

// `taddr` = TMEM address of the head's 72 columns in this thread's lane; `xch` = the warp's exchange tile; `release` is
// called once the accumulator has been read completely (the caller hands the TMEM stage back to the MMA issuer there).
// All 32 lanes of the warp must call it.
template <typename Release>
__device__ __forceinline__ void attend_row(uint32_t taddr, const float* bias, const float* cs, const float* sn,
                                           const float* bias_kr, const float* bias_vs, float valid, int lane, bool rope,
                                           uint8_t* xch, float* out, Release release) {
  const int t = lane & 3, base = lane & ~3;
  const float4* bias4 = reinterpret_cast<const float4*>(bias);
  float4 c4[3], s4[3];
#pragma unroll
  for (int g = 0; g < 3; ++g) {
    c4[g] = reinterpret_cast<const float4*>(cs + t * (HD / 2))[g];
    s4[g] = reinterpret_cast<const float4*>(sn + t * (HD / 2))[g];
  }
  float4* mine = reinterpret_cast<float4*>(xch + lane * XCH_PITCH);
  int pos[XCH_CHUNKS];                   // position of chunk c in the rows of this lane's sequence
#pragma unroll
  for (int c = 0; c < XCH_CHUNKS; ++c) {
    const int v = c + (lane >> 2);
    pos[c] = v >= XCH_CHUNKS ? v - XCH_CHUNKS : v;
  }
  float q[HD];
  {
    uint32_t acc[HD], accv[HD];
    tmem_ld24(taddr + HD, acc);          // k
    tmem_ld24(taddr + 2 * HD, accv);     // v (in flight while k is rotated)
    tmem_ld_wait();
    __syncwarp();                        // every lane has finished reading the previous tile's exchange rows
    {
      float k[HD];
      bias_rope24(acc, bias4 + HD / 4, c4, s4, rope, k);
#pragma unroll
      for (int g = 0; g < HD / 4; ++g) mine[pos[g]] = make_float4(k[4 * g], k[4 * g + 1], k[4 * g + 2], k[4 * g + 3]);
    }
    tmem_ld24(taddr, acc);               // q (in flight while v is stored)
#pragma unroll
    for (int g = 0; g < HD / 4; ++g) {
      const float4 b = bias4[2 * (HD / 4) + g];
      mine[pos[HD / 4 + g]] = make_float4(__uint_as_float(accv[4 * g]) + b.x, __uint_as_float(accv[4 * g + 1]) + b.y,
                                     __uint_as_float(accv[4 * g + 2]) + b.z, __uint_as_float(accv[4 * g + 3]) + b.w);
    }
    tmem_ld_wait();
    release();
    bias_rope24(acc, bias4, c4, s4, rope, q);
  }
  __syncwarp();                          // the 4 rows of every sequence are visible
  // ---- scores against the 4 tokens of the sequence and the bias token; fp32 softmax (mha.py:359-381)
  float sc[SEQ + 1];
#pragma unroll
  for (int j = 0; j < SEQ; ++j) {
    const float4* kj = reinterpret_cast<const float4*>(xch + (base + j) * XCH_PITCH);
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
    for (int g = 0; g < HD / 4; ++g) {
      const float4 b = kj[pos[g]];
      s0 = fmaf(q[4 * g], b.x, s0);
      s1 = fmaf(q[4 * g + 1], b.y, s1);
      s2 = fmaf(q[4 * g + 2], b.z, s2);
      s3 = fmaf(q[4 * g + 3], b.w, s3);
    }
    sc[j] = (s0 + s1) + (s2 + s3);
  }
  {
    const float4* bk4 = reinterpret_cast<const float4*>(bias_kr);
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
    for (int g = 0; g < HD / 4; ++g) {
      const float4 b = bk4[g];
      s0 = fmaf(q[4 * g], b.x, s0);
      s1 = fmaf(q[4 * g + 1], b.y, s1);
      s2 = fmaf(q[4 * g + 2], b.z, s2);
      s3 = fmaf(q[4 * g + 3], b.w, s3);
    }
    sc[SEQ] = (s0 + s1) + (s2 + s3);
  }
  float m = sc[SEQ];
#pragma unroll
  for (int j = 0; j < SEQ; ++j) {
    const float vj = __shfl_sync(0xffffffffu, valid, base + j);
    sc[j] = vj != 0.f ? sc[j] : -INFINITY;
    m = fmaxf(m, sc[j]);
  }
  float l = 0.f;
#pragma unroll
  for (int j = 0; j <= SEQ; ++j) {
    sc[j] = __expf(sc[j] - m);
    l += sc[j];
  }
  const float inv = __frcp_rn(l);
#pragma unroll
  for (int j = 0; j <= SEQ; ++j) sc[j] *= inv;
  // ---- P.V (24 independent accumulators)
  {
    const float4* bv4 = reinterpret_cast<const float4*>(bias_vs);
#pragma unroll
    for (int g = 0; g < HD / 4; ++g) {
      const float4 b = bv4[g];
      out[4 * g] = sc[SEQ] * b.x;
      out[4 * g + 1] = sc[SEQ] * b.y;
      out[4 * g + 2] = sc[SEQ] * b.z;
      out[4 * g + 3] = sc[SEQ] * b.w;
    }
  }
#pragma unroll
  for (int j = 0; j < SEQ; ++j) {
    const float4* vj = reinterpret_cast<const float4*>(xch + (base + j) * XCH_PITCH);
#pragma unroll
    for (int g = 0; g < HD / 4; ++g) {
      const float4 b = vj[pos[HD / 4 + g]];
      out[4 * g] = fmaf(sc[j], b.x, out[4 * g]);
      out[4 * g + 1] = fmaf(sc[j], b.y, out[4 * g + 1]);
      out[4 * g + 2] = fmaf(sc[j], b.z, out[4 * g + 2]);
      out[4 * g + 3] = fmaf(sc[j], b.w, out[4 * g + 3]);
    }
  }
}


__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS, 1)
gemm_qkv_attn4_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapW, int M, int K,
                      int num_n, Params p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) __trap();     // dynamic shared memory starts at the window base (no static shared)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  float* rope_cs = reinterpret_cast<float*>(smem + OFF_ROPE);
  float* rope_sn = rope_cs + (SEQ + 1) * (HD / 2);
  float* bias_kr = reinterpret_cast<float*>(smem + OFF_BK);
  float* bias_vs = reinterpret_cast<float*>(smem + OFF_BV);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, tid = threadIdx.x;
  const uint32_t rank = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  const int num_m = (M + TILE_M - 1) / TILE_M;
  const int num_tiles = num_m * num_n;
  const int num_kb = (K + BLOCK_K - 1) / BLOCK_K;
  const bool rope = p.cs != nullptr;
  // n fastest: the clusters running at the same time share A rows through L2
  auto get_tile = [&](int i, int& mb, int& nb) -> bool {
    const int tile = cluster_id + i * num_clusters;
    mb = tile / num_n;
    nb = tile % num_n;
    return tile < num_tiles;
  };

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapW) : "memory");
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 2 * EPI_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {   // the same warp of both CTAs allocates the pair's tensor memory (all 512 columns)
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_slot)), "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  pdl_wait();          // common.cuh: nothing in global memory is touched before this
  // step-invariant tables: RoPE cos / sin of positions 0..4, the bias token (key rotated to position 4, value)
  for (int i = tid; i < (SEQ + 1) * (HD / 2); i += THREADS) {
    rope_cs[i] = rope ? p.cs[i] : 1.f;
    rope_sn[i] = rope ? p.sn[i] : 0.f;
  }
  for (int i = tid; i < p.H * HD; i += THREADS) bias_vs[i] = p.bias_v[i];
  {
    float* bias_all = reinterpret_cast<float*>(smem + OFF_BIAS);
    for (int i = tid; i < num_n * TILE_N; i += THREADS) bias_all[i] = p.bias[i];
  }
  if (tid >= 64 && tid < 64 + p.H) {
    const int h = tid - 64;
    float kk[HD];
#pragma unroll
    for (int i = 0; i < HD; ++i) kk[i] = p.bias_k[h * HD + i];
    if (rope) {
#pragma unroll
      for (int i = 0; i < HD / 2; ++i) {
        const float c = p.cs[SEQ * (HD / 2) + i], s = p.sn[SEQ * (HD / 2) + i];
        const float a = kk[i], b = kk[i + HD / 2];
        kk[i] = a * c - b * s;
        kk[i + HD / 2] = b * c + a * s;
      }
    }
#pragma unroll
    for (int i = 0; i < HD; ++i) bias_kr[h * HD + i] = kk[i];
  }
  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();       // the peer's barriers are initialised before anything is signalled across the pair
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;
  pdl_trigger();            // after the TMEM allocation

  if (warp == 0) {
    // ===== TMA producer: own 128 rows of A, own half (112 rows) of the head-major W tile =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int mb, nb;
      for (int i = 0; get_tile(i, mb, nb); ++i) {
        const int arow = mb * TILE_M + (int)rank * BLOCK_M;
        const int wrow = nb * TILE_N + (int)rank * HALF_N;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
          const uint32_t fb = map_to_rank(smem_u32(&full_bar[stage]), 0);
          tma_load_2d_pair(&tmapA, fb, sa, kb * BLOCK_K, arow);
          tma_load_2d_pair(&tmapW, fb, sa + A_STAGE_BYTES, kb * BLOCK_K, wrow);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: warp 1 of the leader CTA walks the loop, one elected lane drives both tensor cores (tc::elect_one) =====
    if (rank == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      int mb, nb;
      const uint32_t idesc = make_idesc_bf16(TILE_M, TILE_N);
      for (int i = 0; get_tile(i, mb, nb); ++i) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tcgen05_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * ACC_STRIDE);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tcgen05_fence_after();
          const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
          const uint64_t adesc = make_smem_desc_sw128(sa);
          const uint64_t bdesc = make_smem_desc_sw128(sa + A_STAGE_BYTES);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
              umma_bf16_pair(d_tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) != 0 ? 1u : 0u);
            tcgen05_commit_pair(&empty_bar[stage]);   // frees this stage in both CTAs
            if (kb == num_kb - 1) tcgen05_commit_pair(&tmem_full[acc]);
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
    // ===== epilogue warps 2..13: TMEM lane quarter = warp % 4, head of the tile = (warp - 2) / 4 =====
    const int quarter = warp & 3, hh = (warp - 2) >> 2;
    const float* bias_all = reinterpret_cast<const float*>(smem + OFF_BIAS);
    int acc = 0;
    uint32_t acc_phase = 0;
    int mb, nb;
    // key-valid flag of this thread's row, prefetched one tile ahead (the only per-tile global read of the epilogue)
    auto row_of = [&](int mbb) { return mbb * TILE_M + (int)rank * BLOCK_M + quarter * 32 + lane; };
    auto valid_of = [&](int rw) { return rw < M ? __ldg(p.key_valid + (size_t)rw % (size_t)p.mask_mod) : 0.f; };
    float valid_next = get_tile(0, mb, nb) ? valid_of(row_of(mb)) : 0.f;
    for (int i = 0; get_tile(i, mb, nb); ++i) {
      const int head = nb * HEADS_PER_TILE + hh;
      const int row = row_of(mb);
      const uint32_t te = map_to_rank(smem_u32(&tmem_empty[acc]), 0);
      const bool live = head < p.H;                        // the last tile of 16 heads holds a single head
      const float valid = valid_next;
      const float* bias_s = bias_all + nb * TILE_N + hh * HEAD_COLS;
      {
        int mb2, nb2;
        valid_next = get_tile(i + 1, mb2, nb2) ? valid_of(row_of(mb2)) : 0.f;
      }
      mbar_wait(&tmem_full[acc], acc_phase);
      tcgen05_fence_after();
      auto release = [&]() {                               // accumulator read: hand the TMEM stage back to the leader
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(te);
      };
      if (!live || p.dbg_skip) release();
      if (live && !p.dbg_skip) {
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * ACC_STRIDE + hh * HEAD_COLS);
        float o[HD];
        attend_row(taddr, bias_s, rope_cs, rope_sn, bias_kr + head * HD, bias_vs + head * HD, valid, lane, rope,
                   smem + OFF_XCH + (warp - 2) * XCH_WARP_BYTES, o, release);
        if (row < M) {
          uint4* dst = reinterpret_cast<uint4*>(p.att + (size_t)row * p.ldo + head * HD);
#pragma unroll
          for (int g2 = 0; g2 < 3; ++g2)
            dst[g2] = make_uint4(pack2(o[8 * g2], o[8 * g2 + 1]), pack2(o[8 * g2 + 2], o[8 * g2 + 3]),
                                 pack2(o[8 * g2 + 4], o[8 * g2 + 5]), pack2(o[8 * g2 + 6], o[8 * g2 + 7]));
        }
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();       // the peer's shared memory and barriers stay valid until both CTAs are done
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

// fp32 packed qkv weight [3C, C] (q rows pre-scaled) / bias [3C] -> head-major bf16 [num_n * 224, C] / fp32 [num_n * 224]
__global__ void pack_qkv_headmajor_kernel(const float* __restrict__ w, const float* __restrict__ b, int H, int C,
                                          bf16* __restrict__ wp, float* __restrict__ bp, int rows_out) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)rows_out * C) return;
  const int ro = (int)(i / C), c = (int)(i % C);
  const int tile = ro / TILE_N, rt = ro % TILE_N;
  const int hh = rt / HEAD_COLS, e = rt % HEAD_COLS, head = tile * HEADS_PER_TILE + hh;
  float v = 0.f, bv = 0.f;
  if (rt < HEADS_PER_TILE * HEAD_COLS && head < H) {
    const int src = (e / HD) * C + head * HD + (e % HD);        // part (q / k / v) * C + head * 24 + element
    v = w[(size_t)src * C + c];
    bv = b[src];
  }
  wp[i] = __float2bfloat16_rn(v);
  if (c == 0) bp[ro] = bv;
}

}  // namespace tqa
}  // namespace mdgen
