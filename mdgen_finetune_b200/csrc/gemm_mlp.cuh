// gemm_mlp.cuh -- the whole MLP block of a LatentMDGenLayer in ONE kernel (latent_model.py:393, layers.py:78-85):
//   hid      = GELU(h . W1^T + b1)                                     h = the LayerNorm + modulate output of the previous kernel
//   x[m,:]  += gate[b] * (hid . W2^T + b2)
//   h'[m,:]  = LN(x[m,:]) * (1 + scale'[b]) + shift'[b]   as bf16      (the next block's LayerNorm + modulate)
// for C = 384, F = 1536.  Why: as two kernels (fc1 + GELU on the pair GEMM, fc2 + residual + LayerNorm in gemm_ln.cuh) the
// block costs 47.5 + 52 us on config 3 for 2 x 27 us of MMA: fc1 is bound by its GELU epilogue, fc2 cannot overlap its
// LayerNorm epilogue with MMAs, and the hidden activations make a 196 MB round trip through HBM.  Here the hidden
// activations never leave the SM and the GELU of one 128-column chunk runs under the MMAs of its neighbours.
//
// One CTA pair (cta_group::2, M = 256 rows) per tile:
//   * the tile's h rows stay resident in shared memory (96 KB per CTA, loaded once per tile);
//   * the hidden dimension is walked in 12 chunks of 128 columns.  Per chunk j:
//       fc1(j):  acc1[128 cols] = h . W1[j]^T            K = 384, N = 128      TMEM columns 384..511
//       GELU(j): acc1 -> + b1 -> erf-GELU -> bf16 -> shared memory (K-major SWIZZLE_128B, 32 KB per CTA): the A operand of
//       fc2(j):  acc2[384 cols] += hid[j] . W2[:, j]^T   K = 128, N = 256 + 128 TMEM columns 0..383
//     issued as fc1(0), fc1(1), fc2(0), fc1(2), fc2(1), ...: the tensor pipe always has the other GEMM's chunk to work on
//     while the 8 epilogue warps turn acc1 into the next A operand (both acc1 and the hidden chunk are single-buffered);
//   * W1 and W2 stream through two independent rings of six 8 KB stages (64 rows x 64 k each; a W1 k-block is one
//     stage, a W2 k-block three: rows [0, 128) of this CTA's half for the N = 256 MMA, 64 rows for the N = 128 MMA), each
//     fed by its own producer thread (warp 0: h rows + W1, warp 10: W2) -- one whole chunk of look-ahead per GEMM.  One
//     shared ring (3 x 24 KB, then 9 x 8 KB) stalled every chunk: the W2 stages parked in it until the GELU had published
//     its chunk kept the next W1 chunk from being prefetched (tools/mlp_trace.py: 4800 / 6000 clocks per chunk);
//   * after chunk 11 the residual + LayerNorm epilogue of gemm_ln.cuh (tcln::LnEpilogue) runs on acc2; its 64 KB of I/O
//     tiles and its per-column vectors alias the h rows, which are dead by then (the next tile's h load waits for the
//     epilogue; the vectors are re-staged every tile).
// Shared-memory traffic per chunk per CTA: 256 KB of MMA operand reads (fc1: h 96 + W1 48; fc2: the hidden chunk twice --
// once per MMA shape -- 64 + W2 48) + 96 KB of TMA writes + 32 KB of hidden-chunk stores = 384 KB in 3072 MMA clocks =
// 125 B/clk of the ~128 B/clk port: the A rows being resident is what makes this fit at all, and the measured period is
// 3400 clocks per chunk (fc1 runs at 82 % of its MMA rate while the port is that busy; tools/mlp_trace.py).
#pragma once
#include "gemm_ln.cuh"

namespace mdgen {
namespace tcmlp {

using namespace tc;
using tc2::cluster_ctarank;
using tc2::cluster_sync_all;
using tc2::map_to_rank;
using tc2::mbar_arrive_cluster;
using tc2::tcgen05_commit_pair;
using tc2::tma_load_2d_pair;
using tc2::umma_bf16_pair;
using tc2::stamp;
using tcln::CV_BYTES;
using tcln::EBUF_BYTES;
using tcln::LnFuse;
using tcln::STATS_BYTES;

constexpr int C = 384, F = 1536, TILE_M = 256;
constexpr int CHUNK = 128, NCHUNK = F / CHUNK;               // hidden columns per step
constexpr int KB1 = C / BLOCK_K;                             // 6 k-blocks of fc1
constexpr int KB2 = CHUNK / BLOCK_K;                         // 2 k-blocks of fc2 per chunk
constexpr int A1_BYTES = KB1 * A_STAGE_BYTES;                // 96 KB: this CTA's 128 rows of h
constexpr int H_BYTES = KB2 * A_STAGE_BYTES;                 // 32 KB: this CTA's 128 rows of the hidden chunk
constexpr int W1_ROWS = CHUNK / 2;                           // 64 rows of the W1 chunk per CTA
constexpr int W2A_ROWS = 128, W2B_ROWS = 64;                 // W2 rows per CTA for the N = 256 and the N = 128 MMA
constexpr int BOX_ROWS = 64;                                 // every W load is a 64-row x 64-k box = one stage
constexpr int STAGE_BYTES = BOX_ROWS * BLOCK_K * 2, STAGES = 6;   // per ring
constexpr int W2_STAGES = (W2A_ROWS + W2B_ROWS) / BOX_ROWS;  // 3 consecutive stages per W2 k-block
static_assert(W1_ROWS == BOX_ROWS && STAGES % W2_STAGES == 0, "a W2 k-block never wraps around its ring");
constexpr int THREADS = NUM_THREADS + 32;                    // + warp 10: the W2 producer
constexpr int OFF_A1 = 0, OFF_EBUF = 0;                      // the epilogue's I/O tiles and vectors alias the (dead) h rows
constexpr int OFF_CV = OFF_EBUF + NUM_EPI_WARPS * 2 * EBUF_BYTES;
constexpr int OFF_H = OFF_A1 + A1_BYTES;
constexpr int OFF_RING1 = OFF_H + H_BYTES;
constexpr int OFF_RING2 = OFF_RING1 + STAGES * STAGE_BYTES;
constexpr int OFF_B1 = OFF_RING2 + STAGES * STAGE_BYTES;     // [8 warps][64] the current chunk's fc1 bias
constexpr int OFF_BAR = OFF_B1 + NUM_EPI_WARPS * 256;
constexpr int OFF_TRACE = OFF_BAR + 512;                      // 64 clock stamps of the traced epilogue thread (tools/mlp_trace.py)
constexpr int SMEM_BYTES = OFF_TRACE + 512;
static_assert(OFF_CV + CV_BYTES + STATS_BYTES <= A1_BYTES, "epilogue tiles and vectors must fit in the h rows they alias");
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
static_assert(OFF_H % 1024 == 0 && OFF_RING1 % 1024 == 0 && OFF_RING2 % 1024 == 0 && STAGE_BYTES % 1024 == 0, "SWIZZLE_128B bases");
constexpr uint32_t ACC1_COL = 384;

__device__ __forceinline__ void mbar_wait_cluster(uint64_t* bar, uint32_t parity) {      // acquire at cluster scope
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "DONE:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// Publishes this thread's (proxy-fenced) shared-memory writes to the MMA issuer in the leader CTA: release at CTA scope on a
// barrier that may be remote -- the form CUTLASS's ClusterBarrier::arrive(cta_id) emits.  The data itself never crosses
// the pair: each SM's tensor core reads its OWN CTA's rows of the A operand, and fence.proxy.async has made this thread's
// stores visible to that SM's async proxy before the arrival is sent.  A release at CLUSTER scope costs a GPU-scope
// membar here: 1850 clocks per chunk on the critical path (tools/mlp_trace.py: 4250 -> 3400 clocks per chunk without it).
__device__ __forceinline__ void mbar_arrive_release_remote(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cta.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS, 1)
gemm_mlp_ln_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapW1,
                   const __grid_constant__ CUtensorMap tmapW2, const __grid_constant__ CUtensorMap tmapX, const __grid_constant__ CUtensorMap tmapH, int M,
                   const float* __restrict__ b1, EpiParams ep, LnFuse ln) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  float* cv = reinterpret_cast<float*>(smem + OFF_CV);
  float* b1s = reinterpret_cast<float*>(smem + OFF_B1);
  uint64_t* full1 = reinterpret_cast<uint64_t*>(smem + OFF_BAR);      // W1 ring
  uint64_t* empty1 = full1 + STAGES;
  uint64_t* full2 = empty1 + STAGES;                                  // W2 ring
  uint64_t* empty2 = full2 + STAGES;
  uint64_t* a1_full = empty2 + STAGES;
  uint64_t* a1_empty = a1_full + 1;
  uint64_t* acc1_full = a1_empty + 1;
  uint64_t* acc1_empty = acc1_full + 1;
  uint64_t* h_full = acc1_empty + 1;
  uint64_t* h_empty = h_full + 1;
  uint64_t* tmem_full = h_empty + 1;                         // acc2
  uint64_t* tmem_empty = tmem_full + 1;
  uint64_t* xbar = tmem_empty + 1;                           // [8 warps][2] residual-tile arrival barriers
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(xbar + 2 * NUM_EPI_WARPS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  const int num_tiles = (M + TILE_M - 1) / TILE_M;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapW1) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapW2) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapX) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapH) : "memory");
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full1[s], 1);
      mbar_init(&empty1[s], 1);
      mbar_init(&full2[s], 1);
      mbar_init(&empty2[s], 1);
    }
    mbar_init(a1_full, 1);
    mbar_init(a1_empty, 1 + NUM_EPI_WARPS);      // the commit after fc1(11) + this CTA's epilogue warps after the LayerNorm
    mbar_init(acc1_full, 1);
    mbar_init(acc1_empty, 2 * NUM_EPI_WARPS);
    mbar_init(h_full, 2 * NUM_EPI_WARPS);
    mbar_init(h_empty, 1);
    mbar_init(tmem_full, 1);
    mbar_init(tmem_empty, 2 * NUM_EPI_WARPS);
    for (int i = 0; i < 2 * NUM_EPI_WARPS; ++i) mbar_init(&xbar[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_slot)), "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;
  pdl_trigger();       // common.cuh: after the TMEM allocation
  pdl_wait();          // the previous kernel's outputs (h, the residual stream) are complete from here on

  if (warp == 0) {
    // ===== TMA producer 1: this CTA's 128 rows of h once per tile, its half of every W1 chunk =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0, a1_phase = 0;
      auto load_w1 = [&](int j) {          // chunk j of W1: rows j*128 + rank*64 .. +64, one stage per k-block
        for (int kb = 0; kb < KB1; ++kb) {
          mbar_wait(&empty1[stage], phase ^ 1);
          if (rank == 0) mbar_expect_tx(&full1[stage], 2 * STAGE_BYTES);
          tma_load_2d_pair(&tmapW1, map_to_rank(smem_u32(&full1[stage]), 0), smem + OFF_RING1 + stage * STAGE_BYTES, kb * BLOCK_K,
                           j * CHUNK + (int)rank * W1_ROWS);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      };
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        const int arow = tile * TILE_M + (int)rank * BLOCK_M;
        load_w1(0);                         // may run ahead of the previous tile's LayerNorm epilogue
        mbar_wait(a1_empty, a1_phase ^ 1);  // h rows (and the epilogue tiles aliasing them) are free
        if (rank == 0) mbar_expect_tx(a1_full, 2 * A1_BYTES);
        {
          const uint32_t fb = map_to_rank(smem_u32(a1_full), 0);
          for (int kb = 0; kb < KB1; ++kb) tma_load_2d_pair(&tmapA, fb, smem + OFF_A1 + kb * A_STAGE_BYTES, kb * BLOCK_K, arow);
        }
        a1_phase ^= 1;
        for (int j = 1; j < NCHUNK; ++j) load_w1(j);
      }
    }
  } else if (warp == 10) {
    // ===== TMA producer 2: k columns j*128 .. +128 of W2, rows split as in gemm_ln.cuh, three stages per k-block =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        for (int j = 0; j < NCHUNK; ++j) {
          for (int kb = 0; kb < KB2; ++kb) {
            for (int i = 0; i < W2_STAGES; ++i) {
              mbar_wait(&empty2[stage], phase ^ 1);
              if (rank == 0) mbar_expect_tx(&full2[stage], 2 * STAGE_BYTES);
              const int row = i < 2 ? (int)rank * W2A_ROWS + i * BOX_ROWS : 256 + (int)rank * W2B_ROWS;
              tma_load_2d_pair(&tmapW2, map_to_rank(smem_u32(&full2[stage]), 0), smem + OFF_RING2 + stage * STAGE_BYTES,
                               j * CHUNK + kb * BLOCK_K, row);
              if (++stage == STAGES) {
                stage = 0;
                phase ^= 1;
              }
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA): the whole warp walks the loop, one elected lane issues =====
    if (rank == 0) {
      constexpr uint32_t idesc_fc1 = make_idesc_bf16(TILE_M, CHUNK);
      constexpr uint32_t idesc_a = make_idesc_bf16(TILE_M, 256), idesc_b = make_idesc_bf16(TILE_M, 128);
      const uint32_t sA1 = smem_u32(smem + OFF_A1), sH = smem_u32(smem + OFF_H);
      const uint32_t sRing1 = smem_u32(smem + OFF_RING1), sRing2 = smem_u32(smem + OFF_RING2);
      int stage1 = 0, stage2 = 0;
      uint32_t phase1 = 0, phase2 = 0, a1_phase = 0, acc1_phase = 0, h_phase = 0, acc2_phase = 0;
      int ti = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++ti) {
        mbar_wait(a1_full, a1_phase);
        a1_phase ^= 1;
        tcgen05_fence_after();
        // developer timeline (tools/mlp_trace.py), second tile of CTA ep.dbg_cta: row 0 fc1(j) issue, row 1 fc2(j) issue,
        // row 2 GELU(j) sees acc1, row 3 GELU(j) published, row 4 [h landed, MMAs issued, acc2 seen, LayerNorm done]
        const bool tr = ep.dbg && (int)blockIdx.x == ep.dbg_cta && ti == 1 && lane == 0;
        if (tr) ep.dbg[4 * 16 + 0] = clock64();
        for (int j = 0; j <= NCHUNK; ++j) {
          if (j < NCHUNK) {
            // ---- fc1(j) into acc1
            mbar_wait(acc1_empty, acc1_phase ^ 1);          // GELU(j-1) has read acc1
            acc1_phase ^= 1;
            tcgen05_fence_after();
            if (tr) ep.dbg[0 * 16 + j] = clock64();
            for (int kb = 0; kb < KB1; ++kb) {
              mbar_wait(&full1[stage1], phase1);
              tcgen05_fence_after();
              const uint64_t adesc = make_smem_desc_sw128(sA1 + kb * A_STAGE_BYTES);
              const uint64_t bdesc = make_smem_desc_sw128(sRing1 + stage1 * STAGE_BYTES);
              if (elect_one()) {
#pragma unroll
                for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
                  umma_bf16_pair(tmem_base + ACC1_COL, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc_fc1,
                                 (kb | k) != 0 ? 1u : 0u);
                tcgen05_commit_pair(&empty1[stage1]);
              }
              __syncwarp();
              if (++stage1 == STAGES) {
                stage1 = 0;
                phase1 ^= 1;
              }
            }
            if (elect_one()) {
              tcgen05_commit_pair(acc1_full);
              if (j == NCHUNK - 1) tcgen05_commit_pair(a1_empty);   // every read of the h rows has retired
            }
            __syncwarp();
            if (tr) ep.dbg[10 * 16 + j] = clock64();
          }
          if (j >= 1) {
            // ---- fc2(j-1) from the hidden chunk in shared memory into acc2
            if (j == 1) {
              mbar_wait(tmem_empty, acc2_phase ^ 1);        // the previous tile's LayerNorm epilogue has read acc2
              acc2_phase ^= 1;
            }
            mbar_wait_cluster(h_full, h_phase);             // both CTAs' halves of the hidden chunk are in shared memory
            h_phase ^= 1;
            tcgen05_fence_after();
            if (tr) ep.dbg[1 * 16 + j - 1] = clock64();
            for (int kb = 0; kb < KB2; ++kb) {
              const uint32_t sw = sRing2 + stage2 * STAGE_BYTES;     // three consecutive stages, never wrapping
              for (int i = 0; i < W2_STAGES; ++i) mbar_wait(&full2[stage2 + i], phase2);
              tcgen05_fence_after();
              const uint64_t adesc = make_smem_desc_sw128(sH + kb * A_STAGE_BYTES);
              const uint64_t b1desc = make_smem_desc_sw128(sw);
              const uint64_t b2desc = make_smem_desc_sw128(sw + W2A_ROWS * BLOCK_K * 2);
              if (elect_one()) {
#pragma unroll
                for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                  const uint32_t accum = ((j - 1) | kb | k) != 0 ? 1u : 0u;
                  umma_bf16_pair(tmem_base, adesc + (uint64_t)(k * 2), b1desc + (uint64_t)(k * 2), idesc_a, accum);
                  umma_bf16_pair(tmem_base + 256, adesc + (uint64_t)(k * 2), b2desc + (uint64_t)(k * 2), idesc_b, accum);
                }
#pragma unroll
                for (int i = 0; i < W2_STAGES; ++i) tcgen05_commit_pair(&empty2[stage2 + i]);
              }
              __syncwarp();
              if ((stage2 += W2_STAGES) == STAGES) {
                stage2 = 0;
                phase2 ^= 1;
              }
            }
            if (elect_one()) tcgen05_commit_pair(h_empty);
            __syncwarp();
            if (tr) ep.dbg[11 * 16 + j - 1] = clock64();
          }
        }
        if (elect_one()) tcgen05_commit_pair(tmem_full);
        __syncwarp();
        if (tr) ep.dbg[4 * 16 + 1] = clock64();
      }
    }
  } else if (warp < 2 + NUM_EPI_WARPS) {
    // ===== epilogue warps 2..9: GELU of every hidden chunk, then the residual + LayerNorm epilogue =====
    tcln::LnEpilogue epi;
    epi.init(warp, lane, tmem_base, smem + OFF_EBUF, cv, xbar, tmem_full, tmem_empty);
    float* bj = b1s + (warp - 2) * 64;                       // this warp's 64 fc1 bias values of the current chunk
    long long* trace = reinterpret_cast<long long*>(smem + OFF_TRACE);   // stamps stay in shared memory until the tile is done:
                                                                         // global stores would sit in front of the release-arrive
    const int quarter = warp & 3, half = (warp - 2) >> 2;
    const uint32_t tacc1 = tmem_base + ((uint32_t)(quarter * 32) << 16) + ACC1_COL + (uint32_t)(half * 64);
    const uint32_t a1e = map_to_rank(smem_u32(acc1_empty), 0), hf = map_to_rank(smem_u32(h_full), 0);
    // this thread's row of the hidden chunk: k-block `half` (its 64 columns), row quarter*32 + lane, SWIZZLE_128B
    uint8_t* hrow = smem + OFF_H + half * A_STAGE_BYTES + (quarter * 32 + lane) * 128;
    const uint32_t swz = (uint32_t)(lane & 7);
    uint32_t acc1_phase = 0, h_phase = 0;
    int ti = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++ti) {
      const int cta_row0 = tile * TILE_M + (int)rank * BLOCK_M;
      const bool tr = ep.dbg && (int)blockIdx.x == ep.dbg_cta && ti == 1 && warp == 2 && lane == 0;
      if (lane < 16)
        reinterpret_cast<float4*>(bj)[lane] = b1 ? __ldg(reinterpret_cast<const float4*>(b1 + half * 64) + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
      __syncwarp();
#pragma unroll 1
      for (int j = 0; j < NCHUNK; ++j) {
        // the NEXT chunk's 64 bias values travel from global memory while this chunk is processed (lanes 0..15 hold them)
        mbar_wait(acc1_full, acc1_phase);
        acc1_phase ^= 1;
        tcgen05_fence_after();
        if (tr) trace[0 * NCHUNK + j] = clock64();
        // the last fc1 has retired: the h rows are dead, and the first residual tiles of the LayerNorm epilogue (whose
        // buffers alias them) can start moving under the last GELU and the last fc2
        if (j == NCHUNK - 1) epi.prefetch(tmapX, cta_row0);
        uint32_t r[64];
        tmem_ld_32x32b_x32(tacc1, r);
        tmem_ld_32x32b_x32(tacc1 + 32, r + 32);
        float4 bnext = make_float4(0.f, 0.f, 0.f, 0.f);
        if (lane < 16 && b1 && j + 1 < NCHUNK) bnext = __ldg(reinterpret_cast<const float4*>(b1 + (j + 1) * CHUNK + half * 64) + lane);
        tmem_ld_wait();
        tcgen05_fence_before();          // acc1 is in registers: fc1(j+1) may overwrite it
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(a1e);
        // bias, erf-GELU, bf16 into registers while fc2(j-1) still reads the previous hidden chunk ...
        uint32_t pk[32];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float4 ba = *reinterpret_cast<const float4*>(bj + c * 8), bb = *reinterpret_cast<const float4*>(bj + c * 8 + 4);
          const float v0 = gelu_erf_fast(__uint_as_float(r[8 * c]) + ba.x), v1 = gelu_erf_fast(__uint_as_float(r[8 * c + 1]) + ba.y);
          const float v2 = gelu_erf_fast(__uint_as_float(r[8 * c + 2]) + ba.z), v3 = gelu_erf_fast(__uint_as_float(r[8 * c + 3]) + ba.w);
          const float v4 = gelu_erf_fast(__uint_as_float(r[8 * c + 4]) + bb.x), v5 = gelu_erf_fast(__uint_as_float(r[8 * c + 5]) + bb.y);
          const float v6 = gelu_erf_fast(__uint_as_float(r[8 * c + 6]) + bb.z), v7 = gelu_erf_fast(__uint_as_float(r[8 * c + 7]) + bb.w);
          const __nv_bfloat162 p0 = __floats2bfloat162_rn(v0, v1), p1 = __floats2bfloat162_rn(v2, v3);
          const __nv_bfloat162 p2 = __floats2bfloat162_rn(v4, v5), p3 = __floats2bfloat162_rn(v6, v7);
          pk[4 * c] = *reinterpret_cast<const uint32_t*>(&p0);
          pk[4 * c + 1] = *reinterpret_cast<const uint32_t*>(&p1);
          pk[4 * c + 2] = *reinterpret_cast<const uint32_t*>(&p2);
          pk[4 * c + 3] = *reinterpret_cast<const uint32_t*>(&p3);
        }
        if (tr) trace[2 * NCHUNK + j] = clock64();
        // ... then this thread's row of the A operand (8 chunks of 16 bytes) once fc2(j-1) has retired
        mbar_wait(h_empty, h_phase ^ 1);
        h_phase ^= 1;
#pragma unroll
        for (int c = 0; c < 8; ++c)
          *reinterpret_cast<uint4*>(hrow + (((uint32_t)c ^ swz) << 4)) = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
        fence_async_smem();                    // visible to the tensor core's (async proxy) reads
        if (tr) trace[1 * NCHUNK + j] = clock64();
        __syncwarp();                          // ... and every lane is done with this chunk's bias
        if (lane == 0) mbar_arrive_release_remote(hf);
        if (tr) trace[3 * NCHUNK + j] = clock64();
        if (lane < 16) reinterpret_cast<float4*>(bj)[lane] = bnext;
        __syncwarp();
        if (tr) trace[4 * NCHUNK + j] = clock64();
      }
      // ---- acc2 complete (which also means every MMA that read the h rows has retired: the I/O tiles may overwrite them)
      mbar_wait(tmem_full, epi.acc_phase);
      tcgen05_fence_after();
      if (tr) ep.dbg[4 * 16 + 2] = clock64();
      epi.staged_q = -1;                       // the per-column vectors share the h rows' memory: staged again every tile
      epi.tile(tmapX, tmapH, ep, ln, cta_row0, ti + 8);       // its stamps go to rows 8.. of the timeline buffer
      if (tr) {
        ep.dbg[4 * 16 + 3] = clock64();
        const int rows[5] = {2, 5, 7, 6, 3};
        for (int q = 0; q < 5; ++q)
          for (int j = 0; j < NCHUNK; ++j) ep.dbg[rows[q] * 16 + j] = trace[q * NCHUNK + j];
      }
      if (lane == 0) {
        bulk_wait_read<0>();                   // the last h' tiles have left shared memory: the next tile's h may land
        mbar_arrive(a1_empty);
      }
      __syncwarp();
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

}  // namespace tcmlp
}  // namespace mdgen
