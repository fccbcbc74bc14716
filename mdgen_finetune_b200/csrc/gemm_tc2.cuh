// gemm_tc2.cuh -- the CTA-pair (tcgen05 cta_group::2) variant of gemm_tc.cuh for the large projections.
//   C[M,N] = A[M,K] . W[N,K]^T (+ fused epilogue), 256 x 256 output tiles owned by a cluster of two CTAs.
// A single CTA issuing 128 x 128 MMAs has to pull (128 + 128) * 32 B of operands out of shared memory per 64-clock
// MMA while TMA writes the same amount in: the mainloop saturates at ~50 % of the tensor peak (measured 690 TFLOP/s).
// In a pair each CTA stages only its own 128 rows of A and its own half (128 rows) of the W tile, the leader's
// MMA (M = 256, N = 256) reads both CTAs' shared memory, and each CTA's TMEM receives its own 128 accumulator rows:
// operand traffic per SM halves.
// Roles per CTA (320 threads): warp 0 = TMA producer (both CTAs), warp 1 = MMA issuer (leader CTA only),
// warps 2..9 = epilogue (both CTAs, own TMEM).  Barrier protocol:
//   full[s]      lives in the leader; leader's producer arms it with the bytes of BOTH CTAs, both CTAs' TMA complete on it
//   empty[s]     one per CTA; the leader's tcgen05.commit multicasts the arrival to both
//   tmem_full[a] one per CTA; multicast commit after the last MMA of a tile
//   tmem_empty[a] lives in the leader; the 8 epilogue warps of both CTAs arrive on it (remote arrive from the peer)
// Only the staged (shared-memory transposed, coalesced) epilogues of gemm_tc.cuh are instantiated here.
#pragma once
#include "gemm_tc.cuh"

namespace mdgen {
namespace tc2 {

using namespace tc;

constexpr int TILE_M = 256, TILE_N = 256;
constexpr int HALF_N = TILE_N / 2;                       // W rows staged per CTA
constexpr int STAGES2 = 5;
constexpr int STAGE2_BYTES = A_STAGE_BYTES + HALF_N * BLOCK_K * 2;   // 32 KB per CTA per stage
constexpr int SMEM2_BYTES = STAGES2 * STAGE2_BYTES + EPI_SMEM_BYTES + 1024 + 256;
static_assert(SMEM2_BYTES <= 227 * 1024, "shared memory budget");

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local` (a shared::cta address) inside CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_rank(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
// Relaxed on purpose: the arrival only hands a TMEM accumulator stage back (ordered by tcgen05.wait::ld +
// tcgen05.fence::before_thread_sync).  A release at cluster scope would first drain the warp's outstanding global
// stores -- measured as the dominant epilogue stall (ERRBAR / membar) in the first version of this kernel.
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load into this CTA's shared memory, completion bytes signalled on a barrier that may live in the peer CTA
__device__ __forceinline__ void tma_load_2d_pair(const CUtensorMap* map, uint32_t bar_cluster_addr, void* dst, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst)),
      "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at the same shared-memory offset in both CTAs of the pair once all prior MMAs retire
__device__ __forceinline__ void tcgen05_commit_pair(uint64_t* bar) {
  const uint16_t mask = 3;
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"(mask)
               : "memory");
}

// kernel-developer timeline (tools/gemm_trace.py): 16 slots per tile for the first 16 tiles of CTA ep.dbg_cta
__device__ __forceinline__ void stamp(const EpiParams& ep, int tile_i, int slot) {
  if (ep.dbg && (int)blockIdx.x == ep.dbg_cta && tile_i < 16) ep.dbg[tile_i * 16 + slot] = clock64();
}
__device__ __forceinline__ void stamp_ns(const EpiParams& ep, int tile_i, int slot) {     // wall clock (globaltimer)
  if (ep.dbg && (int)blockIdx.x == ep.dbg_cta) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    ep.dbg[tile_i * 16 + slot] = (long long)t;
  }
}

// EW = epilogue warps per CTA.  8: two per TMEM lane quarter, 128 columns each (all epilogues).  16 (bf16 store / GELU
// only): four per quarter, 64 columns each as two 32-column chunks through two alternating 32 x 32 SWIZZLE_64B store tiles.
// Why: with K = 384 draining a 256 x 256 accumulator through the 64 B/clk TMEM read port takes 2048 cycles per tile against
// 3072 cycles of MMA; on 8 warps that drain, the GELU (~11 instructions + 1 MUFU per element) and the staging run one
// after the other in each warp and the epilogue, not the tensor pipe, sets the tile time.  (16 warps on the SINGLE-CTA
// kernel change nothing -- gemm_tc.cuh STG 3 -- because its 128 x 128 mainloop is shared-memory-bandwidth bound.)
constexpr int STAGES2_EW16 = 4;      // one operand stage goes to the second store tile of every warp
constexpr int SMEM2_EW16_BYTES = STAGES2_EW16 * STAGE2_BYTES + 1024 + EPI16_WARPS * (2 * TMA_TILE32_BYTES + 256) + 1024 + 256;
static_assert(SMEM2_EW16_BYTES <= 227 * 1024, "shared memory budget");
__host__ __device__ constexpr int num_threads2(int ew) { return 64 + 32 * ew; }

template <int EPI, typename OT, int EW = NUM_EPI_WARPS>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(num_threads2(EW), 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapW,
                const __grid_constant__ CUtensorMap tmapC, int M, int N, int K, EpiParams ep) {
  constexpr bool kResid = (EPI == EPI_GATE_RESID || EPI == EPI_RESID);
  constexpr bool kBf16 = (EPI == EPI_STORE || EPI == EPI_GELU) && sizeof(OT) == 2;
  static_assert(kResid || kBf16, "gemm_tc2: only the staged epilogues exist");
  static_assert(EW == NUM_EPI_WARPS || (EW == EPI16_WARPS && kBf16), "16 epilogue warps: bf16 store / GELU only");
  constexpr int STAGES2 = EW == EPI16_WARPS ? STAGES2_EW16 : tc2::STAGES2;
  // No integer round trip on this pointer: the compiler must keep seeing a shared-space address, otherwise every
  // staging access becomes a generic LD.E / ST.E (long-scoreboard latency instead of LDS / STS).
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) __trap();     // dynamic shared memory starts at the window base (no static shared)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES2 * STAGE2_BYTES);
  uint64_t* empty_bar = full_bar + STAGES2;
  uint64_t* tmem_full = empty_bar + STAGES2;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { stamp(ep, 15, 0); stamp_ns(ep, 15, 3); }      // kernel entry
  if (threadIdx.x == 0 && ep.dbg) {       // every CTA: wall-clock entry (exit below) at dbg[256 + 2 * blockIdx.x]
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    ep.dbg[256 + 2 * blockIdx.x] = (long long)t;
  }
  const uint32_t rank = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  const int num_m = (M + TILE_M - 1) / TILE_M, num_n = (N + TILE_N - 1) / TILE_N;
  const int num_tiles = num_m * num_n;
  const int num_kb = (K + BLOCK_K - 1) / BLOCK_K;
  // n fastest: the clusters running at the same time share A rows through L2
  auto get_tile = [&](int i, int& mb, int& nb, int& n_cur) -> bool {
    const int tile = cluster_id + i * num_clusters;
    mb = tile / num_n;
    nb = tile % num_n;
    n_cur = min(TILE_N, N - nb * TILE_N);       // host guarantees N % 128 == 0: 128 or 256
    return tile < num_tiles;
  };

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapW) : "memory");
    if constexpr (kBf16) asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapC) : "memory");
    for (int s = 0; s < STAGES2; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 2 * EW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {   // the same warp of both CTAs allocates the pair's tensor memory (all 512 columns)
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_slot)), "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();       // the peer's barriers are initialised before anything is signalled across the pair
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;
  pdl_trigger();       // common.cuh: after the TMEM allocation
  pdl_wait();          // the previous kernel's outputs are complete from here on
  if (threadIdx.x == 0) stamp(ep, 15, 1);      // setup done

  if (warp == 0) {
    // ===== TMA producer: own 128 rows of A, own half of the W tile =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int mb, nb, n_cur;
      for (int i = 0; get_tile(i, mb, nb, n_cur); ++i) {
        const int arow = mb * TILE_M + (int)rank * BLOCK_M;
        const int wrow = nb * TILE_N + (int)rank * (n_cur >> 1);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (kb == 0) stamp(ep, i, 12);
          uint8_t* sa = smem + stage * STAGE2_BYTES;
          if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * STAGE2_BYTES);
          const uint32_t fb = map_to_rank(smem_u32(&full_bar[stage]), 0);
          tma_load_2d_pair(&tmapA, fb, sa, kb * BLOCK_K, arow);
          tma_load_2d_pair(&tmapW, fb, sa + A_STAGE_BYTES, kb * BLOCK_K, wrow);
          if (kb == num_kb - 1) stamp(ep, i, 13);
          if (++stage == STAGES2) {
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
      int mb, nb, n_cur;
      for (int i = 0; get_tile(i, mb, nb, n_cur); ++i) {
        const uint32_t idesc = make_idesc_bf16(TILE_M, n_cur);
        if (lane == 0) stamp(ep, i, 0);
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tcgen05_fence_after();
        if (lane == 0) stamp(ep, i, 1);
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * TILE_N);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tcgen05_fence_after();
          if (kb < 6 && lane == 0) stamp(ep, i, 2 + kb);
          const uint32_t sa = smem_u32(smem + stage * STAGE2_BYTES);
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
          if (++stage == STAGES2) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (lane == 0) stamp(ep, i, 8);
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else if constexpr (EW == EPI16_WARPS) {
    // ===== 16 epilogue warps: TMEM lane quarter = warp % 4, column quarter of the tile = (warp - 2) / 4 =====
    const int quarter = warp & 3, part = (warp - 2) >> 2;
    uint8_t* epi = smem + STAGES2 * STAGE2_BYTES + 1024;
    uint8_t* tiles = epi + (warp - 2) * (2 * TMA_TILE32_BYTES);
    float* cv = reinterpret_cast<float*>(epi + EW * 2 * TMA_TILE32_BYTES + (warp - 2) * 256);
    int acc = 0;
    uint32_t acc_phase = 0;
    int mb, nb, n_cur;
    for (int i = 0; get_tile(i, mb, nb, n_cur); ++i) {
      const int cpw = n_cur >> 7;                         // 32-column chunks per warp: 2 (n_cur 256) or 1 (128)
      const int mrow0 = mb * TILE_M + (int)rank * BLOCK_M + quarter * 32;
      const int col_in_tile = part * cpw * 32;
      const int ncol0 = nb * TILE_N + col_in_tile;
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * TILE_N + col_in_tile);
      const uint32_t te = map_to_rank(smem_u32(&tmem_empty[acc]), 0);
      __syncwarp();
      for (int c = lane; c < cpw * 32; c += 32) cv[c] = (ep.bias && ncol0 + c < N) ? __ldg(ep.bias + ncol0 + c) : 0.f;
      __syncwarp();
      mbar_wait(&tmem_full[acc], acc_phase);
      tcgen05_fence_after();
      uint32_t r0[32], r1[32];
      tmem_ld_32x32b_x32(taddr, r0);
      if (cpw == 2) tmem_ld_32x32b_x32(taddr + 32, r1);
      tmem_ld_wait();
      tcgen05_fence_before();                             // accumulator is in registers: hand the TMEM stage back
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(te);
      epilogue_tma_bf16_32<EPI, 1>(&tmapC, tiles, cv, lane, mrow0, ncol0, r0);
      if (cpw == 2) epilogue_tma_bf16_32<EPI, 1>(&tmapC, tiles + TMA_TILE32_BYTES, cv + 32, lane, mrow0, ncol0 + 32, r1);
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (lane == 0) bulk_wait_read<0>();       // shared memory must outlive the last bulk store's read
  } else {
    // ===== epilogue warps 2..9: TMEM lane quarter = warp % 4, column half = (warp - 2) / 4 =====
    const int quarter = warp & 3, half = (warp - 2) >> 2;
    // bf16 outputs leave through TMA stores from 1024-byte aligned SWIZZLE_128B tiles; the residual epilogue keeps its
    // padded transposition tiles
    uint8_t* epi = smem + STAGES2 * STAGE2_BYTES + (kBf16 ? 1024 : 256);
    uint8_t* stg = epi + (warp - 2) * (kBf16 ? TMA_TILE_BYTES : STG_BYTES);
    float* cv = reinterpret_cast<float*>(epi + NUM_EPI_WARPS * (kBf16 ? TMA_TILE_BYTES : STG_BYTES) + (warp - 2) * CV_BYTES_PER_WARP);
    int acc = 0;
    uint32_t acc_phase = 0;
    int mb, nb, n_cur;
    for (int i = 0; get_tile(i, mb, nb, n_cur); ++i) {
      const int cpw = n_cur >> 6;                         // 32-column chunks per warp: 4 (n_cur 256) or 2 (128)
      const int mrow0 = mb * TILE_M + (int)rank * BLOCK_M + quarter * 32;
      const int col_in_tile = half * cpw * 32;
      const int ncol0 = nb * TILE_N + col_in_tile;
      const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * TILE_N + col_in_tile);
      const uint32_t te = map_to_rank(smem_u32(&tmem_empty[acc]), 0);
      if constexpr (kResid) {
        // rolling residual prefetch: chunks c and c+1 are in flight while chunk c is combined and written back
        const ResidLane rl = resid_lane<EPI>(ep, lane, mrow0, ncol0, M, N);
        stage_resid_colvecs<EPI>(ep, cv, lane, mrow0, ncol0, N);
        __syncwarp();
        float4 xv[2][8];
        residual_prefetch(rl, lane, 0, N, xv[0]);
        residual_prefetch(rl, lane, 32, N, xv[1]);
        mbar_wait(&tmem_full[acc], acc_phase);
        tcgen05_fence_after();
#pragma unroll 1
        for (int c = 0; c < cpw; c += 2) {
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            uint32_t rr[32];
            tmem_ld_32x32b_x32(taddr + (c + u) * 32, rr);
            tmem_ld_wait();
            if (c + u == cpw - 1) {                 // accumulator fully read: hand the TMEM stage back to the leader
              tcgen05_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive_cluster(te);
            }
            epilogue_staged_resid<EPI>(rl, cv, stg, lane, (c + u) * 32, N, rr, xv[u]);
            if (c + u + 2 < cpw) residual_prefetch(rl, lane, (c + u + 2) * 32, N, xv[u]);
          }
        }
      } else {
        stage_colvec(cv, ep.bias, ncol0, N, lane, 0.f);
        __syncwarp();
        mbar_wait(&tmem_full[acc], acc_phase);
        tcgen05_fence_after();
        if (warp == 2 && lane == 0) stamp(ep, i, 9);
        // 64 columns (one 128-byte bf16 row segment) per round, the next pair's tcgen05.ld in flight meanwhile
        uint32_t ra[2][32], rb[2][32];
        tmem_ld_32x32b_x32(taddr, ra[0]);
        tmem_ld_32x32b_x32(taddr + 32, ra[1]);
#pragma unroll 1
        for (int c = 0; c < cpw; c += 4) {
          tmem_ld_wait();
          if (c + 2 < cpw) {
            tmem_ld_32x32b_x32(taddr + (c + 2) * 32, rb[0]);
            tmem_ld_32x32b_x32(taddr + (c + 3) * 32, rb[1]);
          } else {                                    // accumulator fully read: hand the TMEM stage back to the leader
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(te);
            if (warp == 2 && lane == 0) stamp(ep, i, 10);
          }
          epilogue_tma_bf16<EPI>(&tmapC, stg, cv + c * 32, lane, mrow0, ncol0 + c * 32, ra);
          if (c + 2 < cpw) {
            tmem_ld_wait();
            if (c + 4 < cpw) {
              tmem_ld_32x32b_x32(taddr + (c + 4) * 32, ra[0]);
              tmem_ld_32x32b_x32(taddr + (c + 5) * 32, ra[1]);
            } else {
              tcgen05_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive_cluster(te);
              if (warp == 2 && lane == 0) stamp(ep, i, 10);
            }
            epilogue_tma_bf16<EPI>(&tmapC, stg, cv + (c + 2) * 32, lane, mrow0, ncol0 + (c + 2) * 32, rb);
          }
        }
      }
      if (warp == 2 && lane == 0) stamp(ep, i, 11);
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if constexpr (kBf16) {
      if (lane == 0) bulk_wait_read<0>();       // shared memory must outlive the last bulk store's read
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();       // the peer's shared memory and barriers stay valid until both CTAs are done
  if (threadIdx.x == 0) { stamp(ep, 15, 2); stamp_ns(ep, 15, 4); }      // kernel exit
  if (threadIdx.x == 0 && ep.dbg) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    ep.dbg[256 + 2 * blockIdx.x + 1] = (long long)t;
  }
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

}  // namespace tc2
}  // namespace mdgen
