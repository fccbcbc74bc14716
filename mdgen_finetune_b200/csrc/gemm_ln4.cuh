// gemm_ln4.cuh -- the fused projection + residual + next LayerNorm of gemm_ln.cuh for SMALL token counts (< 8192 rows:
// BASELINE configs 1 and 2, i.e. one batch element per GPU under strong scaling):
//   x[m,:]  += gate[b] * (A[m,:] . W^T + bias)
//   h[m,:]   = LN(x[m,:]) * (1 + scale'[b]) + shift'[b]   as bf16
// Why a second kernel: gemm_ln_kernel gives a 256-row tile to one CTA pair, whose 8 epilogue warps walk 6 column chunks each
// through a strictly serial load -> update -> store chain (~8 us); at 4000 rows that is 16 tiles on 148 SMs and the
// kernel was SLOWER than a plain GEMM (11.5 us) + the stand-alone LayerNorm (4.8 us).  Here the ROW is split instead:
//   * a cluster of 4 CTAs owns a 128-row tile, CTA r the columns [96 r, 96 r + 96): 4 x as many CTAs (config 1: 128),
//     each with a 128 x 96 accumulator (tcgen05 cta_group::1, M = 128, N = 96);
//   * 12 epilogue warps per CTA = TMEM lane quarter x one 32-column chunk: every warp handles exactly ONE 32 x 32 fp32
//     residual tile (requested by TMA before the MMAs have even started), keeps its 32 new values in registers, stores
//     the tile back through TMA;
//   * LayerNorm statistics: every thread publishes (shifted sum, shifted sum of squares, shift) of its 32 columns to all
//     four CTAs through distributed shared memory, one cluster barrier, then each thread combines the 12 triples of its
//     row (same shifted one-pass formula as tcln::LnEpilogue) and normalises its 32 values straight from registers.
#pragma once
#include "gemm_ln.cuh"

namespace mdgen {
namespace tcln4 {

using namespace tc;
using tc2::cluster_ctarank;
using tc2::cluster_sync_all;
using tc2::map_to_rank;
using tcln::LnFuse;
using tcln::named_bar_sync;

constexpr int CL = 4, TILE_M = 128, LN_N = 384, N_CTA = LN_N / CL;       // 96 columns per CTA
constexpr int CHUNKS = N_CTA / 32;                                          // 3 chunks of 32 columns
constexpr int EPI_WARPS = 4 * CHUNKS, THREADS = 64 + 32 * EPI_WARPS;       // 12 epilogue warps, 448 threads
constexpr int STAGE_BYTES = A_STAGE_BYTES + N_CTA * BLOCK_K * 2;           // 16 KB + 12 KB
constexpr int STAGES = 5;
static_assert(STAGE_BYTES % 1024 == 0, "SWIZZLE_128B tiles need 1024-byte aligned bases");
constexpr int EBUF_BYTES = 4096;                                            // one 32 x 32 fp32 tile (SWIZZLE_128B)
constexpr int NVEC = 7;                                                     // bias, gate[2], 1 + scale[2], shift[2]
constexpr int OFF_EBUF = STAGES * STAGE_BYTES;
constexpr int OFF_CV = OFF_EBUF + EPI_WARPS * EBUF_BYTES;
constexpr int SLOTS = CL * CHUNKS;                                          // 12 statistic triples per row
constexpr int OFF_XCH = OFF_CV + NVEC * N_CTA * 4;                          // [SLOTS][128 rows] float4 (s1, q1, shift, -)
constexpr int OFF_BAR = OFF_XCH + SLOTS * TILE_M * 16;
constexpr int SMEM_BYTES = OFF_BAR + 256;
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");

__device__ __forceinline__ void st_cluster_f4(uint32_t cluster_addr, float4 v) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(cluster_addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__global__ void __cluster_dims__(CL, 1, 1) __launch_bounds__(THREADS, 1)
gemm_ln4_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapW,
                const __grid_constant__ CUtensorMap tmapX, const __grid_constant__ CUtensorMap tmapH, int M, int K, EpiParams ep,
                LnFuse ln) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  float* cv = reinterpret_cast<float*>(smem + OFF_CV);                      // [NVEC][96]
  float4* xch = reinterpret_cast<float4*>(smem + OFF_XCH);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* xbar = tmem_full + 1;                                           // [12 warps] residual-tile arrival
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(xbar + EPI_WARPS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int tile = blockIdx.x / CL;
  const int num_kb = (K + BLOCK_K - 1) / BLOCK_K;
  const int row0 = tile * TILE_M, col0 = (int)rank * N_CTA;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapW) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapX) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapH) : "memory");
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full, 1);
    for (int i = 0; i < EPI_WARPS; ++i) mbar_init(&xbar[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_slot)), "r"(128)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();       // every CTA of the cluster is running before anyone writes into a peer's shared memory
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_base_slot;
  pdl_trigger();            // common.cuh: after the TMEM allocation
  pdl_wait();               // the previous kernel's outputs (A, the residual stream) are complete from here on

  if (warp == 0) {
    // ===== TMA producer: the tile's 128 rows of A and this CTA's 96 rows of W, k-block by k-block =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem + stage * STAGE_BYTES;
        mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
        tma_load_2d(&tmapA, &full_bar[stage], sa, kb * BLOCK_K, row0);
        tma_load_2d(&tmapW, &full_bar[stage], sa + A_STAGE_BYTES, kb * BLOCK_K, col0);
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
    __syncwarp();
    cluster_sync_all();     // the statistics exchange barrier counts every thread of the cluster
  } else if (warp == 1) {
    // ===== MMA issuer (warp-uniform, one elected lane issues) =====
    constexpr uint32_t idesc = make_idesc_bf16(TILE_M, N_CTA);
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < num_kb; ++kb) {
      mbar_wait(&full_bar[stage], phase);
      tcgen05_fence_after();
      const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
      const uint64_t adesc = make_smem_desc_sw128(sa), bdesc = make_smem_desc_sw128(sa + A_STAGE_BYTES);
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
          umma_bf16(tmem_base, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) != 0 ? 1u : 0u);
        tcgen05_commit(&empty_bar[stage]);
        if (kb == num_kb - 1) tcgen05_commit(tmem_full);
      }
      __syncwarp();
      if (++stage == STAGES) {
        stage = 0;
        phase ^= 1;
      }
    }
    cluster_sync_all();
  } else {
    // ===== epilogue warps 2..13: TMEM lane quarter = warp % 4, 32-column chunk = (warp - 2) / 4 =====
    const int e = warp - 2, quarter = warp & 3, chunk = e >> 2;
    const int row_in_cta = quarter * 32 + lane;
    uint8_t* tile_buf = smem + OFF_EBUF + e * EBUF_BYTES;
    uint8_t* myrow = tile_buf + lane * 128;
    const uint32_t swz = (uint32_t)(lane & 7);
    const int gcol = col0 + chunk * 32, grow = row0 + quarter * 32;
    // ---- this warp's residual tile starts moving now, long before the accumulator is ready
    if (lane == 0) {
      mbar_expect_tx(&xbar[e], EBUF_BYTES);
      tma_load_2d(&tmapX, &xbar[e], tile_buf, gcol, grow);
    }
    // ---- per-column vectors of the (at most two) batch elements the tile's 128 rows touch, this CTA's 96 columns
    const int q0 = row0 / ep.rows_per_batch;
    const int split = (q0 + 1) * ep.rows_per_batch - row0;                // rows >= split belong to the next batch element
    {
      const int b0 = q0 % ep.nbatch, b1 = (q0 + 1) % ep.nbatch;
      for (int i = e * 32 + lane; i < NVEC * (N_CTA / 4); i += 32 * EPI_WARPS) {
        const int v = i / (N_CTA / 4), cc = (i % (N_CTA / 4)) * 4;
        float4 val;
        if (v == 0) {
          val = ep.bias ? __ldg(reinterpret_cast<const float4*>(ep.bias + col0 + cc)) : make_float4(0.f, 0.f, 0.f, 0.f);
        } else if (v <= 2) {
          val = ep.gate ? __ldg(reinterpret_cast<const float4*>(ep.gate + (size_t)(v == 1 ? b0 : b1) * ep.gate_stride + col0 + cc))
                        : make_float4(1.f, 1.f, 1.f, 1.f);
        } else if (v <= 4) {
          val = __ldg(reinterpret_cast<const float4*>(ln.scale + (size_t)(v == 3 ? b0 : b1) * ep.gate_stride + col0 + cc));
          val = make_float4(1.f + val.x, 1.f + val.y, 1.f + val.z, 1.f + val.w);
        } else {
          val = __ldg(reinterpret_cast<const float4*>(ln.shift + (size_t)(v == 5 ? b0 : b1) * ep.gate_stride + col0 + cc));
        }
        *reinterpret_cast<float4*>(cv + v * N_CTA + cc) = val;
      }
      named_bar_sync(5, 32 * EPI_WARPS);
    }
    const int sel = row_in_cta < split ? 0 : 1;
    const float* cbias = cv + chunk * 32;
    const float* cgate = cv + (1 + sel) * N_CTA + chunk * 32;
    const float* cmul = cv + (3 + sel) * N_CTA + chunk * 32;
    const float* cadd = cv + (5 + sel) * N_CTA + chunk * 32;
    // ---- x_new = x + gate * (acc + bias): 32 values per thread, kept in registers for the normalisation
    mbar_wait(tmem_full, 0);
    tcgen05_fence_after();
    uint32_t r[32];
    tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(chunk * 32), r);
    mbar_wait(&xbar[e], 0);
    float4 xr[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) xr[j] = *reinterpret_cast<const float4*>(myrow + ((j ^ swz) << 4));
    tmem_ld_wait();
    float o[32];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float4 bb = *reinterpret_cast<const float4*>(cbias + j * 4), gg = *reinterpret_cast<const float4*>(cgate + j * 4);
      o[4 * j] = fmaf(gg.x, __uint_as_float(r[4 * j]) + bb.x, xr[j].x);
      o[4 * j + 1] = fmaf(gg.y, __uint_as_float(r[4 * j + 1]) + bb.y, xr[j].y);
      o[4 * j + 2] = fmaf(gg.z, __uint_as_float(r[4 * j + 2]) + bb.z, xr[j].z);
      o[4 * j + 3] = fmaf(gg.w, __uint_as_float(r[4 * j + 3]) + bb.w, xr[j].w);
      *reinterpret_cast<float4*>(myrow + ((j ^ swz) << 4)) = make_float4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
    }
    fence_async_smem();
    __syncwarp();
    if (lane == 0) {
      tma_store_2d(&tmapX, tile_buf, gcol, grow);
      bulk_commit();
    }
    // ---- shifted statistics of these 32 columns, published to every CTA of the cluster
    const float shiftv = o[0];
    float s1 = 0.f, q1 = 0.f;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const float d = o[j] - shiftv;
      s1 += d;
      q1 = fmaf(d, d, q1);
    }
    {
      const uint32_t slot_addr = smem_u32(&xch[((int)rank * CHUNKS + chunk) * TILE_M + row_in_cta]);
      const float4 t = make_float4(s1, q1, shiftv, 0.f);
#pragma unroll
      for (uint32_t pr = 0; pr < (uint32_t)CL; ++pr) st_cluster_f4(map_to_rank(slot_addr, pr), t);
    }
    cluster_sync_all();       // release / acquire at cluster scope: all 12 triples of every row are in place
    float mean = 0.f;
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
      const float4 t = xch[s * TILE_M + row_in_cta];
      mean += fmaf(32.0f, t.z, t.x);
    }
    mean *= 1.0f / LN_N;
    float var = 0.f;
#pragma unroll
    for (int s = 0; s < SLOTS; ++s) {
      const float4 t = xch[s * TILE_M + row_in_cta];
      const float dm = mean - t.z;
      var += fmaf(32.0f * dm, dm, fmaf(-2.0f * dm, t.x, t.y));
    }
    const float rstd = rsqrtf(var * (1.0f / LN_N) + ln.eps);
    // ---- normalise + modulate from registers -> bf16 tile (32 columns = 64 B per row, SWIZZLE_64B) -> bulk store
    if (lane == 0) bulk_wait_read<0>();                     // the residual store has read the tile
    __syncwarp();
    uint8_t* hrow = tile_buf + lane * 64;
    const uint32_t swz64 = (uint32_t)((lane >> 1) & 3);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint32_t pk[4];
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {
        const float4 mu = *reinterpret_cast<const float4*>(cmul + j * 8 + h2 * 4), ad = *reinterpret_cast<const float4*>(cadd + j * 8 + h2 * 4);
        const float y0 = fmaf((o[8 * j + 4 * h2] - mean) * rstd, mu.x, ad.x), y1 = fmaf((o[8 * j + 4 * h2 + 1] - mean) * rstd, mu.y, ad.y);
        const float y2 = fmaf((o[8 * j + 4 * h2 + 2] - mean) * rstd, mu.z, ad.z), y3 = fmaf((o[8 * j + 4 * h2 + 3] - mean) * rstd, mu.w, ad.w);
        const __nv_bfloat162 lo = __floats2bfloat162_rn(y0, y1), hi = __floats2bfloat162_rn(y2, y3);
        pk[2 * h2] = *reinterpret_cast<const uint32_t*>(&lo);
        pk[2 * h2 + 1] = *reinterpret_cast<const uint32_t*>(&hi);
      }
      *reinterpret_cast<uint4*>(hrow + (((uint32_t)j ^ swz64) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
    }
    fence_async_smem();
    __syncwarp();
    if (lane == 0) {
      tma_store_2d(&tmapH, tile_buf, gcol, grow);
      bulk_commit();
      bulk_wait_read<0>();                                  // shared memory must outlive the bulk store's read
    }
    tcgen05_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(128) : "memory");
  }
}

}  // namespace tcln4
}  // namespace mdgen
