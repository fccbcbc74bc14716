// gemm_ln.cuh -- output projection fused with the residual update AND the next block's LayerNorm + adaLN modulate:
//   x[m,:]  += gate[b] * (A[m,:] . W^T + bias)                        (latent_model.py:387,392,486,491)
//   h[m,:]   = LN(x[m,:]) * (1 + scale'[b]) + shift'[b]   as bf16      (the NEXT block's latent_model.py:385,390,467,475,489)
// for the C = 384 wide residual stream (N = 384 = one full row per tile, K = 384 or 1536).
// Why: the stand-alone LayerNorm kernel re-reads the fp32 stream the projection has just written (49 MB in, 24.5 MB out,
// 14-17 us) and costs a kernel boundary; here the new row is still on chip when it is normalised.
// Structure: the CTA-pair machinery of gemm_tc2.cuh (cluster of 2, tcgen05 cta_group::2, M = 256 rows per pair), but a
// tile spans the whole row: two MMAs per k-step (N = 256 into TMEM columns 0..255, N = 128 into 256..383).  The 384
// accumulator columns leave no room for a second accumulator stage, so the epilogue does not overlap the next tile's
// MMAs -- the TMA producer keeps prefetching the next tile's operands meanwhile.
// Epilogue (warps 2..9, thread = row, two warps per TMEM lane quarter each owning 6 of the 12 32-column chunks):
//   pass 1  acc -> x_new = x + gate * (acc + bias) -> global x, x_new written back over the accumulator in TMEM, row sum
//   (the centred second moment comes from shifted sums accumulated in pass 1; round 1 re-read the row from TMEM for it)
//   pass 3  normalise, modulate, bf16 -> global h
// Global x / h traffic of the epilogue moves by TMA: per warp two SWIZZLE_128B tiles (residual tile in, x_new in place and
// the bf16 h tile out through bulk tensor stores) -- the LDG/STG path reached only ~25 B/clk/SM here.
// Row statistics of the two column halves are combined through shared memory (one named barrier per lane quarter).
#pragma once
#include "gemm_tc2.cuh"

namespace mdgen {
namespace tcln {

using namespace tc;
using tc2::cluster_ctarank;
using tc2::cluster_sync_all;
using tc2::map_to_rank;
using tc2::mbar_arrive_cluster;
using tc2::tcgen05_commit_pair;
using tc2::tma_load_2d_pair;
using tc2::umma_bf16_pair;
using tc2::stamp;

constexpr int LN_N = 384, TILE_M = 256;
constexpr int W1_ROWS = 128, W2_ROWS = 64;      // W rows staged per CTA for the N = 256 and the N = 128 MMA
constexpr int STAGE_BYTES = A_STAGE_BYTES + (W1_ROWS + W2_ROWS) * BLOCK_K * 2;   // 40 KB
constexpr int STAGES = 3;
constexpr int NVEC = 7;                          // bias, gate[2], 1 + scale[2], shift[2]   (two batch elements per slab)
constexpr int CV_BYTES = NVEC * LN_N * 4;
constexpr int STATS_BYTES = 2 * 2 * 128 * 4;     // [sum | centred square][column half][row] partial statistics
constexpr int EBUF_BYTES = 4096;                 // one epilogue I/O tile: 32 rows x 128 B (32 fp32 or 64 bf16 columns), SWIZZLE_128B
constexpr int OFF_EBUF = STAGES * STAGE_BYTES;   // [8 warps][2] tiles, 1024-byte aligned (TMA swizzle atom)
constexpr int OFF_CV = OFF_EBUF + NUM_EPI_WARPS * 2 * EBUF_BYTES;
constexpr int OFF_BAR = OFF_CV + CV_BYTES + STATS_BYTES;
constexpr int SMEM_BYTES = OFF_BAR + 512;
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");
static_assert(STAGE_BYTES % 1024 == 0, "SWIZZLE_128B tiles need 1024-byte aligned bases");

struct LnFuse {
  bf16* out;            // [M, 384] normalised + modulated rows for the next block
  const float* shift;   // next block's adaLN shift / scale rows, indexed like the gate: + b * gate_stride
  const float* scale;
  float eps;
};

__device__ __forceinline__ void tmem_st_32x32b_x32(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
      "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]),
      "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]),
      "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// The residual + LayerNorm epilogue of one 128-row x 384-column accumulator (TMEM columns 0..383 of this CTA), shared by
// gemm_ln_kernel and gemm_mlp_ln_kernel (gemm_mlp.cuh).  One instance per epilogue thread (warps 2..9, thread = row, two
// warps per TMEM lane quarter each owning 6 of the 12 32-column chunks).
//   All arithmetic happens in the TMEM row layout (per-row statistics need no shuffles, x_new can live in TMEM between
//   the passes); every global access goes through TMA: per warp two 32 x 128 B tiles in the SWIZZLE_128B layout, which a
//   thread = row reads / writes conflict-free (16-byte chunk j of row r sits at chunk j ^ (r & 7)).  Row-per-thread global
//   accesses were measured 10x slower (no L1 left to merge the 16-byte pieces of a line with the shared-memory carve-out at
//   its maximum), and the LSU path (LDG/STG + shared-memory transposition) moved only ~25 B/clk/SM.
struct LnEpilogue {
  static constexpr int CH = 6;                             // 32-column chunks per warp
  int lane, quarter, half, et, row_in_cta, col0;
  uint32_t trow, te, swz;
  uint8_t *ebuf, *myrow0;
  uint64_t *xb, *tmem_full;
  float *cv, *stats, *stats2;
  bool st0;
  int staged_q;
  uint32_t acc_phase, xph0, xph1;

  // ebuf_all: [8 warps][2] tiles of EBUF_BYTES, 1024-byte aligned; cv_all: [NVEC][384] + STATS_BYTES; xbar: [8 warps][2]
  __device__ __forceinline__ void init(int warp, int lane_, uint32_t tmem_base, uint8_t* ebuf_all, float* cv_all, uint64_t* xbar,
                                       uint64_t* tmem_full_, uint64_t* tmem_empty) {
    lane = lane_;
    quarter = warp & 3;
    half = (warp - 2) >> 2;
    et = (warp - 2) * 32 + lane;                           // 0..255 among the epilogue threads
    row_in_cta = quarter * 32 + lane;
    trow = tmem_base + ((uint32_t)(quarter * 32) << 16);
    te = map_to_rank(smem_u32(tmem_empty), 0);
    ebuf = ebuf_all + (warp - 2) * 2 * EBUF_BYTES;
    xb = xbar + (warp - 2) * 2;
    swz = (uint32_t)(lane & 7);
    myrow0 = ebuf + lane * 128;
    col0 = half * CH * 32;
    cv = cv_all;
    stats = cv + NVEC * LN_N;                              // row sums [2][128]
    stats2 = stats + 2 * 128;                              // centred squares [2][128]
    tmem_full = tmem_full_;
    st0 = (warp == 2 && lane == 0);
    staged_q = -1;
    acc_phase = xph0 = xph1 = 0;
  }

  // the first two residual tiles of this slab start moving (the previous tile's stores have left the buffers)
  __device__ __forceinline__ void prefetch(const CUtensorMap& tmapX, int cta_row0) {
    const int slab_row0 = cta_row0 + quarter * 32;
    if (lane == 0) {
      bulk_wait_read<0>();
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        mbar_expect_tx(&xb[u], EBUF_BYTES);
        tma_load_2d(&tmapX, &xb[u], ebuf + u * EBUF_BYTES, col0 + u * 32, slab_row0);
      }
    }
    __syncwarp();
  }

  // waits for the accumulator, runs the three passes, hands the accumulator back (arrival on `te`)
  __device__ __forceinline__ void tile(const CUtensorMap& tmapX, const CUtensorMap& tmapH, const EpiParams& ep, const LnFuse& ln,
                                       int cta_row0, int ti) {
    const int slab_row0 = cta_row0 + quarter * 32;
    // ---- per-column vectors of the (at most two) batch elements this CTA's 128 rows touch
    const int q0 = cta_row0 / ep.rows_per_batch;
    const int split = (q0 + 1) * ep.rows_per_batch - cta_row0;        // rows >= split belong to the next batch element
    if (q0 != staged_q) {
      named_bar_sync(5, 32 * NUM_EPI_WARPS);                          // everyone is done with the previous vectors
      const int b0 = q0 % ep.nbatch, b1 = (q0 + 1) % ep.nbatch;
      for (int i = et; i < NVEC * (LN_N / 4); i += 32 * NUM_EPI_WARPS) {
        const int v = i / (LN_N / 4), cc = (i % (LN_N / 4)) * 4;
        float4 val;
        if (v == 0) {
          val = ep.bias ? __ldg(reinterpret_cast<const float4*>(ep.bias + cc)) : make_float4(0.f, 0.f, 0.f, 0.f);
        } else if (v <= 2) {
          val = ep.gate ? __ldg(reinterpret_cast<const float4*>(ep.gate + (size_t)(v == 1 ? b0 : b1) * ep.gate_stride + cc))
                        : make_float4(1.f, 1.f, 1.f, 1.f);
        } else if (v <= 4) {
          val = __ldg(reinterpret_cast<const float4*>(ln.scale + (size_t)(v == 3 ? b0 : b1) * ep.gate_stride + cc));
          val = make_float4(1.f + val.x, 1.f + val.y, 1.f + val.z, 1.f + val.w);
        } else {
          val = __ldg(reinterpret_cast<const float4*>(ln.shift + (size_t)(v == 5 ? b0 : b1) * ep.gate_stride + cc));
        }
        *reinterpret_cast<float4*>(cv + v * LN_N + cc) = val;
      }
      named_bar_sync(5, 32 * NUM_EPI_WARPS);
      staged_q = q0;
    }
    const int sel = row_in_cta < split ? 0 : 1;
    const float* cbias = cv + col0;
    const float* cgate = cv + (1 + sel) * LN_N + col0;
    const float* cmul = cv + (3 + sel) * LN_N + col0;
    const float* cadd = cv + (5 + sel) * LN_N + col0;
    if (st0) stamp(ep, ti, 2);
    mbar_wait(tmem_full, acc_phase);
    tcgen05_fence_after();
    if (st0) stamp(ep, ti, 3);
    // ---- pass 1: residual update, x_new back into TMEM and (in place in the tile) out to global; row statistics in the
    // same pass as SHIFTED sums (shift = the row's first new value in this column half: s1 = sum(x - shift), q1 = sum((x -
    // shift)^2), so that the centred second moment q1 - 2 (mean - shift) s1 + n (mean - shift)^2 loses no digits to a
    // large row mean).  A separate pass over TMEM for the centred squares cost 1300 of the epilogue's 15400 clocks.
    float s1 = 0.f, q1 = 0.f, shiftv = 0.f;
#pragma unroll 1
    for (int c2 = 0; c2 < CH; c2 += 2) {
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int c = c2 + u;
        uint8_t* row = myrow0 + u * EBUF_BYTES;
        uint32_t r[32];
        tmem_ld_32x32b_x32(trow + col0 + c * 32, r);
        float4 bb[8], gg[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          bb[j] = *reinterpret_cast<const float4*>(cbias + c * 32 + j * 4);
          gg[j] = *reinterpret_cast<const float4*>(cgate + c * 32 + j * 4);
        }
        mbar_wait(&xb[u], u ? xph1 : xph0);               // this chunk's residual tile has landed
        if (u) xph1 ^= 1; else xph0 ^= 1;
        float4 xr[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) xr[j] = *reinterpret_cast<const float4*>(row + ((j ^ swz) << 4));
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float o0 = fmaf(gg[j].x, __uint_as_float(r[4 * j]) + bb[j].x, xr[j].x);
          const float o1 = fmaf(gg[j].y, __uint_as_float(r[4 * j + 1]) + bb[j].y, xr[j].y);
          const float o2 = fmaf(gg[j].z, __uint_as_float(r[4 * j + 2]) + bb[j].z, xr[j].z);
          const float o3 = fmaf(gg[j].w, __uint_as_float(r[4 * j + 3]) + bb[j].w, xr[j].w);
          if (u == 0 && j == 0) shiftv = c2 == 0 ? o0 : shiftv;
          const float d0 = o0 - shiftv, d1 = o1 - shiftv, d2 = o2 - shiftv, d3 = o3 - shiftv;
          s1 += (d0 + d1) + (d2 + d3);
          q1 = fmaf(d0, d0, q1);
          q1 = fmaf(d1, d1, q1);
          q1 = fmaf(d2, d2, q1);
          q1 = fmaf(d3, d3, q1);
          r[4 * j] = __float_as_uint(o0);
          r[4 * j + 1] = __float_as_uint(o1);
          r[4 * j + 2] = __float_as_uint(o2);
          r[4 * j + 3] = __float_as_uint(o3);
          *reinterpret_cast<uint4*>(row + ((j ^ swz) << 4)) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
        }
        tmem_st_32x32b_x32(trow + col0 + c * 32, r);
        fence_async_smem();                               // this thread's row is visible to the bulk-copy engine
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmapX, ebuf + u * EBUF_BYTES, col0 + c * 32, slab_row0);
          bulk_commit();
          if (c + 2 < CH) {                               // refill the tile as soon as the store has read it
            bulk_wait_read<0>();
            mbar_expect_tx(&xb[u], EBUF_BYTES);
            tma_load_2d(&tmapX, &xb[u], ebuf + u * EBUF_BYTES, col0 + (c + 2) * 32, slab_row0);
          }
        }
        __syncwarp();
      }
    }
    tmem_st_wait();
    if (st0) stamp(ep, ti, 4);
    constexpr float NH = (float)(CH * 32);                 // columns of this half
    const float sum = fmaf(NH, shiftv, s1);
    stats[half * 128 + row_in_cta] = sum;
    named_bar_sync(1 + quarter, 64);
    const float mean = (sum + stats[(half ^ 1) * 128 + row_in_cta]) * (1.0f / LN_N);
    named_bar_sync(1 + quarter, 64);                      // both partial sums consumed before the slots are reused
    // ---- centred second moment of this half from the shifted sums
    const float dm = mean - shiftv;
    const float sq = fmaf(NH * dm, dm, fmaf(-2.0f * dm, s1, q1));
    if (st0) stamp(ep, ti, 5);
    stats2[half * 128 + row_in_cta] = sq;
    named_bar_sync(1 + quarter, 64);
    const float rstd = rsqrtf((sq + stats2[(half ^ 1) * 128 + row_in_cta]) * (1.0f / LN_N) + ln.eps);
    // ---- pass 3: normalise + modulate -> bf16 tile (64 columns = 128 B per row) -> bulk store
    auto pack_h = [&](int c, int pbase, uint8_t* row, const uint32_t (&r)[32]) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint32_t pk[4];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const float4 mu = *reinterpret_cast<const float4*>(cmul + c * 32 + j * 8 + e * 4);
          const float4 ad = *reinterpret_cast<const float4*>(cadd + c * 32 + j * 8 + e * 4);
          const float y0 = fmaf((__uint_as_float(r[8 * j + 4 * e]) - mean) * rstd, mu.x, ad.x);
          const float y1 = fmaf((__uint_as_float(r[8 * j + 4 * e + 1]) - mean) * rstd, mu.y, ad.y);
          const float y2 = fmaf((__uint_as_float(r[8 * j + 4 * e + 2]) - mean) * rstd, mu.z, ad.z);
          const float y3 = fmaf((__uint_as_float(r[8 * j + 4 * e + 3]) - mean) * rstd, mu.w, ad.w);
          const __nv_bfloat162 lo = __floats2bfloat162_rn(y0, y1), hi = __floats2bfloat162_rn(y2, y3);
          pk[2 * e] = *reinterpret_cast<const uint32_t*>(&lo);
          pk[2 * e + 1] = *reinterpret_cast<const uint32_t*>(&hi);
        }
        *reinterpret_cast<uint4*>(row + (((uint32_t)(pbase + j) ^ swz) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      }
    };
    {
      uint32_t ra[32], rb[32];
      tmem_ld_32x32b_x32(trow + col0, ra);
      tmem_ld_32x32b_x32(trow + col0 + 32, rb);
#pragma unroll 1
      for (int c = 0; c < CH; c += 2) {
        const int u = (c >> 1) & 1;
        uint8_t* row = myrow0 + u * EBUF_BYTES;
        if (lane == 0) {                                    // the tile's previous store (pass 1 / pair c - 4) has been read out
          if (c == 0) bulk_wait_read<0>(); else bulk_wait_read<1>();
        }
        __syncwarp();
        tmem_ld_wait();
        pack_h(c, 0, row, ra);
        pack_h(c + 1, 4, row, rb);
        if (c + 2 < CH) {
          tmem_ld_32x32b_x32(trow + col0 + (c + 2) * 32, ra);
          tmem_ld_32x32b_x32(trow + col0 + (c + 3) * 32, rb);
        } else {                                            // every accumulator column has been read: free the stage
          tcgen05_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(te);
        }
        fence_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmapH, ebuf + u * EBUF_BYTES, col0 + c * 32, slab_row0);
          bulk_commit();
        }
      }
    }
    if (st0) stamp(ep, ti, 6);
    acc_phase ^= 1;
  }
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
gemm_ln_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapW1,
               const __grid_constant__ CUtensorMap tmapW2, const __grid_constant__ CUtensorMap tmapX,
               const __grid_constant__ CUtensorMap tmapH, int M, int K, EpiParams ep, LnFuse ln) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  float* cv = reinterpret_cast<float*>(smem + OFF_CV);                        // [NVEC][384]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 1;
  uint64_t* xbar = tmem_empty + 1;                                            // [8 warps][2] residual-tile arrival barriers
  uint32_t* tmem_base_slot = reinterpret_cast<uint32_t*>(xbar + 2 * NUM_EPI_WARPS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  const int num_tiles = (M + TILE_M - 1) / TILE_M;
  const int num_kb = (K + BLOCK_K - 1) / BLOCK_K;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapW1) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapW2) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapX) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmapH) : "memory");
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
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
  pdl_wait();          // the previous kernel's outputs (A, the residual stream) are complete from here on

  if (warp == 0) {
    // ===== TMA producer: own 128 rows of A, own halves of the two W slices =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        const int arow = tile * TILE_M + (int)rank * BLOCK_M;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);
          const uint32_t fb = map_to_rank(smem_u32(&full_bar[stage]), 0);
          tma_load_2d_pair(&tmapA, fb, sa, kb * BLOCK_K, arow);
          tma_load_2d_pair(&tmapW1, fb, sa + A_STAGE_BYTES, kb * BLOCK_K, (int)rank * W1_ROWS);
          tma_load_2d_pair(&tmapW2, fb, sa + A_STAGE_BYTES + W1_ROWS * BLOCK_K * 2, kb * BLOCK_K, 256 + (int)rank * W2_ROWS);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA, warp-uniform, one elected lane issues): per k-step one N = 256 and one N = 128 MMA =====
    if (rank == 0) {
      constexpr uint32_t idesc1 = make_idesc_bf16(TILE_M, 256), idesc2 = make_idesc_bf16(TILE_M, 128);
      int stage = 0;
      uint32_t phase = 0, acc_phase = 0;
      int ti = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++ti) {
        mbar_wait(tmem_empty, acc_phase ^ 1);
        tcgen05_fence_after();
        if (lane == 0) stamp(ep, ti, 0);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tcgen05_fence_after();
          const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
          const uint64_t adesc = make_smem_desc_sw128(sa);
          const uint64_t b1desc = make_smem_desc_sw128(sa + A_STAGE_BYTES);
          const uint64_t b2desc = make_smem_desc_sw128(sa + A_STAGE_BYTES + W1_ROWS * BLOCK_K * 2);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
              const uint32_t accum = (kb | k) != 0 ? 1u : 0u;
              umma_bf16_pair(tmem_base, adesc + (uint64_t)(k * 2), b1desc + (uint64_t)(k * 2), idesc1, accum);
              umma_bf16_pair(tmem_base + 256, adesc + (uint64_t)(k * 2), b2desc + (uint64_t)(k * 2), idesc2, accum);
            }
            tcgen05_commit_pair(&empty_bar[stage]);
            if (kb == num_kb - 1) tcgen05_commit_pair(tmem_full);
          }
          __syncwarp();
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (lane == 0) stamp(ep, ti, 1);
        acc_phase ^= 1;
      }
    }
  } else {
    // ===== epilogue warps 2..9 =====
    LnEpilogue epi;
    epi.init(warp, lane, tmem_base, smem + OFF_EBUF, cv, xbar, tmem_full, tmem_empty);
    int ti = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters, ++ti) {
      const int cta_row0 = tile * TILE_M + (int)rank * BLOCK_M;
      epi.prefetch(tmapX, cta_row0);
      epi.tile(tmapX, tmapH, ep, ln, cta_row0, ti);
    }
    if (lane == 0) bulk_wait_read<0>();                       // shared memory must outlive the last bulk stores' reads
  }
  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tcgen05_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

}  // namespace tcln
}  // namespace mdgen
