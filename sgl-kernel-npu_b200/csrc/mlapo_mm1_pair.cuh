// k2p MM1 on CTA pairs (tcgen05 cta_group::2) for decode calls of more than 128 tokens: a0 . Wdqkv^T -> split-K int32 partial sums.
//
// A pair of CTAs on neighbouring SMs owns one work item = (128-column weight tile, K-split, block of 256 tokens).  CTA r of the
// pair holds token tile r (128 rows of a0) and HALF of the weight tile's rows (64); one MMA (M = 256, N = 128) spans both SMs.
// Against the single-CTA form (mlapo_mm1.cuh: both token tiles and the whole weight tile in one CTA, 8 K-splits) a pair covers
// twice the K range with the same bytes per CTA (16 KB + 8 KB per k-block instead of 32 KB + 16 KB), so 4 K-splits fill the
// chip: each CTA writes 64 KB of partial sums instead of 128 KB and the norm kernel sums 4 slabs instead of 8.
// 10 warps per CTA: w0 TMA producer (own a0 tile + own half of the weight rows; completions are counted on the LEADER's
// barriers), w1 MMA issuer (leader CTA only; commits arrive in both CTAs) + TMEM alloc, w2..w9 two epilogue groups.
#pragma once
#include "mlapo_mm1.cuh"

namespace mlapo {

// BN = 128 (one block of 256 tokens: 17 tiles x 4 K-splits = 68 pairs) or 256 (several token blocks, where the K range of a
// pair is long and the a0 bytes per weight byte matter more than the partial sums: 9 tiles x 2 blocks x 4 K-splits at 512 tokens).
template <int BN>
struct Mm1Pair {
  static constexpr int HALF_B = BN / 2;                                  // weight rows per CTA
  static constexpr int STAGE_BYTES = kTile16K + HALF_B * kBlockKBytes;   // 24576 / 32768
  static constexpr int STAGES = BN == 128 ? 8 : 6;
  static constexpr int BAR_OFF = STAGES * STAGE_BYTES;                   // 196608 for both
  static constexpr int SMEM_BYTES = BAR_OFF + 256;
  static constexpr int GRANULES = BN / 32;                               // 32 int32 columns = one 128-byte row per token
  static_assert(STAGE_BYTES % 1024 == 0, "swizzled tiles need 1024-byte alignment");
  static_assert(GRANULES * kTile16K <= BAR_OFF, "the epilogue stages its granule tiles in the drained pipeline");
};

template <int BN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
mm1_pair_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                const __grid_constant__ CUtensorMap tm_p, int m_total, int kb_total, long long* tl) {
  if (threadIdx.x == 0) { tl_mark_global(tl, 0); tl_mark(tl, 1); }
  using P = Mm1Pair<BN>;
  constexpr int kM1PStages = P::STAGES, kM1PStageBytes = P::STAGE_BYTES, kM1PHalfB = P::HALF_B;
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
  uint64_t* const bars = reinterpret_cast<uint64_t*>(smem + P::BAR_OFF);
  uint64_t* const full_bar = bars;                    // [8] leader's copy used: TMA of both CTAs -> MMA
  uint64_t* const empty_bar = bars + kM1PStages;      // [8] MMA commit (both CTAs) -> TMA
  uint64_t* const done_bar = bars + 2 * kM1PStages;   // [1] MMA commit (both CTAs) -> epilogue
  uint32_t* const tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kM1PStages + 2);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const bool leader = rank == 0;
  const int n0 = (int)(blockIdx.x >> 1) * BN;
  const int split = blockIdx.y, splits = gridDim.y;
  const int m0 = blockIdx.z * (2 * kBlockM) + (int)rank * kBlockM;      // this CTA's token tile
  const int kb0 = (int)(((long long)split * kb_total) / splits);
  const int nkb = (int)(((long long)(split + 1) * kb_total) / splits) - kb0;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_b);
    ptx::prefetch_tmap(&tm_a);
    ptx::prefetch_tmap(&tm_p);
    for (int i = 0; i < 2 * kM1PStages + 1; ++i) ptx::mbar_init(ptx::smem_u32(bars + i), 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc_pair(ptx::smem_u32(tmem_slot), BN);
    ptx::tmem_relinquish_pair();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_arrive();  // the peer's barriers must exist before anything is sent to them
  ptx::cluster_wait();
  ptx::tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot;
  ptx::griddep_launch_dependents();
  if (threadIdx.x == 0) tl_mark(tl, 2);  // prologue done

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      constexpr uint32_t stage_tx = 2 * kM1PStageBytes;  // both CTAs' bytes land on the leader's barrier
      const int pre = nkb < kM1PStages ? nkb : kM1PStages;
      auto load_b = [&](int i, int s) {
        ptx::tma_load_2d_pair(ptx::smem_u32(smem + s * kM1PStageBytes + kTile16K), &tm_b, ptx::mapa(ptx::smem_u32(full_bar + s), 0),
                              (kb0 + i) * kBlockKBytes, n0 + (int)rank * kM1PHalfB, ptx::kEvictFirst);
      };
      auto load_a = [&](int i, int s) {
        ptx::tma_load_2d_pair(ptx::smem_u32(smem + s * kM1PStageBytes), &tm_a, ptx::mapa(ptx::smem_u32(full_bar + s), 0),
                              (kb0 + i) * kBlockKBytes, m0, ptx::kEvictLast);
      };
      for (int i = 0; i < pre; ++i) {  // the weight rows depend on nothing upstream
        if (leader) ptx::mbar_arrive_expect_tx(ptx::smem_u32(full_bar + i), stage_tx);
        load_b(i, i);
      }
      ptx::griddep_wait();  // a0 is produced by the quantise kernel
      tl_mark(tl, 3);
      for (int i = 0; i < pre; ++i) load_a(i, i);
      uint32_t stage = 0, phase = 1;
      for (int i = pre; i < nkb; ++i) {
        ptx::mbar_wait(ptx::smem_u32(empty_bar + stage), phase ^ 1);
        if (leader) ptx::mbar_arrive_expect_tx(ptx::smem_u32(full_bar + stage), stage_tx);
        load_b(i, stage);
        load_a(i, stage);
        if (++stage == kM1PStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA, one thread, for both SMs)
    if (leader) {
      constexpr uint32_t idesc = ptx::make_idesc(ptx::kCFmtS32, ptx::kFmtS8, ptx::kFmtS8, 0, 0, 2 * kBlockM, BN);
      uint32_t stage = 0, phase = 0;
      for (int i = 0; i < nkb; ++i) {
        ptx::mbar_wait(ptx::smem_u32(full_bar + stage), phase);
        __syncwarp();
        ptx::tc_fence_after();
        if (lane == 0) {
          if (i == 0) tl_mark(tl, 4);
          if (i == nkb - 1) tl_mark(tl, 5);
        }
        {  // the converged warp issues (one elected lane inside the asm blocks): see sm100_ptx.cuh, umma_i8_elect
          const uint32_t a_base = ptx::smem_u32(smem + stage * kM1PStageBytes);
          const uint64_t da0 = ptx::make_smem_desc(a_base, 1024, 0);
          const uint64_t db0 = ptx::make_smem_desc(a_base + kTile16K, 1024, 0);
#pragma unroll
          for (int k = 0; k < kBlockKBytes / kUmmaKBytes; ++k)
            ptx::umma_i8_pair_elect(tmem_acc, ptx::desc_advance(da0, k * kUmmaKBytes), ptx::desc_advance(db0, k * kUmmaKBytes), idesc,
                                    (i > 0 || k > 0) ? 1u : 0u);
          ptx::umma_commit_pair_elect(ptx::smem_u32(empty_bar + stage));
        }
        __syncwarp();
        if (++stage == kM1PStages) { stage = 0; phase ^= 1; }
      }
      ptx::umma_commit_pair_elect(ptx::smem_u32(done_bar));
      __syncwarp();
    }
  } else if (warp < 10) {
    // ------------------------------------------------------------------ epilogue: this CTA's 128 rows x BN int32 columns
    // Two groups of four warps, half of the 32-column granules each: TMEM -> swizzled smem tile -> its own TMA store.
    const int quarter = warp & 3, group = (warp - 2) >> 2;
    const int row = quarter * 32 + lane;
    const uint32_t t_base = tmem_acc + ((uint32_t)(quarter * 32) << 16);
    const bool issuer = threadIdx.x == 64 + group * 128;
    ptx::mbar_wait(ptx::smem_u32(done_bar), 0);   // all MMAs retired => every pipeline stage has been consumed
    __syncwarp();
    ptx::tc_fence_after();
    if (issuer && group == 0) tl_mark(tl, 6);
    __syncwarp();
#pragma unroll 1
    for (int j = 0; j < P::GRANULES / 2; ++j) {
      const int g = group * (P::GRANULES / 2) + j;
      uint32_t r[32];
      ptx::tmem_ld_32x32b_x32(t_base + g * 32, r);
      ptx::tmem_ld_wait();
      const uint32_t tile = ptx::smem_u32(smem + g * kTile16K);
#pragma unroll
      for (int q = 0; q < 8; ++q)
        ptx::st_shared_v4(tile + ptx::sw128_offset(row, q), r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3]);
      ptx::fence_proxy_async_smem();
      ptx::named_bar_sync(1 + group, 128);
      if (issuer) {
        if (n0 + g * 32 < MLAPO_MM1_OUT) ptx::tma_store_3d(&tm_p, tile, n0 + g * 32, m0, split);
        ptx::tma_store_commit();
      }
    }
    if (issuer) {
      if (group == 0) tl_mark(tl, 7);
      ptx::tma_store_wait_read<0>();
      if (group == 0) tl_mark(tl, 8);
    }
  }

  // nobody leaves while the peer may still address this CTA's barriers or tensor memory
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_arrive();
  ptx::cluster_wait();
  if (warp == 1) ptx::tmem_dealloc_pair(tmem_acc, BN);
  if (threadIdx.x == 32) { tl_mark(tl, 9); tl_mark_global(tl, 15); }
}

}  // namespace mlapo
