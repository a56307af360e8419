// k7: the q path on CTA pairs (tcgen05 cta_group::2) — same arithmetic as k4 (mlapo_q.cuh), different ownership of the data.
//
// A pair of CTAs on neighbouring SMs owns one work unit = (head, block of 256 tokens) at a time; CTA r of the pair holds
// token tile r (128 rows).  One MMA instruction spans both SMs, so per k-block each CTA ingests only ITS a1 tile (16 KB) and
// HALF of the head's W_UQ rows (12 KB) — k4 ingests 2 x 16 KB + 24 KB per CTA for the same 256 tokens — and each CTA needs
// only 192 TMEM columns per unit.  That leaves room for a second accumulator set, so a pair walks a list of units and
// MM2 of unit i+1 streams in while unit i is dequantised, absorbed and stored:
//   MM2    acc2 = a1 . Wuq[head]^T        kind::i8, M=256 (2x128), N=192, TMEM region (i&1)*256 + [0,192)
//   epi2   dequant set (warps 4..11, two groups of 128 threads, thread = token row): group g takes nope columns
//          [64g,64g+64) -> swizzled smem (absorb A operand, k-block g) and rope dims [16g,16g+16) with their partners +32
//          -> RoPE -> q_out1
//   BMM3   acc3 = q_nope . Wuk[head][:, chunk c]   kind::f16, M=256, N=128 (each CTA holds 64 of the 128 columns),
//          buffer c&1 of the same TMEM region: [0,128) (the nope columns, already drained) / [128,256)
//   epi3   drain set (warps 12..19, two groups): act out — group g drains columns [64g,64g+64) of every chunk through its
//          16 KB staging tile -> TMA store; int8 out — group g drains whole chunks c = g, g+2 (128 int8 = one 128-byte row).
//          The two sets work on different units at the same time (dequant of unit i+1 under the drains of unit i).
// Warps: 0 TMA a1+W_UQ | 1 MM2 issuer (leader CTA), TMEM alloc | 2 TMA W_UK | 3 absorb issuer (leader CTA) | 4..19 epilogue.
// Barriers live at the same offset in both CTAs.  TMA completions of both CTAs are counted on the LEADER's full
// barriers; tcgen05.commit multicasts to both CTAs; epilogue -> MMA signals are remote arrives on the leader's copies.
#pragma once
#include "mlapo_q.cuh"

namespace mlapo {

constexpr int kMagic = 0x4B400000;      // bits of 1.5 * 2^23: for an int v with |v| < 2^22, float(v) == as_float(kMagic + v) - kMagicF
constexpr float kMagicF = 12582912.f;
constexpr int kPThreads = 640;
constexpr int kPStages = 4;
constexpr int kPHalfB = MLAPO_Q_HEAD / 2;                            // 96 W_UQ rows per CTA
constexpr int kPStageBytes = kTile16K + kPHalfB * kBlockKBytes;      // 28672
constexpr int kPNopeOff = kPStages * kPStageBytes;                   // 114688: 2 x 16 KB (k-block 0 / 1 of this CTA's token tile)
constexpr int kPWukSlots = 3;
constexpr int kPWukOff = kPNopeOff + 2 * kTile16K;                   // 147456: 3 x 16 KB ([128 k x 64 n] per slot)
constexpr int kPStageOff = kPWukOff + kPWukSlots * kTile16K;         // 196608: 2 x 16 KB output staging (one per drain group)
constexpr int kPBarOff = kPStageOff + 2 * kTile16K;                  // 229376
constexpr int kPCstPerGroup = 96;
constexpr int kPSmemBytes = kPBarOff + 256 + 2 * kPCstPerGroup * 8 + 1024;  // 232192
static_assert(kPStageBytes % 1024 == 0 && kPNopeOff % 1024 == 0, "swizzled tiles need 1024-byte alignment");
static_assert(kPSmemBytes <= 232448, "q_pair smem");

template <class T, bool kInt8Out>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPThreads, 1)
q_pair_kernel(const __grid_constant__ CUtensorMap tm_a1, const __grid_constant__ CUtensorMap tm_wuq,
              const __grid_constant__ CUtensorMap tm_wuk, const __grid_constant__ CUtensorMap tm_out0, QArgs qa, void* q_out1,
              int total_units, long long* tl) {
  if (threadIdx.x == 0) { tl_mark_global(tl, 0); tl_mark(tl, 1); }
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kPBarOff);
  uint64_t* full_bar = bars;            // [4] leader's copy used: TMA of both CTAs -> MM2
  uint64_t* empty_bar = bars + 4;       // [4] MM2 commit (both CTAs) -> TMA
  uint64_t* mm2_done = bars + 8;        // [2] per TMEM region: MM2 commit (both) -> epilogue
  uint64_t* region_free = bars + 10;    // [2] leader's: remote arrives after the last two chunks -> MM2 of the unit after next
  uint64_t* acc_read = bars + 12;       // [1] leader's: 4 remote arrives (CTA x dequant group): acc2 read, q_nope tiles written -> absorb
  uint64_t* qn_free = bars + 13;        // [1] absorb commit (both) -> epilogue may overwrite the q_nope tiles
  uint64_t* wuk_full = bars + 14;       // [3] leader's: TMA of both CTAs -> absorb
  uint64_t* wuk_empty = bars + 17;      // [3] absorb commit (both) -> TMA
  // unit_full is indexed by TMEM region AND buffer.  One barrier per buffer for both regions deadlocked (once in ~20 000
  // launches): the commits of unit i+1's first chunks need nothing from unit i's drains, so a drain group that was late for
  // chunk 2 / 3 of unit i could find its barrier two completions on — the parity it waits for is then the current phase's.
  // With one barrier per (region, buffer) the next completion after a unit's two belongs to unit i+2, whose MM2 waits for
  // region_free, i.e. for every drain group to be past these waits.
  uint64_t* unit_full = bars + 20;      // [4] (region, buffer): absorb commit (both) -> drains
  uint64_t* buf_empty = bars + 24;      // [2] leader's: remote arrives after chunks 0 / 1 -> absorb chunk c+2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 26);
  int2* s_cst_all = reinterpret_cast<int2*>(smem + kPBarOff + 256);
  // arrivals per completion: act out — both drain groups take part in every chunk; int8 out — group g alone drains chunks g, g+2
  constexpr uint32_t kBufEmptyCount = kInt8Out ? 2u : 4u;
  constexpr uint32_t kRegionFreeCount = kInt8Out ? 4u : 8u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int n_tb = (qa.m_total + 255) / 256;  // token blocks; units are head-major: u = head * n_tb + token block

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_wuq);
    ptx::prefetch_tmap(&tm_a1);
    ptx::prefetch_tmap(&tm_wuk);
    ptx::prefetch_tmap(&tm_out0);
    for (int i = 0; i < 26; ++i) {
      const uint32_t cnt = (i == 10 || i == 11) ? kRegionFreeCount : i == 12 ? 4u : (i == 24 || i == 25) ? kBufEmptyCount : 1u;
      ptx::mbar_init(ptx::smem_u32(bars + i), cnt);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc_pair(ptx::smem_u32(tmem_slot), 512);
    ptx::tmem_relinquish_pair();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_arrive();  // the peer's barriers must exist before anything is sent to them
  ptx::cluster_wait();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  if (threadIdx.x == 0) tl_mark(tl, 2);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA: own a1 tile + own half of W_UQ per k-block
    if (lane == 0) {
      constexpr uint32_t stage_tx = 2 * kPStageBytes;  // both CTAs' bytes land on the leader's barrier
      uint32_t gk = 0;
      for (int u = pair; u < total_units; u += num_pairs) {
        const int head = u / n_tb, tok0 = (u % n_tb) * 256 + (int)rank * kBlockM;
        for (int kb = 0; kb < kQKBlocks; ++kb, ++gk) {
          const uint32_t s = gk % kPStages;
          if (gk >= kPStages) ptx::mbar_wait(ptx::smem_u32(empty_bar + s), ((gk / kPStages) & 1) ^ 1);
          if (leader) ptx::mbar_arrive_expect_tx(ptx::smem_u32(full_bar + s), stage_tx);
          const uint32_t fb = ptx::mapa(ptx::smem_u32(full_bar + s), 0);
          const uint32_t dst = ptx::smem_u32(smem + s * kPStageBytes);
          ptx::tma_load_2d_pair(dst + kTile16K, &tm_wuq, fb, kb * kBlockKBytes, head * MLAPO_Q_HEAD + (int)rank * kPHalfB,
                                ptx::kEvictLast);
          if (gk == kPStages - 1) {
            // the first W_UQ k-blocks depend on nothing; a1 comes from the norm kernel
            ptx::griddep_wait();
            tl_mark(tl, 3);
            for (int j = 0; j < kPStages; ++j)
              ptx::tma_load_2d_pair(ptx::smem_u32(smem + j * kPStageBytes), &tm_a1, ptx::mapa(ptx::smem_u32(full_bar + j), 0),
                                    j * kBlockKBytes, tok0, ptx::kEvictLast);
          } else if (gk >= kPStages) {
            ptx::tma_load_2d_pair(dst, &tm_a1, fb, kb * kBlockKBytes, tok0, ptx::kEvictLast);
          }
        }
      }
    }
  } else if (warp == 2) {
    // ------------------------------------------------------------------ TMA: own 64 columns of each 128-column W_UK chunk
    if (lane == 0) {
      uint32_t gc = 0;
      for (int u = pair; u < total_units; u += num_pairs) {
        const int head = u / n_tb;
        for (int c = 0; c < 4; ++c, ++gc) {
          const uint32_t w = gc % kPWukSlots, use = gc / kPWukSlots;
          if (use >= 1) ptx::mbar_wait(ptx::smem_u32(wuk_empty + w), (use & 1) ^ 1);
          if (leader) ptx::mbar_arrive_expect_tx(ptx::smem_u32(wuk_full + w), 2 * kTile16K);
          ptx::tma_load_2d_pair(ptx::smem_u32(smem + kPWukOff + w * kTile16K), &tm_wuk, ptx::mapa(ptx::smem_u32(wuk_full + w), 0),
                                c * 128 + (int)rank * 64, head * MLAPO_Q_NOPE, ptx::kEvictLast);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA: MM2 (leader issues for both SMs)
    if (leader) {
      constexpr uint32_t idesc2 = ptx::make_idesc(ptx::kCFmtS32, ptx::kFmtS8, ptx::kFmtS8, 0, 0, 2 * kBlockM, MLAPO_Q_HEAD);
      uint32_t gk = 0, it = 0;
      for (int u = pair; u < total_units; u += num_pairs, ++it) {
        const uint32_t par = it & 1;
        if (it >= 2) ptx::mbar_wait_cluster(ptx::smem_u32(region_free + par), ((it >> 1) - 1) & 1);
        for (int kb = 0; kb < kQKBlocks; ++kb, ++gk) {
          const uint32_t s = gk % kPStages;
          ptx::mbar_wait(ptx::smem_u32(full_bar + s), (gk / kPStages) & 1);
          __syncwarp();
          ptx::tc_fence_after();
          if (lane == 0 && it == 0 && kb == 0) tl_mark(tl, 4);
          {  // the converged warp issues (one elected lane inside the asm blocks): see sm100_ptx.cuh, umma_i8_elect
            const uint32_t a_base = ptx::smem_u32(smem + s * kPStageBytes);
            const uint64_t da0 = ptx::make_smem_desc(a_base, 1024, 0);
            const uint64_t db0 = ptx::make_smem_desc(a_base + kTile16K, 1024, 0);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma_i8_pair_elect(tmem + par * 256, ptx::desc_advance(da0, k * kUmmaKBytes), ptx::desc_advance(db0, k * kUmmaKBytes), idesc2,
                                      (kb > 0 || k > 0) ? 1u : 0u);
            ptx::umma_commit_pair_elect(ptx::smem_u32(empty_bar + s));
          }
          __syncwarp();
        }
        ptx::umma_commit_pair_elect(ptx::smem_u32(mm2_done + par));
        __syncwarp();
      }
    }
  } else if (warp == 3) {
    // ------------------------------------------------------------------ MMA: absorb (leader issues for both SMs)
    if (leader) {
      constexpr uint32_t kAFmt = std::is_same<T, __nv_bfloat16>::value ? ptx::kFmtBF16 : ptx::kFmtF16;
      constexpr uint32_t idesc3 = ptx::make_idesc(ptx::kCFmtF32, kAFmt, kAFmt, 0, 1, 2 * kBlockM, 128);
      uint32_t gc = 0, it = 0;
      for (int u = pair; u < total_units; u += num_pairs, ++it) {
        const uint32_t par = it & 1;
        ptx::mbar_wait_cluster(ptx::smem_u32(acc_read), it & 1);
        for (int c = 0; c < 4; ++c, ++gc) {
          const uint32_t w = gc % kPWukSlots;
          if (c >= 2) ptx::mbar_wait_cluster(ptx::smem_u32(buf_empty + (c & 1)), it & 1);
          ptx::mbar_wait(ptx::smem_u32(wuk_full + w), (gc / kPWukSlots) & 1);
          __syncwarp();
          ptx::tc_fence_after();
          {
            const uint64_t da0 = ptx::make_smem_desc(ptx::smem_u32(smem + kPNopeOff), 1024, 0);
            const uint64_t db0 = ptx::make_smem_desc(ptx::smem_u32(smem + kPWukOff + w * kTile16K), 1024, 0);  // MN-major, one 64-column atom per CTA
#pragma unroll
            for (int k = 0; k < 8; ++k)  // 8 x UMMA_K(16) = 128
              ptx::umma_f16_pair_elect(tmem + par * 256 + (c & 1) * 128, ptx::desc_advance(da0, (k >> 2) * kTile16K + (k & 3) * kUmmaKBytes),
                                       ptx::desc_advance(db0, k * 2048), idesc3, k > 0 ? 1u : 0u);
            ptx::umma_commit_pair_elect(ptx::smem_u32(wuk_empty + w));
            ptx::umma_commit_pair_elect(ptx::smem_u32(unit_full + par * 2 + (c & 1)));
          }
          __syncwarp();
        }
        ptx::umma_commit_pair_elect(ptx::smem_u32(qn_free));  // every MMA that reads the q_nope tiles has retired
        __syncwarp();
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps: thread = token row
    // Two warp sets work on DIFFERENT units at the same time: warps 4..11 dequantise unit i+1 (acc2 -> q_nope tile, RoPE)
    // while warps 12..19 drain the absorb results of unit i.  Each set is two groups of 128 threads that split columns.
    const int eg = (warp - 4) >> 2;        // 0,1: dequant groups; 2,3: drain groups
    const int g = eg & 1;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int et = threadIdx.x - 128 - eg * 128;   // 0..127 inside the group
    const bool issuer = et == 0;
    const int g_bar = 1 + eg;
    const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
    const bool per_token = qa.tok_scale1 != nullptr;
    ptx::griddep_wait();  // per-token scales come from the norm kernel
    ptx::griddep_launch_dependents();  // only after the wait: see q_fused

    if (eg < 2) {
      // ---------------- dequant set.  Group g: nope columns [64g, 64g+64) -> q_nope tile g (absorb A operand, k-block g);
      //                  rope dims [16g, 16g+16) and their partners 32 + the same -> RoPE -> q_out1.
      const uint32_t r_acc_read = ptx::mapa(ptx::smem_u32(acc_read), 0);
      int2* const s_cst = s_cst_all + g * kPCstPerGroup;  // [0,64) nope cols 64g.., [64,80) rope dims 16g.., [80,96) 32+16g..
      // Fetched one unit ahead: the column constants {bias + kMagic, descale} of the group's 96 columns and the token scale.
      float ts = 1.f;
      int bias_next = 0;
      float dsc_next = 0.f;
      auto fetch_unit = [&](int u) {
        const int head = u / n_tb, t = (u % n_tb) * 256 + (int)rank * kBlockM + row;
        const int tc = t < qa.m_total ? t : qa.m_total - 1;  // clamped row for loads
        if (et < kPCstPerGroup) {
          const int col = et < 64 ? 64 * g + et : et < 80 ? 128 + 16 * g + (et - 64) : 160 + 16 * g + (et - 80);
          dsc_next = load_descale(qa.descale1, qa.descale_is_i64, head * MLAPO_Q_HEAD + col);  // consumed when the unit starts
          bias_next = qa.bias1 ? qa.bias1[head * MLAPO_Q_HEAD + col] : 0;
        }
        if (per_token) ts = qa.tok_scale1[tc];
      };
      if (pair < total_units) fetch_unit(pair);

      uint32_t it = 0;
      for (int u = pair; u < total_units; u += num_pairs, ++it) {
        const uint32_t par = it & 1;
        const int head = u / n_tb, tok0 = (u % n_tb) * 256 + (int)rank * kBlockM;
        const int t = tok0 + row;
        const bool valid = t < qa.m_total;
        if (et < kPCstPerGroup) s_cst[et] = make_int2(bias_next + kMagic, __float_as_int(dsc_next));  // fetched one unit ago
        const float ts_u = ts;  // 1 in per-tensor mode
        // this row's cos/sin (dims [16g,16g+16) and 32 + the same), used after the hand-off to the absorb
        const T* cos_row = reinterpret_cast<const T*>(qa.cos) + (long long)(valid ? t : qa.m_total - 1) * 64 + 16 * g;
        const T* sin_row = reinterpret_cast<const T*>(qa.sin) + (long long)(valid ? t : qa.m_total - 1) * 64 + 16 * g;
        ptx::named_bar_sync(g_bar, 128);  // constants visible to the group

        ptx::mbar_wait(ptx::smem_u32(mm2_done + par), (it >> 1) & 1);
        __syncwarp();
        ptx::tc_fence_after();
        if (issuer && g == 0 && it < 3) tl_mark(tl, 5 + 3 * it);  // MM2 of unit `it` retired (as seen by the dequant set)
        const bool tr = issuer && g == 0 && it >= 2 && it < 4;
        const int tb = (it - 2) * 32;
        tl_trace(tl, tr, tb + 0);
        const uint32_t acc2 = lane_base + par * 256;

        // int32 -> fp32 without the quarter-rate I2F: exact while |acc + bias| < 2^22; otherwise (never seen with int8 data,
        // but possible: |acc| can reach 1536 * 2^14) the warp redoes the batch of 8 with the converting instruction.
        // (x * 1.0f is exact, so the per-tensor mode simply carries a token scale of 1; its bias rides in cst.x)
        auto deq16 = [&](const uint32_t* raw, int ci, float (&f)[16]) {
          uint32_t bad = 0;
#pragma unroll
          for (int j = 0; j < 16; j += 2) {
            const int4 cst = *reinterpret_cast<const int4*>(s_cst + ci + j);  // two columns per shared-memory load
            const uint32_t b0 = raw[j] + (uint32_t)cst.x, b1 = raw[j + 1] + (uint32_t)cst.z;
            bad |= (b0 ^ 0x4B000000u) | (b1 ^ 0x4B000000u);
            f[j] = __fmul_rn(__fmul_rn(__fadd_rn(__uint_as_float(b0), -kMagicF), __int_as_float(cst.y)), ts_u);
            f[j + 1] = __fmul_rn(__fmul_rn(__fadd_rn(__uint_as_float(b1), -kMagicF), __int_as_float(cst.w)), ts_u);
          }
          if (__any_sync(0xffffffffu, (bad & 0xFF800000u) != 0)) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int2 cst = s_cst[ci + j];
              f[j] = __fmul_rn(__fmul_rn(__int2float_rn((int)(raw[j] + (uint32_t)cst.x) - kMagic), __int_as_float(cst.y)), ts_u);
            }
          }
        };
        // q_nope tile g is free once the previous unit's absorb MMAs have retired (normally long before this unit's MM2 has)
        if (it > 0) ptx::mbar_wait(ptx::smem_u32(qn_free), (it - 1) & 1);
        __syncwarp();
        tl_trace(tl, tr, tb + 1);
        const uint32_t tile = ptx::smem_u32(smem + kPNopeOff + g * kTile16K);
        float xl[16], xh[16];
        {
          // 16 columns per TMEM load, the next load in flight while the current one is converted (96 registers per thread)
          uint32_t r0[16], r1[16];
          ptx::tmem_ld_32x32b_x16(acc2 + 64 * g, r0);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint32_t (&cur)[16] = (q & 1) ? r1 : r0;
            uint32_t (&nxt)[16] = (q & 1) ? r0 : r1;
            ptx::tmem_ld_wait();
            tl_trace(tl, tr, tb + 2 + 2 * q);
            if (q < 3) ptx::tmem_ld_32x32b_x16(acc2 + 64 * g + 16 * (q + 1), nxt);
            else ptx::tmem_ld_32x32b_x16(acc2 + 128 + 16 * g, nxt);  // rope dims 16g..: lands in r0
            {
              float f[16];
              deq16(cur, 16 * q, f);
#pragma unroll
              for (int b8 = 0; b8 < 2; ++b8)
                ptx::st_shared_v4(tile + ptx::sw128_offset(row, 2 * q + b8), Act<T>::pack(f[8 * b8], f[8 * b8 + 1]), Act<T>::pack(f[8 * b8 + 2], f[8 * b8 + 3]),
                                  Act<T>::pack(f[8 * b8 + 4], f[8 * b8 + 5]), Act<T>::pack(f[8 * b8 + 6], f[8 * b8 + 7]));
            }
            tl_trace(tl, tr, tb + 3 + 2 * q);
          }
          ptx::tmem_ld_wait();
          ptx::tmem_ld_32x32b_x16(acc2 + 160 + 16 * g, r1);  // their partners 32 + 16g..
          deq16(r0, 64, xl);
#pragma unroll
          for (int j = 0; j < 16; ++j) xl[j] = Act<T>::round(xl[j]);
          ptx::tmem_ld_wait();
          deq16(r1, 80, xh);
#pragma unroll
          for (int j = 0; j < 16; ++j) xh[j] = Act<T>::round(xh[j]);
        }
        tl_trace(tl, tr, tb + 10);
        if (u + num_pairs < total_units) fetch_unit(u + num_pairs);  // consumed when the next unit starts
        // cos/sin of this row: an L2 round trip is ~2.5k cycles under this load; requested here (earlier costs spills), used
        // after the hand-off to the absorb
        uint4 cs4[4], sn4[4];
        {
          const uint4* cp = reinterpret_cast<const uint4*>(cos_row);
          const uint4* sp = reinterpret_cast<const uint4*>(sin_row);
          // 32 bytes per request: a warp's rows are 32 different lines either way, so half the requests is half the time the
          // load/store unit spends on them (it also carries the drain set's staging stores)
          ptx::ld_global_nc_v8(cp, cs4[0], cs4[1]);
          ptx::ld_global_nc_v8(sp, sn4[0], sn4[1]);
          ptx::ld_global_nc_v8(cp + 4, cs4[2], cs4[3]);
          ptx::ld_global_nc_v8(sp + 4, sn4[2], sn4[3]);
        }
        ptx::fence_proxy_async_smem();
        ptx::tc_fence_before();
        tl_trace(tl, tr, tb + 11);
        ptx::named_bar_sync(g_bar, 128);
        tl_trace(tl, tr, tb + 12);
        if (issuer) ptx::mbar_arrive_cluster(r_acc_read);  // this group's share of acc2 is read, its q_nope tile is written
        tl_trace(tl, tr, tb + 13);
        if (issuer && g == 1 && it < 3) tl_mark(tl, 6 + 3 * it);

        // RoPE of dims [16g,16g+16) and their partners, 32-byte sectors straight to q_out1 (test_mla:57-67)
        if (valid) {
          T* dst = reinterpret_cast<T*>(q_out1) + ((long long)t * qa.num_heads + head) * MLAPO_ROPE + 16 * g;
          uint32_t wl[8], wh[8];
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            float c_lo[8], s_lo[8], c_hi[8], s_hi[8];
            unpack8<T>(cs4[j], c_lo);
            unpack8<T>(sn4[j], s_lo);
            unpack8<T>(cs4[2 + j], c_hi);
            unpack8<T>(sn4[2 + j], s_hi);
#pragma unroll
            for (int v = 0; v < 4; ++v) {
              const int i0 = 8 * j + 2 * v, i1 = i0 + 1;
              const float a0 = __fadd_rn(__fmul_rn(xl[i0], c_lo[2 * v]), __fmul_rn(-xh[i0], s_lo[2 * v]));
              const float a1 = __fadd_rn(__fmul_rn(xl[i1], c_lo[2 * v + 1]), __fmul_rn(-xh[i1], s_lo[2 * v + 1]));
              const float b0 = __fadd_rn(__fmul_rn(xh[i0], c_hi[2 * v]), __fmul_rn(xl[i0], s_hi[2 * v]));
              const float b1 = __fadd_rn(__fmul_rn(xh[i1], c_hi[2 * v + 1]), __fmul_rn(xl[i1], s_hi[2 * v + 1]));
              wl[4 * j + v] = Act<T>::pack(a0, a1);
              wh[4 * j + v] = Act<T>::pack(b0, b1);
            }
          }
          ptx::st_global_v8(dst, wl[0], wl[1], wl[2], wl[3], wl[4], wl[5], wl[6], wl[7]);
          ptx::st_global_v8(dst + 32, wh[0], wh[1], wh[2], wh[3], wh[4], wh[5], wh[6], wh[7]);
        }
        tl_trace(tl, tr, tb + 14);
      }
    } else {
      // ---------------- drain set.  act out: group g drains columns [64g, 64g+64) of every chunk (one 128-byte row per token);
      //                  int8 out: group g owns TMEM buffer g and drains whole chunks c = g, g+2 (128 int8 = one 128-byte row).
      //                  TMEM -> registers -> this group's swizzled 16 KB staging tile -> TMA store into q_out0.
      const uint32_t st_tile = ptx::smem_u32(smem + kPStageOff + g * kTile16K);
      // test hook (instrumented build only): a drain group that reaches its wait for chunk 2 / 3 late (tests/test_jitter_gpu.py)
      auto late = [&](int c) {
#ifdef MLAPO_INSTRUMENTED
        if (qa.debug_drain_delay && rank == 1 && c >= 2) {
          const long long t0 = clock64();
          while (clock64() - t0 < qa.debug_drain_delay) {}
        }
#else
        (void)c;
#endif
      };
      uint32_t it = 0;
      for (int u = pair; u < total_units; u += num_pairs, ++it) {
        const uint32_t par = it & 1;
        const int head = u / n_tb, tok0 = (u % n_tb) * 256 + (int)rank * kBlockM;
        if (kInt8Out) {
          // S9: int8( fp16( act(acc) * q_nope_scale[h] ) )   (golden test_mla:86-92, 466-470)
          const float qs = Act<T>::to_f(reinterpret_cast<const T*>(qa.q_nope_scale)[head]);
#pragma unroll 1
          for (int cc = 0; cc < 2; ++cc) {
            const int c = g + 2 * cc;
            late(c);
            ptx::mbar_wait(ptx::smem_u32(unit_full + par * 2 + g), cc);  // this barrier's completions 2k, 2k+1 belong to the region's k-th unit
            __syncwarp();
            ptx::tc_fence_after();
            const uint32_t acc3 = lane_base + par * 256 + g * 128;
            uint32_t w[32];
#pragma unroll
            for (int c0 = 0; c0 < 128; c0 += 32) {
              uint32_t r[32];
              ptx::tmem_ld_32x32b_x32(acc3 + c0, r);
              ptx::tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                int q[4];
#pragma unroll
                for (int v = 0; v < 4; ++v) q[v] = sat_i8(__fmul_rn(Act<T>::round(__uint_as_float(r[4 * j + v])), qs));
                w[(c0 >> 2) + j] = pack4_i8(q[0], q[1], q[2], q[3]);
              }
            }
            ptx::tc_fence_before();
            if (issuer) ptx::tma_store_wait_read<0>();   // the group's previous store has finished reading the staging tile
            ptx::named_bar_sync(g_bar, 128);             // ... and every thread of the group is past its TMEM loads of chunk c
            if (issuer) ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(cc == 0 ? buf_empty + g : region_free + par), 0));
#pragma unroll
            for (int j = 0; j < 8; ++j) ptx::st_shared_v4(st_tile + ptx::sw128_offset(row, j), w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
            ptx::fence_proxy_async_smem();
            ptx::named_bar_sync(g_bar, 128);
            if (issuer) {
              ptx::tma_store_2d(&tm_out0, st_tile, head * MLAPO_KV_LORA + c * 128, tok0);
              ptx::tma_store_commit();
            }
          }
        } else {
#pragma unroll 1
          for (int c = 0; c < 4; ++c) {
            late(c);
            ptx::mbar_wait(ptx::smem_u32(unit_full + par * 2 + (c & 1)), c >> 1);  // completions 2k, 2k+1 belong to the region's k-th unit
            __syncwarp();
            ptx::tc_fence_after();
            uint32_t w[32];
            {
              const uint32_t acc3 = lane_base + par * 256 + (c & 1) * 128 + g * 64;
#pragma unroll
              for (int hh = 0; hh < 2; ++hh) {  // 32 columns at a time keeps the live registers under the 96 of a 640-thread CTA
                uint32_t r[32];
                ptx::tmem_ld_32x32b_x32(acc3 + 32 * hh, r);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 16; ++j) w[16 * hh + j] = Act<T>::pack(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
              }
            }
            ptx::tc_fence_before();
            if (issuer) ptx::tma_store_wait_read<0>();
            ptx::named_bar_sync(g_bar, 128);
            if (issuer) ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(c < 2 ? buf_empty + c : region_free + par), 0));
#pragma unroll
            for (int j = 0; j < 8; ++j) ptx::st_shared_v4(st_tile + ptx::sw128_offset(row, j), w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
            ptx::fence_proxy_async_smem();
            ptx::named_bar_sync(g_bar, 128);
            if (issuer) {
              ptx::tma_store_2d(&tm_out0, st_tile, head * MLAPO_KV_LORA + c * 128 + g * 64, tok0);
              ptx::tma_store_commit();
            }
          }
        }
        if (issuer && g == 1 && it < 3) tl_mark(tl, 7 + 3 * it);  // unit `it` drained
      }
      if (issuer) ptx::tma_store_wait_read<0>();
    }
  }

  // nobody leaves while the peer may still address this CTA's shared memory, barriers or tensor memory
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_arrive();
  ptx::cluster_wait();
  if (warp == 1) ptx::tmem_dealloc_pair(tmem, 512);
  if (threadIdx.x == 32) { tl_mark(tl, 14); tl_mark_global(tl, 15); }
}

}  // namespace mlapo
