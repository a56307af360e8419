// k7w: the q path on CTA pairs with 512 tokens per work unit (prefill).
//
// k7 (mlapo_q_pair.cuh) is bound by what the L2 can deliver: every (head, 256 tokens) unit pulls 2 x 409 KB through it
// (a1 tile + half of W_UQ[head] + half of W_UK[head] per CTA) and pushes 2 x 147 KB back.  Here a pair owns
// (head, 512 tokens): CTA r holds tokens [512u + 256r, +256) as two sets of 128 rows, and every weight byte it fetches is
// used by both sets, so per 256 tokens the pair reads 2 x 302 KB instead of 2 x 409 KB.
//   MM2    acc2[s] = a1[set s] . Wuq[head]^T      kind::i8, M=256 (2x128), N=192, TMEM columns 256s + [0,192)
//   epi2   warps 4..19 = four groups of 128 threads (thread = token row); groups 0,1 own set 0, groups 2,3 set 1.  Group g of a
//          set takes nope columns [64g,64g+64) -> bf16 pairs -> TMEM columns 256s + 192 + [32g,32g+32) (the absorb reads its A
//          operand from tensor memory: no shared-memory copy of q_nope) and rope dims [16g,16g+16) with their partners -> RoPE -> q_out1
//   BMM3   acc3[s] = q_nope[s] . Wuk[head][:, chunk c]   kind::f16, A from TMEM, M=256, N=128 (each CTA holds 64 of the 128
//          columns), TMEM columns 256s + [0,128) (the nope accumulators, already read); order (c0,s0) (c0,s1) (c1,s0) ...: a
//          W_UK chunk is fetched once per unit and the two sets' buffers alternate between the tensor pipe and the drains
//   epi3   the same groups then drain their set's chunks: act out — group g drains columns [64g,64g+64) of every chunk through
//          its 16 KB staging tile -> TMA store; int8 out — group g drains chunks g, g+2 whole (128 int8 = one 128-byte row).
//          The staging tiles of groups 2,3 live in the a1 half of pipeline stage 2 (idle until the next unit's third k-block).
// All 512 TMEM columns are in use, so MM2 of the next unit waits for the last drains of this one; its first three k-blocks are
// already in shared memory by then (the pipeline stages are free from the moment MM2 retires).
// Warps: 0 TMA a1+W_UQ | 1 MM2 issuer (leader CTA), TMEM alloc | 2 TMA W_UK | 3 absorb issuer (leader CTA) | 4..19 epilogue.
#pragma once
#include "mlapo_q_pair.cuh"

namespace mlapo {

constexpr int kPWStages = 3;
constexpr int kPWStageBytes = 2 * kTile16K + kPHalfB * kBlockKBytes;   // 45056: a1 set 0 | a1 set 1 | 96 W_UQ rows
constexpr int kPWWukSlots = 3;
constexpr int kPWWukOff = kPWStages * kPWStageBytes;                   // 135168
constexpr int kPWStageOff = kPWWukOff + kPWWukSlots * kTile16K;        // 184320: 2 x 16 KB output staging
constexpr int kPWBarOff = kPWStageOff + 2 * kTile16K;                  // 217088
constexpr int kPWSmemBytes = kPWBarOff + 256 + 4 * kPCstPerGroup * 8;  // 220416
// TMEM columns: [192s, 192s+192) acc2 of set s; [384+64s, +64) q_nope of set s (bf16 pairs); once both sets are dequantised
// the first 384 columns are three 128-column absorb buffers that the eight (chunk, set) products of a unit rotate through
constexpr int kPWAcc = 192;
constexpr int kPWNopeCol = 384;
__host__ __device__ constexpr int pw_buffer(int n) { return (3 - n % 3) % 3; }   // product n = 2 * chunk + set -> buffer 0,2,1,0,2,1,0,2
static_assert(kPWStageBytes % 1024 == 0 && kPWWukOff % 1024 == 0, "swizzled tiles need 1024-byte alignment");
static_assert(kPWSmemBytes <= 232448, "q_pair_wide smem");

template <class T, bool kInt8Out>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kPThreads, 1)
q_pair_wide_kernel(const __grid_constant__ CUtensorMap tm_a1, const __grid_constant__ CUtensorMap tm_wuq,
                   const __grid_constant__ CUtensorMap tm_wuk, const __grid_constant__ CUtensorMap tm_out0,
                   const __grid_constant__ CUtensorMap tm_cos, const __grid_constant__ CUtensorMap tm_sin,
                   const __grid_constant__ CUtensorMap tm_out1, QArgs qa, int total_units, long long* tl) {
  if (threadIdx.x == 0) { tl_mark_global(tl, 0); tl_mark(tl, 1); }
  extern __shared__ __align__(1024) uint8_t smem[];
  if ((ptx::smem_u32(smem) & 1023u) != 0) __trap();
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kPWBarOff);
  uint64_t* full_bar = bars;            // [3] leader's copy used: TMA of both CTAs -> MM2
  uint64_t* empty_bar = bars + 3;       // [3] MM2 commit (both CTAs) -> TMA
  uint64_t* mm2_done = bars + 6;        // [2] per set: MM2 commit (both) -> dequant
  uint64_t* region_free = bars + 8;     // [1] leader's: remote arrives after the drains of products 5..7 -> MM2 of the next unit
  uint64_t* acc_read = bars + 9;        // [2] per set, leader's: 4 remote arrives (CTA x dequant group): acc2 read, q_nope in TMEM
  uint64_t* wuk_full = bars + 11;       // [3] leader's: TMA of both CTAs -> absorb
  uint64_t* wuk_empty = bars + 14;      // [3] absorb commit (both) -> TMA
  uint64_t* unit_full = bars + 17;      // [8] per product (chunk, set): absorb commit (both) -> drains
  uint64_t* buf_empty = bars + 25;      // [3] per buffer, leader's: remote arrives after the drains of products 0..4 -> product n + 3
  uint64_t* stage2_free = bars + 28;    // [1] own CTA: the set-1 drain groups' stores have been read out of pipeline stage 2
  uint64_t* rope_full = bars + 29;      // [2] per set, own CTA: the set's cos rows and sin rows have landed in its two staging tiles
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 31);
  int2* s_cst_all = reinterpret_cast<int2*>(smem + kPWBarOff + 256);
  // arrivals per completion: act out — both groups of a set, in both CTAs, take part in every chunk; int8 out — one group per chunk
  constexpr uint32_t kDrainArrivals = kInt8Out ? 2u : 4u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int n_tb = (qa.m_total + 511) / 512;  // token blocks; units are head-major: u = head * n_tb + token block

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_wuq);
    ptx::prefetch_tmap(&tm_a1);
    ptx::prefetch_tmap(&tm_wuk);
    ptx::prefetch_tmap(&tm_out0);
    ptx::prefetch_tmap(&tm_cos);
    ptx::prefetch_tmap(&tm_sin);
    ptx::prefetch_tmap(&tm_out1);
    for (int i = 0; i < 31; ++i) {
      const uint32_t cnt = i == 8 ? 3u * kDrainArrivals : (i == 9 || i == 10) ? 4u : (i >= 25 && i <= 27) ? kDrainArrivals : i >= 28 ? 2u : 1u;
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
    // ------------------------------------------------------------------ TMA: own 256 rows of a1 + own half of W_UQ per k-block
    if (lane == 0) {
      constexpr uint32_t stage_tx = 2 * kPWStageBytes;  // both CTAs' bytes land on the leader's barrier
      uint32_t gk = 0, it = 0;
      for (int u = pair; u < total_units; u += num_pairs, ++it) {
        const int head = u / n_tb, tok0 = (u % n_tb) * 512 + (int)rank * 256;
        for (int kb = 0; kb < kQKBlocks; ++kb, ++gk) {
          const uint32_t s = gk % kPWStages;
          if (gk >= kPWStages) ptx::mbar_wait(ptx::smem_u32(empty_bar + s), ((gk / kPWStages) & 1) ^ 1);
          // stage 2 doubles as the staging tiles of the set-1 drain groups of the previous unit
          if (it >= 1 && kb == kPWStages - 1) ptx::mbar_wait(ptx::smem_u32(stage2_free), (it - 1) & 1);
          if (leader) ptx::mbar_arrive_expect_tx(ptx::smem_u32(full_bar + s), stage_tx);
          const uint32_t fb = ptx::mapa(ptx::smem_u32(full_bar + s), 0);
          const uint32_t dst = ptx::smem_u32(smem + s * kPWStageBytes);
          ptx::tma_load_2d_pair(dst + 2 * kTile16K, &tm_wuq, fb, kb * kBlockKBytes, head * MLAPO_Q_HEAD + (int)rank * kPHalfB,
                                ptx::kEvictLast);
          if (gk == kPWStages - 1) {
            // the first W_UQ k-blocks depend on nothing; a1 comes from the norm kernel
            ptx::griddep_wait();
            tl_mark(tl, 3);
            for (int j = 0; j < kPWStages; ++j)
              ptx::tma_load_2d_pair(ptx::smem_u32(smem + j * kPWStageBytes), &tm_a1, ptx::mapa(ptx::smem_u32(full_bar + j), 0),
                                    j * kBlockKBytes, tok0, ptx::kEvictLast);
          } else if (gk >= kPWStages) {
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
          const uint32_t w = gc % kPWWukSlots, use = gc / kPWWukSlots;
          if (use >= 1) ptx::mbar_wait(ptx::smem_u32(wuk_empty + w), (use & 1) ^ 1);
          if (leader) ptx::mbar_arrive_expect_tx(ptx::smem_u32(wuk_full + w), 2 * kTile16K);
          ptx::tma_load_2d_pair(ptx::smem_u32(smem + kPWWukOff + w * kTile16K), &tm_wuk, ptx::mapa(ptx::smem_u32(wuk_full + w), 0),
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
        // all three absorb buffers (= both sets' accumulator columns): the previous unit's last drains have read them
        if (it >= 1) ptx::mbar_wait_cluster(ptx::smem_u32(region_free), (it - 1) & 1);
        tl_trace(tl, lane == 0 && (it == 2 || it == 3), 72 + (it - 2));
        for (int kb = 0; kb < kQKBlocks; ++kb, ++gk) {
          const uint32_t s = gk % kPWStages;
          ptx::mbar_wait(ptx::smem_u32(full_bar + s), (gk / kPWStages) & 1);
          __syncwarp();
          ptx::tc_fence_after();
          if (lane == 0) {
            if (it == 0 && kb == 0) tl_mark(tl, 4);
            const uint32_t a_base = ptx::smem_u32(smem + s * kPWStageBytes);
            const uint32_t b_base = a_base + 2 * kTile16K;
            // the two sets' accumulation chains alternate: consecutive instructions never wait on each other's accumulator
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t db = ptx::make_smem_desc(b_base + k * kUmmaKBytes, 1024, 0);
#pragma unroll
              for (int set = 0; set < 2; ++set) {
                const uint64_t da = ptx::make_smem_desc(a_base + set * kTile16K + k * kUmmaKBytes, 1024, 0);
                ptx::umma_i8_pair(tmem + set * kPWAcc, da, db, idesc2, (kb > 0 || k > 0) ? 1u : 0u);
              }
            }
            if (kb == kQKBlocks - 1) {
              ptx::umma_commit_pair(ptx::smem_u32(mm2_done + 0));
              ptx::umma_commit_pair(ptx::smem_u32(mm2_done + 1));
            }
            ptx::umma_commit_pair(ptx::smem_u32(empty_bar + s));
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 3) {
    // ------------------------------------------------------------------ MMA: absorb (leader issues for both SMs), A from TMEM
    if (leader) {
      constexpr uint32_t kAFmt = std::is_same<T, __nv_bfloat16>::value ? ptx::kFmtBF16 : ptx::kFmtF16;
      constexpr uint32_t idesc3 = ptx::make_idesc(ptx::kCFmtF32, kAFmt, kAFmt, 0, 1, 2 * kBlockM, 128);
      uint32_t gc = 0, it = 0;
      uint32_t emp[3] = {0, 0, 0};  // completions of buf_empty[b] consumed so far
      for (int u = pair; u < total_units; u += num_pairs, ++it) {
#pragma unroll
        for (int c = 0; c < 4; ++c, ++gc) {
          const uint32_t w = gc % kPWWukSlots;
          ptx::mbar_wait(ptx::smem_u32(wuk_full + w), (gc / kPWWukSlots) & 1);
          // chunk 0: both sets' accumulators are read and their q_nope is in TMEM (buffer 1, first used by product 2, spans both
          // sets' columns); products 3..7: the buffer's previous product is drained
          if (c == 0) {
            ptx::mbar_wait_cluster(ptx::smem_u32(acc_read + 0), it & 1);
            ptx::mbar_wait_cluster(ptx::smem_u32(acc_read + 1), it & 1);
          }
#pragma unroll
          for (int set = 0; set < 2; ++set) {
            const int n = 2 * c + set, b = pw_buffer(n);
            if (n >= 3) { ptx::mbar_wait_cluster(ptx::smem_u32(buf_empty + b), emp[b] & 1); ++emp[b]; }
          }
          __syncwarp();
          ptx::tc_fence_after();
          tl_trace(tl, lane == 0 && it == 2, 64 + 2 * c);
          if (lane == 0) {
            const uint32_t b_base = ptx::smem_u32(smem + kPWWukOff + w * kTile16K);
            // the two sets' products of this chunk alternate instruction by instruction (independent accumulators)
#pragma unroll
            for (int k = 0; k < 8; ++k) {  // 8 x UMMA_K(16) = 128; 16 bf16 of a row = 8 TMEM columns
              const uint64_t db = ptx::make_smem_desc(b_base + k * 2048, 1024, 0);  // MN-major, one 64-column atom per CTA
#pragma unroll
              for (int set = 0; set < 2; ++set)
                ptx::umma_f16_pair_ts(tmem + pw_buffer(2 * c + set) * 128, tmem + kPWNopeCol + 64 * set + k * 8, db, idesc3, k > 0 ? 1u : 0u);
            }
            ptx::umma_commit_pair(ptx::smem_u32(unit_full + 2 * c));
            ptx::umma_commit_pair(ptx::smem_u32(unit_full + 2 * c + 1));
            ptx::umma_commit_pair(ptx::smem_u32(wuk_empty + w));
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps: thread = token row of ONE set
    // Four groups of 128 threads: groups 0,1 own token set 0, groups 2,3 set 1; inside a set group g takes half of the columns.
    // A group first dequantises its share of the set's MM2 accumulators, then drains its share of the set's absorb chunks.
    const int eg = (warp - 4) >> 2;        // 0..3
    const int set = eg >> 1, g = eg & 1;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int et = threadIdx.x - 128 - eg * 128;   // 0..127 inside the group
    const bool issuer = et == 0;
    const int g_bar = 1 + eg;
    const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);   // this thread's TMEM lane
    const uint32_t region = lane_base + set * kPWAcc;                      // ... and its set's MM2 accumulators
    const uint32_t nope_col = lane_base + kPWNopeCol + 64 * set + 32 * g;  // ... and where its q_nope pairs go
    const bool per_token = qa.tok_scale1 != nullptr;
    // staging tile of the group: groups 0,1 have their own; groups 2,3 borrow the a1 half of pipeline stage 2, which is idle
    // between the last MMA of a unit and the producer's refill (held back until these groups' stores have been read)
    const uint32_t st_tile = ptx::smem_u32(smem + (eg < 2 ? kPWStageOff + eg * kTile16K : 2 * kPWStageBytes + (eg - 2) * kTile16K));
    const uint32_t r_acc_read = ptx::mapa(ptx::smem_u32(acc_read + set), 0);
    const uint32_t r_buf_empty = ptx::mapa(ptx::smem_u32(buf_empty), 0);
    const uint32_t r_region_free = ptx::mapa(ptx::smem_u32(region_free), 0);
    ptx::griddep_wait();  // per-token scales come from the norm kernel
    ptx::griddep_launch_dependents();  // only after the wait: see q_fused

    int2* const s_cst = s_cst_all + eg * kPCstPerGroup;  // [0,64) nope cols 64g.., [64,80) rope dims 16g.., [80,96) 32+16g..
    // Fetched one unit ahead: the column constants {bias + kMagic, descale} of the group's 96 columns and the token scale.
    float ts = 1.f;
    int bias_next = 0;
    float dsc_next = 0.f;
    auto fetch_unit = [&](int u) {
      const int head = u / n_tb, t = (u % n_tb) * 512 + (int)rank * 256 + set * kBlockM + row;
      if (et < kPCstPerGroup) {
        const int col = et < 64 ? 64 * g + et : et < 80 ? 128 + 16 * g + (et - 64) : 160 + 16 * g + (et - 80);
        dsc_next = load_descale(qa.descale1, qa.descale_is_i64, head * MLAPO_Q_HEAD + col);  // consumed when the unit starts
        bias_next = qa.bias1 ? qa.bias1[head * MLAPO_Q_HEAD + col] : 0;
      }
      if (per_token) ts = qa.tok_scale1[t < qa.m_total ? t : qa.m_total - 1];
    };
    if (pair < total_units) fetch_unit(pair);

    uint32_t it = 0;
    for (int u = pair; u < total_units; u += num_pairs, ++it) {
      const int head = u / n_tb, tok0 = (u % n_tb) * 512 + (int)rank * 256 + set * kBlockM;
      // (the group left the previous unit's constants behind at the barrier before its hand-off to the absorb)
      if (et < kPCstPerGroup) s_cst[et] = make_int2(bias_next + kMagic, __float_as_int(dsc_next));  // fetched one unit ago
      const float ts_u = ts;  // 1 in per-tensor mode
      ptx::named_bar_sync(g_bar, 128);  // constants visible to the group
      if (u + num_pairs < total_units) fetch_unit(u + num_pairs);  // consumed when the next unit starts
      const bool tr = issuer && (eg == 0 || eg == 2) && it == 2;  // fine-grained trace of two threads (debug timeline only)
      const int tb = eg * 16;
      tl_trace(tl, tr, tb + 0);

      // ================ dequant: nope columns [64g, 64g+64) -> packed pairs -> TMEM (absorb A operand, K = 64g..64g+63);
      //                  rope dims [16g, 16g+16) and their partners 32 + the same -> rounded to T, kept packed until the
      //                  drains are done (the RoPE itself waits on cos/sin from L2: not on the path to the absorb)
      {
        ptx::mbar_wait(ptx::smem_u32(mm2_done + set), it & 1);
        __syncwarp();
        ptx::tc_fence_after();
        // cos rows (group 0 of the set) / sin rows (group 1) of the set's 128 tokens -> the group's own staging tile: idle since
        // its last store of the previous unit was read, and (tiles in pipeline stage 2) since the MMAs just retired
        if (issuer) {
          ptx::mbar_arrive_expect_tx(ptx::smem_u32(rope_full + set), kTile16K);
          ptx::tma_load_2d(st_tile, g == 0 ? &tm_cos : &tm_sin, ptx::smem_u32(rope_full + set), 0, tok0, ptx::kEvictLast);
        }
        if (issuer && eg == 0 && it < 3) tl_mark(tl, 5 + 3 * it);  // MM2 of unit `it` (set 0) retired
        tl_trace(tl, tr, tb + 1);

        // int32 -> fp32 without the quarter-rate I2F: exact while |acc + bias| < 2^22; otherwise the warp redoes the batch
        // with the converting instruction (see mlapo_q_pair.cuh)
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
        float xl[16], xh[16];
        {
          // 16 columns per TMEM load, the next load in flight while the current one is converted
          uint32_t r0[16], r1[16];
          ptx::tmem_ld_32x32b_x16(region + 64 * g, r0);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint32_t (&cur)[16] = (q & 1) ? r1 : r0;
            uint32_t (&nxt)[16] = (q & 1) ? r0 : r1;
            ptx::tmem_ld_wait();
            if (q < 3) ptx::tmem_ld_32x32b_x16(region + 64 * g + 16 * (q + 1), nxt);
            else ptx::tmem_ld_32x32b_x16(region + 128 + 16 * g, nxt);  // rope dims 16g..: lands in r0
            float f[16];
            deq16(cur, 16 * q, f);
            uint32_t pk[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) pk[j] = Act<T>::pack(f[2 * j], f[2 * j + 1]);
            ptx::tmem_st_32x32b_x8(nope_col + 8 * q, pk);  // K = 64g + 16q .. +15 of this row
          }
          ptx::tmem_ld_wait();
          tl_trace(tl, tr, tb + 2);
          ptx::tmem_ld_32x32b_x16(region + 160 + 16 * g, r1);  // their partners 32 + 16g..
          deq16(r0, 64, xl);
#pragma unroll
          for (int j = 0; j < 16; ++j) xl[j] = Act<T>::round(xl[j]);
          ptx::tmem_ld_wait();
          deq16(r1, 80, xh);
#pragma unroll
          for (int j = 0; j < 16; ++j) xh[j] = Act<T>::round(xh[j]);
        }
        tl_trace(tl, tr, tb + 3);
        ptx::tmem_st_wait();
        ptx::tc_fence_before();
        tl_trace(tl, tr, tb + 4);
        ptx::named_bar_sync(g_bar, 128);
        tl_trace(tl, tr, tb + 5);
        if (issuer) ptx::mbar_arrive_cluster(r_acc_read);  // this group's share of acc2 is read, its q_nope columns are written
        if (issuer && eg == 3 && it < 3) tl_mark(tl, 6 + 3 * it);

        // RoPE of dims [16g,16g+16) and their partners (test_mla:57-67): cos/sin from the set's two tiles, results into the
        // first of them, one TMA store of the set's 128 x 64 block of q_out1.  (Per-thread global loads and stores here — 32
        // different lines per warp instruction — kept the load/store unit busy for thousands of cycles and held up the
        // TMA issue behind them.)
        {
          const uint32_t cos_tile = ptx::smem_u32(smem + (set == 0 ? kPWStageOff : 2 * kPWStageBytes));
          const uint32_t sin_tile = cos_tile + kTile16K;
          ptx::mbar_wait(ptx::smem_u32(rope_full + set), it & 1);
          uint32_t wl[8], wh[8];
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            float c_lo[8], s_lo[8], c_hi[8], s_hi[8];
            unpack8<T>(ptx::ld_shared_v4(cos_tile + ptx::sw128_offset(row, 2 * g + j)), c_lo);
            unpack8<T>(ptx::ld_shared_v4(sin_tile + ptx::sw128_offset(row, 2 * g + j)), s_lo);
            unpack8<T>(ptx::ld_shared_v4(cos_tile + ptx::sw128_offset(row, 4 + 2 * g + j)), c_hi);
            unpack8<T>(ptx::ld_shared_v4(sin_tile + ptx::sw128_offset(row, 4 + 2 * g + j)), s_hi);
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
          ptx::named_bar_sync(5 + set, 256);  // both groups of the set have read cos and sin
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            ptx::st_shared_v4(cos_tile + ptx::sw128_offset(row, 2 * g + j), wl[4 * j], wl[4 * j + 1], wl[4 * j + 2], wl[4 * j + 3]);
            ptx::st_shared_v4(cos_tile + ptx::sw128_offset(row, 4 + 2 * g + j), wh[4 * j], wh[4 * j + 1], wh[4 * j + 2], wh[4 * j + 3]);
          }
          ptx::fence_proxy_async_smem();
          ptx::named_bar_sync(5 + set, 256);
          if (issuer && g == 0) {
            ptx::tma_store_2d(&tm_out1, cos_tile, head * MLAPO_ROPE, tok0);
            ptx::tma_store_commit();
          }
        }
      }
      tl_trace(tl, tr, tb + 6);
      // ================ drain: this set's absorb chunks, TMEM -> registers -> the group's swizzled 16 KB staging tile -> TMA store
      //                  act out: columns [64g, 64g+64) of every chunk (one 128-byte row per token);
      //                  int8 out: chunks c = g, g+2 whole (128 int8 = one 128-byte row)
      if (kInt8Out) {
        // S9: int8( fp16( act(acc) * q_nope_scale[h] ) )   (golden test_mla:86-92, 466-470)
        const float qs = Act<T>::to_f(reinterpret_cast<const T*>(qa.q_nope_scale)[head]);
#pragma unroll 1
        for (int cc = 0; cc < 2; ++cc) {
          const int c = g + 2 * cc, n = 2 * c + set, b = pw_buffer(n);
          ptx::mbar_wait(ptx::smem_u32(unit_full + n), it & 1);
          __syncwarp();
          ptx::tc_fence_after();
          uint32_t w[32];
#pragma unroll
          for (int c0 = 0; c0 < 128; c0 += 32) {
            uint32_t r[32];
            ptx::tmem_ld_32x32b_x32(lane_base + b * 128 + c0, r);
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
          ptx::named_bar_sync(g_bar, 128);             // ... and every thread of the group is past its TMEM loads of this chunk
          if (issuer) ptx::mbar_arrive_cluster(n <= 4 ? r_buf_empty + 8 * b : r_region_free);  // product n + 3 reuses the buffer
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
          const int n = 2 * c + set, b = pw_buffer(n);
          ptx::mbar_wait(ptx::smem_u32(unit_full + n), it & 1);
          __syncwarp();
          ptx::tc_fence_after();
          tl_trace(tl, tr, tb + 7 + 6 * c);
          uint32_t w[32];
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {  // 32 columns at a time keeps the live registers under the 96 of a 640-thread CTA
            uint32_t r[32];
            ptx::tmem_ld_32x32b_x32(lane_base + b * 128 + g * 64 + 32 * hh, r);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) w[16 * hh + j] = Act<T>::pack(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
          }
          ptx::tc_fence_before();
          tl_trace(tl, tr, tb + 8 + 6 * c);
          if (issuer) ptx::tma_store_wait_read<0>();
          tl_trace(tl, tr, tb + 9 + 6 * c);
          ptx::named_bar_sync(g_bar, 128);
          if (issuer) ptx::mbar_arrive_cluster(n <= 4 ? r_buf_empty + 8 * b : r_region_free);  // product n + 3 reuses the buffer
          tl_trace(tl, tr, tb + 10 + 6 * c);
#pragma unroll
          for (int j = 0; j < 8; ++j) ptx::st_shared_v4(st_tile + ptx::sw128_offset(row, j), w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
          ptx::fence_proxy_async_smem();
          tl_trace(tl, tr, tb + 11 + 6 * c);
          ptx::named_bar_sync(g_bar, 128);
          if (issuer) {
            ptx::tma_store_2d(&tm_out0, st_tile, head * MLAPO_KV_LORA + c * 128 + g * 64, tok0);
            ptx::tma_store_commit();
          }
          tl_trace(tl, tr, tb + 12 + 6 * c);
        }
      }
      if (issuer) {  // the tile takes cos / sin rows next; groups 2,3 also hand pipeline stage 2 back to the producer
        ptx::tma_store_wait_read<0>();
        if (eg >= 2) ptx::mbar_arrive(ptx::smem_u32(stage2_free));
      }
      if (issuer && eg == 3 && it < 3) tl_mark(tl, 7 + 3 * it);  // unit `it` drained

    }
    if (issuer) ptx::tma_store_wait_read<0>();
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
