// k7: the q path on CTA pairs (tcgen05 cta_group::2) — same arithmetic as k4 (mlapo_q.cuh), different ownership of the data.
//
// A pair of CTAs on neighbouring SMs owns one work unit = (head, block of 256 tokens) at a time; CTA r of the pair holds
// token tile r (128 rows).  One MMA instruction spans both SMs, so per k-block each CTA ingests only ITS a1 tile (16 KB) and
// HALF of the head's W_UQ rows (12 KB) — k4 ingests 2 x 16 KB + 24 KB per CTA for the same 256 tokens — and each CTA needs
// only 192 TMEM columns per unit.  That leaves room for a second accumulator set, so a pair walks a list of units and
// MM2 of unit i+1 streams in while unit i is dequantised, absorbed and stored:
//   MM2    acc2 = a1 . Wuq[head]^T        kind::i8, M=256 (2x128), N=192, TMEM region (i&1)*256 + [0,192)
//   epi2   16 epilogue warps = 4 groups of 128 threads (thread = token row).  Group q takes nope columns [32q,32q+32)
//          -> swizzled smem (absorb A operand) and rope dims [8q,8q+8) with their partners +32 -> RoPE -> q_out1
//   BMM3   acc3 = q_nope . Wuk[head][:, chunk c]   kind::f16, M=256, N=128 (each CTA holds 64 of the 128 columns),
//          buffer c&1 of the same TMEM region: [0,128) (the nope columns, already drained) / [128,256)
//   epi3   act out: groups {0,1} / {2,3} form two teams; team t drains columns [64t,64t+64) of every chunk through one
//          16 KB staging tile -> TMA store.   int8 out: team t drains whole chunks c = t, t+2 (128 int8 = one 128-byte row)
// Warps: 0 TMA a1+W_UQ | 1 MM2 issuer (leader CTA), TMEM alloc | 2 TMA W_UK | 3 absorb issuer (leader CTA) | 4..19 epilogue.
// Barriers live at the same offset in both CTAs.  TMA completions of both CTAs are counted on the LEADER's full
// barriers; tcgen05.commit multicasts to both CTAs; epilogue -> MMA signals are remote arrives on the leader's copies.
#pragma once
#include "mlapo_q.cuh"

namespace mlapo {

constexpr int kPThreads = 640;
constexpr int kPStages = 4;
constexpr int kPHalfB = MLAPO_Q_HEAD / 2;                            // 96 W_UQ rows per CTA
constexpr int kPStageBytes = kTile16K + kPHalfB * kBlockKBytes;      // 28672
constexpr int kPNopeOff = kPStages * kPStageBytes;                   // 114688: 2 x 16 KB (k-block 0 / 1 of this CTA's token tile)
constexpr int kPWukSlots = 3;
constexpr int kPWukOff = kPNopeOff + 2 * kTile16K;                   // 147456: 3 x 16 KB ([128 k x 64 n] per slot)
constexpr int kPStageOff = kPWukOff + kPWukSlots * kTile16K;         // 196608: 2 x 16 KB output staging (one per team)
constexpr int kPBarOff = kPStageOff + 2 * kTile16K;                  // 229376
constexpr int kPCstPerGroup = 48;
constexpr int kPSmemBytes = kPBarOff + 256 + 4 * kPCstPerGroup * 8 + 1024;  // 232192
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
  uint64_t* acc_read = bars + 12;       // [1] leader's: 8 remote arrives (CTA x group): acc2 read, q_nope tiles written -> absorb
  uint64_t* qn_free = bars + 13;        // [1] absorb commit (both) -> epilogue may overwrite the q_nope tiles
  uint64_t* wuk_full = bars + 14;       // [3] leader's: TMA of both CTAs -> absorb
  uint64_t* wuk_empty = bars + 17;      // [3] absorb commit (both) -> TMA
  uint64_t* unit_full = bars + 20;      // [2] absorb commit (both) -> drains of TMEM buffer b
  uint64_t* buf_empty = bars + 22;      // [2] leader's: remote arrives after chunks 0 / 1 -> absorb chunk c+2
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 24);
  int2* s_cst_all = reinterpret_cast<int2*>(smem + kPBarOff + 256);
  // arrivals per completion: act out — both teams drain every chunk; int8 out — team t alone drains chunks t, t+2
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
    for (int i = 0; i < 24; ++i) {
      const uint32_t cnt = (i == 10 || i == 11) ? kRegionFreeCount : i == 12 ? 8u : (i == 22 || i == 23) ? kBufEmptyCount : 1u;
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
  ptx::griddep_launch_dependents();
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
          if (lane == 0) {
            if (it == 0 && kb == 0) tl_mark(tl, 4);
            const uint32_t a_base = ptx::smem_u32(smem + s * kPStageBytes);
            const uint32_t b_base = a_base + kTile16K;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t da = ptx::make_smem_desc(a_base + k * kUmmaKBytes, 1024, 0);
              const uint64_t db = ptx::make_smem_desc(b_base + k * kUmmaKBytes, 1024, 0);
              ptx::umma_i8_pair(tmem + par * 256, da, db, idesc2, (kb > 0 || k > 0) ? 1u : 0u);
            }
            ptx::umma_commit_pair(ptx::smem_u32(empty_bar + s));
          }
          __syncwarp();
        }
        if (lane == 0) ptx::umma_commit_pair(ptx::smem_u32(mm2_done + par));
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
          if (lane == 0) {
            const uint32_t b_base = ptx::smem_u32(smem + kPWukOff + w * kTile16K);
#pragma unroll
            for (int k = 0; k < 8; ++k) {  // 8 x UMMA_K(16) = 128
              const uint32_t a_addr = ptx::smem_u32(smem + kPNopeOff + (k >> 2) * kTile16K) + (k & 3) * kUmmaKBytes;
              const uint64_t da = ptx::make_smem_desc(a_addr, 1024, 0);
              const uint64_t db = ptx::make_smem_desc(b_base + k * 2048, 1024, 0);  // MN-major, one 64-column atom per CTA
              ptx::umma_f16_pair(tmem + par * 256 + (c & 1) * 128, da, db, idesc3, k > 0 ? 1u : 0u);
            }
            ptx::umma_commit_pair(ptx::smem_u32(wuk_empty + w));
            ptx::umma_commit_pair(ptx::smem_u32(unit_full + (c & 1)));
          }
          __syncwarp();
        }
        if (lane == 0) ptx::umma_commit_pair(ptx::smem_u32(qn_free));  // every MMA that reads the q_nope tiles has retired
        __syncwarp();
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue: thread = token row, group = column share
    const int grp = (warp - 4) >> 2;       // 0..3
    const int team = grp >> 1, sub = grp & 1;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int et = threadIdx.x - 128 - grp * 128;   // 0..127 inside the group
    const bool g_issuer = et == 0;
    const bool t_issuer = g_issuer && sub == 0;     // one thread per team hands TMEM buffers back
    const int g_bar = 1 + grp, t_bar = 5 + team;
    const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
    const bool per_token = qa.tok_scale1 != nullptr;
    const uint32_t r_acc_read = ptx::mapa(ptx::smem_u32(acc_read), 0);
    constexpr int kMagic = 0x4B400000;             // bits of 1.5 * 2^23: int v, |v| < 2^22  ->  float(v) = as_float(kMagic + v) - 1.5 * 2^23
    constexpr float kMagicF = 12582912.f;
    ptx::griddep_wait();  // per-token scales come from the norm kernel

    // Per-unit inputs are fetched one unit ahead: the column constants {bias + kMagic, descale} of the group's 48 columns go to
    // s_cst[unit parity], this row's cos/sin of dims [8q,8q+8) and 32 + the same and its token scale stay in registers.
    uint4 cs_lo, sn_lo, cs_hi, sn_hi;
    float ts = 1.f;
    int2 cst_next = make_int2(0, 0);
    auto fetch_unit = [&](int u) {
      const int head = u / n_tb, t = (u % n_tb) * 256 + (int)rank * kBlockM + row;
      const int tc = t < qa.m_total ? t : qa.m_total - 1;  // clamped row for loads
      if (et < kPCstPerGroup) {
        const int col = et < 32 ? 32 * grp + et : et < 40 ? 128 + 8 * grp + (et - 32) : 160 + 8 * grp + (et - 40);
        const float dsc = load_descale(qa.descale1, qa.descale_is_i64, head * MLAPO_Q_HEAD + col);
        cst_next = make_int2((qa.bias1 ? qa.bias1[head * MLAPO_Q_HEAD + col] : 0) + kMagic, __float_as_int(dsc));
      }
      if (per_token) ts = qa.tok_scale1[tc];
      const uint4* cp = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(qa.cos) + (long long)tc * 64 + 8 * grp);
      const uint4* sp = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(qa.sin) + (long long)tc * 64 + 8 * grp);
      cs_lo = ptx::ld_global_nc_v4(cp); sn_lo = ptx::ld_global_nc_v4(sp);
      cs_hi = ptx::ld_global_nc_v4(cp + 4); sn_hi = ptx::ld_global_nc_v4(sp + 4);
    };
    if (pair < total_units) fetch_unit(pair);
    int2* const s_cst = s_cst_all + grp * kPCstPerGroup;  // [0,32) nope cols 32q.., [32,40) rope dims 8q.., [40,48) 32+8q..
    const uint32_t st_tile = ptx::smem_u32(smem + kPStageOff + team * kTile16K);

    uint32_t it = 0;
    for (int u = pair; u < total_units; u += num_pairs, ++it) {
      const uint32_t par = it & 1;
      const int head = u / n_tb, tok0 = (u % n_tb) * 256 + (int)rank * kBlockM;
      const int t = tok0 + row;
      const bool valid = t < qa.m_total;
      if (et < kPCstPerGroup) s_cst[et] = cst_next;  // fetched one unit ago; the group's previous unit is completely done
      const float ts_u = ts;
      const uint4 cs_lo_u = cs_lo, sn_lo_u = sn_lo, cs_hi_u = cs_hi, sn_hi_u = sn_hi;
      ptx::named_bar_sync(g_bar, 128);  // constants visible to the group

      ptx::mbar_wait(ptx::smem_u32(mm2_done + par), (it >> 1) & 1);
      __syncwarp();
      ptx::tc_fence_after();
      if (g_issuer && grp == 0 && it < 3) tl_mark(tl, 5 + 3 * it);  // MM2 of unit `it` retired (as seen by the epilogue)
      const uint32_t acc2 = lane_base + par * 256;

      // int32 -> fp32 without the quarter-rate I2F: exact while |acc + bias| < 2^22; otherwise (never seen with int8 data,
      // but possible: |acc| can reach 1536 * 2^14) the warp redoes the batch of 8 with the converting instruction.
      auto deq8 = [&](const uint32_t* raw, int ci, float (&f)[8]) {
        uint32_t bad = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int2 cst = s_cst[ci + j];
          const uint32_t bits = raw[j] + (uint32_t)cst.x;
          bad |= bits ^ 0x4B000000u;
          const float v = __fadd_rn(__uint_as_float(bits), -kMagicF);
          f[j] = per_token ? __fmul_rn(__fmul_rn(v, __int_as_float(cst.y)), ts_u) : __fmul_rn(v, __int_as_float(cst.y));
        }
        if (__any_sync(0xffffffffu, (bad & 0xFF800000u) != 0)) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int2 cst = s_cst[ci + j];
            const float v = __int2float_rn((int)(raw[j] + (uint32_t)cst.x) - kMagic);
            f[j] = per_token ? __fmul_rn(__fmul_rn(v, __int_as_float(cst.y)), ts_u) : __fmul_rn(v, __int_as_float(cst.y));
          }
        }
      };
      float xl[8], xh[8];
      uint32_t wn[16];
      {
        uint32_t ra[32];
        ptx::tmem_ld_32x32b_x32(acc2 + 32 * grp, ra);
        ptx::tmem_ld_wait();
        uint32_t rl[8], rh[8];
        ptx::tmem_ld_32x32b_x8(acc2 + 128 + 8 * grp, rl);
        ptx::tmem_ld_32x32b_x8(acc2 + 160 + 8 * grp, rh);
#pragma unroll
        for (int b8 = 0; b8 < 4; ++b8) {
          float f[8];
          deq8(ra + 8 * b8, 8 * b8, f);
#pragma unroll
          for (int j = 0; j < 4; ++j) wn[4 * b8 + j] = Act<T>::pack(f[2 * j], f[2 * j + 1]);
        }
        ptx::tmem_ld_wait();
        deq8(rl, 32, xl);
        deq8(rh, 40, xh);
#pragma unroll
        for (int j = 0; j < 8; ++j) { xl[j] = Act<T>::round(xl[j]); xh[j] = Act<T>::round(xh[j]); }
      }
      // this group's 64 bytes of every q_nope row (absorb A operand): free once the previous unit's absorb MMAs have retired
      if (it > 0) ptx::mbar_wait(ptx::smem_u32(qn_free), (it - 1) & 1);
      __syncwarp();
      {
        const uint32_t tile = ptx::smem_u32(smem + kPNopeOff + team * kTile16K);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          ptx::st_shared_v4(tile + ptx::sw128_offset(row, sub * 4 + j), wn[4 * j], wn[4 * j + 1], wn[4 * j + 2], wn[4 * j + 3]);
      }
      ptx::fence_proxy_async_smem();
      ptx::tc_fence_before();
      ptx::named_bar_sync(g_bar, 128);
      if (g_issuer) ptx::mbar_arrive_cluster(r_acc_read);  // this group's share of acc2 is read, its q_nope bytes are written
      if (g_issuer && grp == 3 && it < 3) tl_mark(tl, 6 + 3 * it);

      // RoPE of dims [8q,8q+8) and their partners, straight to q_out1 (test_mla:57-67)
      if (valid) {
        T* dst = reinterpret_cast<T*>(q_out1) + ((long long)t * qa.num_heads + head) * MLAPO_ROPE + 8 * grp;
        float c_lo[8], s_lo[8], c_hi[8], s_hi[8];
        unpack8<T>(cs_lo_u, c_lo);
        unpack8<T>(sn_lo_u, s_lo);
        unpack8<T>(cs_hi_u, c_hi);
        unpack8<T>(sn_hi_u, s_hi);
        uint32_t wl[4], wh[4];
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          const int i0 = 2 * v, i1 = i0 + 1;
          const float a0 = __fadd_rn(__fmul_rn(xl[i0], c_lo[i0]), __fmul_rn(-xh[i0], s_lo[i0]));
          const float a1 = __fadd_rn(__fmul_rn(xl[i1], c_lo[i1]), __fmul_rn(-xh[i1], s_lo[i1]));
          const float b0 = __fadd_rn(__fmul_rn(xh[i0], c_hi[i0]), __fmul_rn(xl[i0], s_hi[i0]));
          const float b1 = __fadd_rn(__fmul_rn(xh[i1], c_hi[i1]), __fmul_rn(xl[i1], s_hi[i1]));
          wl[v] = Act<T>::pack(a0, a1);
          wh[v] = Act<T>::pack(b0, b1);
        }
        ptx::st_global_v4(dst, wl[0], wl[1], wl[2], wl[3]);
        ptx::st_global_v4(dst + 32, wh[0], wh[1], wh[2], wh[3]);
      }
      // next unit's inputs: their latency hides behind this unit's drains
      if (u + num_pairs < total_units) fetch_unit(u + num_pairs);

      // drains: TMEM -> registers -> swizzled staging tile of the team -> TMA store into q_out0
      if (kInt8Out) {
        // S9: int8( fp16( act(acc) * q_nope_scale[h] ) )   (golden test_mla:86-92, 466-470).  Team t owns TMEM buffer t:
        // chunks c = t, t+2; group `sub` converts columns [64 sub, 64 sub + 64) = 64 bytes of the 128-byte output row.
        const float qs = Act<T>::to_f(reinterpret_cast<const T*>(qa.q_nope_scale)[head]);
#pragma unroll 1
        for (int cc = 0; cc < 2; ++cc) {
          const int c = team + 2 * cc;
          ptx::mbar_wait(ptx::smem_u32(unit_full + team), (it * 2 + cc) & 1);
          __syncwarp();
          ptx::tc_fence_after();
          const uint32_t acc3 = lane_base + par * 256 + team * 128 + sub * 64;
          uint32_t w[16];
#pragma unroll
          for (int c0 = 0; c0 < 64; c0 += 32) {
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
          if (t_issuer) ptx::tma_store_wait_read<0>();   // the team's previous store has finished reading the staging tile
          ptx::named_bar_sync(t_bar, 256);               // ... and every thread of the team is past its TMEM loads of chunk c
          if (t_issuer) ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(cc == 0 ? buf_empty + team : region_free + par), 0));
#pragma unroll
          for (int j = 0; j < 4; ++j)
            ptx::st_shared_v4(st_tile + ptx::sw128_offset(row, sub * 4 + j), w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
          ptx::fence_proxy_async_smem();
          ptx::named_bar_sync(t_bar, 256);
          if (t_issuer) {
            ptx::tma_store_2d(&tm_out0, st_tile, head * MLAPO_KV_LORA + c * 128, tok0);
            ptx::tma_store_commit();
          }
        }
      } else {
        // Team t drains columns [64t, 64t+64) of every chunk; group `sub` converts 32 of them = 64 bytes of the staging row.
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          ptx::mbar_wait(ptx::smem_u32(unit_full + (c & 1)), (it * 2 + (c >> 1)) & 1);
          __syncwarp();
          ptx::tc_fence_after();
          uint32_t w[16];
          {
            uint32_t r[32];
            ptx::tmem_ld_32x32b_x32(lane_base + par * 256 + (c & 1) * 128 + team * 64 + sub * 32, r);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) w[j] = Act<T>::pack(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
          }
          ptx::tc_fence_before();
          if (t_issuer) ptx::tma_store_wait_read<0>();
          ptx::named_bar_sync(t_bar, 256);
          if (t_issuer) ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(c < 2 ? buf_empty + c : region_free + par), 0));
#pragma unroll
          for (int j = 0; j < 4; ++j)
            ptx::st_shared_v4(st_tile + ptx::sw128_offset(row, sub * 4 + j), w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
          ptx::fence_proxy_async_smem();
          ptx::named_bar_sync(t_bar, 256);
          if (t_issuer) {
            ptx::tma_store_2d(&tm_out0, st_tile, head * MLAPO_KV_LORA + c * 128 + team * 64, tok0);
            ptx::tma_store_commit();
          }
        }
      }
      if (g_issuer && grp == 3 && it < 3) tl_mark(tl, 7 + 3 * it);  // unit `it` drained
    }
    if (t_issuer) ptx::tma_store_wait_read<0>();
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
