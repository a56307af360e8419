// k4 the q path: MM2 (int8) -> dequant -> RoPE -> W_UK absorb (bf16/fp16), one CTA per (head, up to 256 tokens).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "../../include/mla_preprocess_b200.h"
#include "mlapo_numerics.cuh"
#include "mlapo_mm1.cuh"

namespace mlapo {

// ============================================================================================ k4: fused per-head q path
// One CTA per (head, super-tile of up to 256 tokens = 2 token tiles):
//   MM2   acc2[m] (int32 [128 x 192], TMEM) = a1[m-tile] . Wuq[head]^T              (PpMatmulW8a8Aic, bf16.hpp:1515-1975)
//   epi2  dequant (+bias, *descale) -> act; cols 128..191 -> RoPE -> q_out1;         (PpMatmulW8a8Aiv :1981-2361, RopeFp16 :27-244)
//         cols 0..127 -> act q_nope tile written straight into swizzled smem as the next MMA's A operand
//   BMM3  acc3 (f32 [128 x 128] per unit, TMEM) = q_nope[m] . Wuk[head][:, chunk]    (PpMatmulEinSum :1074-1513)
//   epi3  -> act (or int8 re-quant, EinSumQuant :842-1065) -> swizzled smem -> TMA store into q_out0
// 10 warps: w0 TMA producer, w1 MMA issuer, w2..w5 epilogue group 0 (token tile 0), w6..w9 epilogue group 1 (token tile 1).
// Shared memory (224 KB) is one 4-stage MM2 pipeline (A 2x16 KB + B 24 KB per k-block); once MM2 has retired the same
// bytes are re-used as [Wuk chunks 4x32 KB | q_nope tiles 4x16 KB | output staging 2x16 KB].
// TMEM (512 cols): acc2[m] = [256m, 256m+192); absorb buffers of tile m = [256m, 256m+128), [256m+128, 256m+256).
constexpr int kQThreads = 320;
constexpr int kQStages = 4;
constexpr int kQStageBytes = 2 * kTile16K + MLAPO_Q_HEAD * kBlockKBytes;  // 57344
constexpr int kQWukOff = 0;                                 // 4 x 32 KB (two [128 k x 64 n] atoms per 128-column chunk)
constexpr int kQNopeOff = 4 * 32768;                        // 4 x 16 KB: (m, kb) tiles of the absorb A operand
constexpr int kQStageOff = kQNopeOff + 4 * kTile16K;        // 2 x 16 KB output staging (one per epilogue group)
constexpr int kQBarOff = kQStages * kQStageBytes;           // 229376
static_assert(kQStageOff + 2 * kTile16K == kQBarOff, "smem regions must tile the pipeline exactly");
// No slack for aligning the base: the kernel has no static shared memory, its dynamic region is declared 1024-byte aligned and
// starts right behind the 1 KB the system reserves per CTA (checked at run time).  That leaves 1.25 KB of the SM's 228 KB free,
// exactly what lets ONE block of the next call's quantise kernel (no shared memory of its own, 1 KB reserved) share the SM with
// a q CTA — so the quantiser runs while the q kernel is still busy instead of queueing for its SMs.
constexpr int kQSmemBytes = kQBarOff + 256 + MLAPO_Q_HEAD * 8;
static_assert(kQSmemBytes + 1024 + 1024 <= 233472, "q_fused CTA + one quantise block per SM");
constexpr int kQKBlocks = MLAPO_Q_LORA / kBlockKBytes;      // 12

struct QArgs {
  const void* descale1;       // [192*H]
  const int32_t* bias1;       // [192*H] or null (per-token)
  const float* tok_scale1;    // [N] or null
  const void* cos;            // act [N,64]
  const void* sin;
  const void* q_nope_scale;   // act [H] (int8 out)
  int num_heads;
  int m_total;
  int descale_is_i64;
  int cta_tokens;             // 256 (two token tiles per CTA, weights streamed once per 256 tokens) or 128 (more CTAs when heads are few)
  void* q_out1;               // act [N, H, 64]: written with plain stores by the split forms (kSplit > 1)
  int debug_drain_delay;      // test hook (debug bit20): cycles the pair kernel's drain groups of the second CTA idle before chunks 2, 3
};

// Few heads (a TP slice, a small batch) leave most SMs without a (head, token tile) unit.  kSplit = 2 or 4 CTAs then share a unit:
// each recomputes the nope slice of MM2 (all 128 columns: it is the A operand of every absorb chunk), takes 32 / kSplit of the 32
// RoPE pairs (columns i and i + 32 of the rope slice) and 4 / kSplit of the four 128-column absorb chunks.  Per CTA that is
// 196 KB + (98 + 131) KB / kSplit of weights instead of 426 KB, and 1 / kSplit of the outputs.
template <int kSplit> struct QSplit {
  static_assert(kSplit == 1 || kSplit == 2 || kSplit == 4, "kSplit");
  static constexpr int G = 32 / kSplit;        // RoPE pairs of this CTA
  static constexpr int NB = 128 + 2 * G;       // MM2 columns of this CTA (MMA N): 192, 160, 144
  static constexpr int CH = 4 / kSplit;        // absorb chunks of this CTA
};

// Where one CTA's q-path work item lives.
struct QCtx {
  uint8_t* smem;     // 224 KB pipeline / overlay region, 1024-byte aligned
  uint64_t* bars;    // full[4] empty[4] mm2_done wuk_full[4] qn_ready[2] unit_full[4] buf_empty[4] rope_done[2]
  int2* s_cst;       // per column {bias, descale bits}
  uint32_t tmem;     // base of a 512-column TMEM allocation
  int head, tok0, num_mt, load_mt, z;
  static constexpr int NUM_BARS = 25;

  __device__ void set_work(int head_, int tok_tile, int z_, const QArgs& qa) {
    head = head_;
    z = z_;
    tok0 = tok_tile * qa.cta_tokens;
    num_mt = (qa.cta_tokens > kBlockM && (qa.m_total - tok0) > kBlockM) ? 2 : 1;  // token tiles with valid rows in this CTA
    load_mt = (qa.cta_tokens > kBlockM && qa.m_total > kBlockM) ? 2 : 1;          // token tiles fetched per stage (uniform over the grid)
  }
};

__device__ __forceinline__ void q_init_barriers(uint64_t* bars) {  // one thread
  for (int i = 0; i < 13; ++i) ptx::mbar_init(ptx::smem_u32(bars + i), 1);
  ptx::mbar_init(ptx::smem_u32(bars + 13), 128);
  ptx::mbar_init(ptx::smem_u32(bars + 14), 128);
  for (int i = 0; i < 4; ++i) {
    ptx::mbar_init(ptx::smem_u32(bars + 15 + i), 1);
    ptx::mbar_init(ptx::smem_u32(bars + 19 + i), 128);
  }
  ptx::mbar_init(ptx::smem_u32(bars + 23), 128);
  ptx::mbar_init(ptx::smem_u32(bars + 24), 128);
}

// W_UQ rows of k-block kb for this CTA -> the B area of stage `dst_stage`: the whole head (kSplit == 1), or its nope rows plus
// this CTA's RoPE pairs (rows 128 + z G .. and 160 + z G ..), which land behind one another so that the MMA sees NB contiguous rows.
template <int kSplit>
__device__ __forceinline__ void q_load_wuq(const QCtx& c, const CUtensorMap* tm_wuq, const CUtensorMap* tm_wuq_n, const CUtensorMap* tm_wuq_r,
                                           int kb, uint32_t dst, uint32_t bar) {
  using S = QSplit<kSplit>;
  const int row0 = c.head * MLAPO_Q_HEAD;
  if constexpr (kSplit == 1) {
    ptx::tma_load_2d(dst, tm_wuq, bar, kb * kBlockKBytes, row0, ptx::kEvictFirst);
  } else {
    // the nope rows are fetched by every CTA of the unit: keep them in L2 for the siblings
    ptx::tma_load_2d(dst, tm_wuq_n, bar, kb * kBlockKBytes, row0, ptx::kEvictLast);
    ptx::tma_load_2d(dst + 128 * kBlockKBytes, tm_wuq_r, bar, kb * kBlockKBytes, row0 + 128 + c.z * S::G, ptx::kEvictFirst);
    ptx::tma_load_2d(dst + (128 + S::G) * kBlockKBytes, tm_wuq_r, bar, kb * kBlockKBytes, row0 + 160 + c.z * S::G, ptx::kEvictFirst);
  }
}

// One thread: arm the four MM2 stages and request their W_UQ k-blocks (which depend on nothing upstream).
template <int kSplit>
__device__ __forceinline__ void q_issue_weights(const QCtx& c, const CUtensorMap* tm_wuq, const CUtensorMap* tm_wuq_n, const CUtensorMap* tm_wuq_r) {
  const uint32_t stage_tx = (uint32_t)(c.load_mt * kTile16K + QSplit<kSplit>::NB * kBlockKBytes);
  for (int kb = 0; kb < kQStages; ++kb) {
    const uint32_t fb = ptx::smem_u32(c.bars + kb);
    ptx::mbar_arrive_expect_tx(fb, stage_tx);
    q_load_wuq<kSplit>(c, tm_wuq, tm_wuq_n, tm_wuq_r, kb, ptx::smem_u32(c.smem + kb * kQStageBytes + 2 * kTile16K), fb);
  }
}

// Warp roles of the q path: warp 0 = TMA producer, warp 1 = MMA issuer, warps 2..5 / 6..9 = epilogue groups of token tile 0 / 1.
// The caller has already issued q_issue_weights().  Under programmatic dependent launch the roles wait for the upstream kernel
// before touching a1 / the per-token scales.
template <class T, bool kInt8Out, int kSplit>
__device__ __forceinline__ void q_roles(const QCtx& c, const CUtensorMap* tm_a1, const CUtensorMap* tm_wuq, const CUtensorMap* tm_wuq_n,
                                        const CUtensorMap* tm_wuq_r, const CUtensorMap* tm_wuk, const CUtensorMap* tm_out0,
                                        const CUtensorMap* tm_out1, const QArgs& qa, int warp, int lane, long long* tl) {
  using S = QSplit<kSplit>;
  constexpr int G = S::G, NB = S::NB, CH = S::CH;
  uint8_t* const smem = c.smem;
  uint64_t* const full_bar = c.bars;        // [4]  TMA -> MMA
  uint64_t* const empty_bar = c.bars + 4;   // [4]  MMA -> TMA
  uint64_t* const mm2_done = c.bars + 8;    // [1]
  uint64_t* const wuk_full = c.bars + 9;    // [4]
  uint64_t* const qn_ready = c.bars + 13;   // [2]  epilogue group m -> MMA: q_nope tile m in smem and acc2[m] drained
  uint64_t* const unit_full = c.bars + 15;  // [4]  (m*2 + buffer) MMA -> epilogue
  uint64_t* const buf_empty = c.bars + 19;  // [4]  (m*2 + buffer) epilogue -> MMA
  uint64_t* const rope_done = c.bars + 23;  // [2]  epilogue group m -> MMA: acc2[m] columns 128..191 drained as well
  int2* const s_cst = c.s_cst;
  const uint32_t tmem = c.tmem;
  const int head = c.head, tok0 = c.tok0, num_mt = c.num_mt, load_mt = c.load_mt, z = c.z;
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      const uint32_t stage_tx = (uint32_t)(load_mt * kTile16K + NB * kBlockKBytes);
      auto load_b = [&](int kb) {
        const int s = kb & 3;
        q_load_wuq<kSplit>(c, tm_wuq, tm_wuq_n, tm_wuq_r, kb, ptx::smem_u32(smem + s * kQStageBytes + 2 * kTile16K), ptx::smem_u32(full_bar + s));
      };
      auto load_a = [&](int kb) {  // one box = all token tiles of the stage (box rows = 128 * load_mt)
        const int s = kb & 3;
        ptx::tma_load_2d(ptx::smem_u32(smem + s * kQStageBytes), tm_a1, ptx::smem_u32(full_bar + s), kb * kBlockKBytes, tok0,
                         ptx::kEvictLast);
      };
      ptx::griddep_wait();  // a1 is produced by the norm kernel (the first W_UQ k-blocks were requested at kernel start)
      tl_mark(tl, 3);
      for (int kb = 0; kb < kQStages; ++kb) load_a(kb);
      for (int kb = kQStages; kb < kQKBlocks; ++kb) {
        const int s = kb & 3;
        ptx::mbar_wait(ptx::smem_u32(empty_bar + s), ((kb >> 2) & 1) ^ 1);
        ptx::mbar_arrive_expect_tx(ptx::smem_u32(full_bar + s), stage_tx);
        load_b(kb);
        load_a(kb);
      }
      // This CTA's Wuk chunk cl (columns 128 (z CH + cl) ..) -> bytes [32K*cl, 32K*cl+32K): needs the last use (3rd release,
      // parity 0) of the stages it overlaps
      auto load_wuk = [&](int cl) {
        const uint32_t fb = ptx::smem_u32(wuk_full + cl);
        ptx::mbar_arrive_expect_tx(fb, 2 * kTile16K);
        for (int at = 0; at < 2; ++at)
          ptx::tma_load_2d(ptx::smem_u32(smem + kQWukOff + cl * 32768 + at * kTile16K), tm_wuk, fb, (z * CH + cl) * 128 + at * 64,
                           head * MLAPO_Q_NOPE, ptx::kEvictFirst);
      };
      ptx::mbar_wait(ptx::smem_u32(empty_bar + 0), 0);
      load_wuk(0);                                   // [0, 32K)    in stage 0
      if constexpr (CH > 1) {
        ptx::mbar_wait(ptx::smem_u32(empty_bar + 1), 0);
        load_wuk(1);                                 // [32K, 64K)  in stages 0, 1
      }
      if constexpr (CH > 2) {
        load_wuk(2);                                 // [64K, 96K)  in stage 1
        ptx::mbar_wait(ptx::smem_u32(empty_bar + 2), 0);
        load_wuk(3);                                 // [96K, 128K) in stages 1, 2
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc2 = ptx::make_idesc(ptx::kCFmtS32, ptx::kFmtS8, ptx::kFmtS8, 0, 0, kBlockM, NB);
    constexpr uint32_t kAFmt = std::is_same<T, __nv_bfloat16>::value ? ptx::kFmtBF16 : ptx::kFmtF16;
    constexpr uint32_t idesc3 = ptx::make_idesc(ptx::kCFmtF32, kAFmt, kAFmt, 0, 1, kBlockM, 128);
    for (int kb = 0; kb < kQKBlocks; ++kb) {
      const int s = kb & 3;
      ptx::mbar_wait(ptx::smem_u32(full_bar + s), (kb >> 2) & 1);
      __syncwarp();
      ptx::tc_fence_after();
      if (lane == 0) {
        if (kb == 0) tl_mark(tl, 4);
        if (kb == kQKBlocks - 1) tl_mark(tl, 5);
      }
      {  // issued by the converged warp (one elected lane inside the asm blocks): see sm100_ptx.cuh, umma_i8_elect
        const uint32_t a_base = ptx::smem_u32(smem + s * kQStageBytes);
        const uint64_t da0 = ptx::make_smem_desc(a_base, 1024, 0);
        const uint64_t db0 = ptx::make_smem_desc(a_base + 2 * kTile16K, 1024, 0);
        for (int m = 0; m < num_mt; ++m) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            ptx::umma_i8_elect(tmem + m * 256, ptx::desc_advance(da0, m * kTile16K + k * kUmmaKBytes), ptx::desc_advance(db0, k * kUmmaKBytes), idesc2,
                               (kb > 0 || k > 0) ? 1u : 0u);
        }
        ptx::umma_commit_elect(ptx::smem_u32(empty_bar + s));
      }
      __syncwarp();
    }
    ptx::umma_commit_elect(ptx::smem_u32(mm2_done));
    __syncwarp();
    // absorb units (c, m): token tile m ping-pongs its own two TMEM buffers over this CTA's Wuk chunks
    for (int c = 0; c < CH; ++c) {
      for (int m = 0; m < num_mt; ++m) {
        const int ub = m * 2 + (c & 1);
        if (c == 0) ptx::mbar_wait(ptx::smem_u32(qn_ready + m), 0);   // buffer 0 = acc2[m] columns 0..127 (nope, drained)
        if (c == 1) ptx::mbar_wait(ptx::smem_u32(rope_done + m), 0);  // buffer 1 overlaps acc2[m] columns 128.. (rope)
        if (c >= 2) ptx::mbar_wait(ptx::smem_u32(buf_empty + ub), 0);
        if (m == 0) ptx::mbar_wait(ptx::smem_u32(wuk_full + c), 0);
        __syncwarp();
        ptx::tc_fence_after();
        {
          const uint64_t da0 = ptx::make_smem_desc(ptx::smem_u32(smem + kQNopeOff + m * 2 * kTile16K), 1024, 0);
          const uint64_t db0 = ptx::make_smem_desc(ptx::smem_u32(smem + kQWukOff + c * 32768), 1024, kTile16K);  // MN-major: SBO 8 k-rows, LBO next 64-col atom
#pragma unroll
          for (int k = 0; k < 8; ++k)  // 8 x UMMA_K(16) = 128
            ptx::umma_f16_elect(tmem + m * 256 + (c & 1) * 128, ptx::desc_advance(da0, (k >> 2) * kTile16K + (k & 3) * kUmmaKBytes),
                                ptx::desc_advance(db0, k * 2048), idesc3, k > 0 ? 1u : 0u);
          ptx::umma_commit_elect(ptx::smem_u32(unit_full + ub));
        }
        __syncwarp();
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue groups (128 threads each, thread = token row)
    const int g = (warp - 2) >> 2;           // group = token tile
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int et = threadIdx.x - 64 - g * 128;
    const bool issuer = et == 0;
    const int bar_id = 1 + g;
    const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
    const uint32_t st_tile = ptx::smem_u32(smem + kQStageOff + g * kTile16K);
    auto stage_acquire = [&]() {  // previous TMA store of this group has finished reading the staging tile
      if (issuer) ptx::tma_store_wait_read<0>();
      ptx::named_bar_sync(bar_id, 128);
    };
    auto stage_acquire_pp = [&]() {  // ping-pong variant: only the store before the most recent one must have been read
      if (issuer) ptx::tma_store_wait_read<1>();
      ptx::named_bar_sync(bar_id, 128);
    };
    auto stage_publish = [&]() {  // all rows written -> visible to the async proxy
      ptx::fence_proxy_async_smem();
      ptx::named_bar_sync(bar_id, 128);
    };

    if (g == 0) {
      for (int cc = et; cc < NB; cc += 128) {  // local MM2 column -> column of the head: nope | this CTA's pairs, low halves | high halves
        const int col = head * MLAPO_Q_HEAD + (cc < 128 ? cc : cc < 128 + G ? 128 + z * G + (cc - 128) : 160 + z * G + (cc - 128 - G));
        const float dsc = load_descale(qa.descale1, qa.descale_is_i64, col);
        s_cst[cc] = make_int2(qa.bias1 ? qa.bias1[col] : 0, __float_as_int(dsc));
      }
    }
    ptx::griddep_wait();  // tok_scale1 comes from the norm kernel (and keeps the PDL completion chain transitive)
    // dependents (the next call's quantise kernel, which may store to a0 without waiting) are signalled only now: the norm kernel,
    // and with it this call's MM1, the last reader of a0, are complete
    ptx::griddep_launch_dependents();
    ptx::named_bar_sync(3, 256);

    if (g < num_mt) {
      const int m = g;
      const bool per_token = qa.tok_scale1 != nullptr;
      const int t = tok0 + m * kBlockM + row;
      const int tc = t < qa.m_total ? t : qa.m_total - 1;  // clamped row for loads; TMA stores clip rows >= N
      const float ts = per_token ? qa.tok_scale1[tc] : 1.f;
      // y = fp32(acc + bias) * descale * token scale for columns c, c+1 (c even): the per-tensor mode carries a token scale of 1
      // (x * 1.0f is exact), the per-token mode a bias of 0, so one expression serves both; one 16-byte load brings both columns'
      // constants
      auto deq2 = [&](uint32_t acc0, uint32_t acc1, int c, float& f0, float& f1) {
        const int4 cst = *reinterpret_cast<const int4*>(s_cst + c);
        f0 = __fmul_rn(__fmul_rn(__int2float_rn((int)acc0 + cst.x), __int_as_float(cst.y)), ts);
        f1 = __fmul_rn(__fmul_rn(__int2float_rn((int)acc1 + cst.z), __int_as_float(cst.w)), ts);
      };
      // cos/sin of this token (issued before the MM2 wait; consumed after the nope slice): the whole rows, or — split forms — the
      // G values of this CTA's pairs and the G of their partners 32 columns on
      constexpr int CSV = kSplit == 1 ? 8 : 2 * G / 8;   // 16-byte vectors per table
      uint4 cs[CSV], sn[CSV];
      {
        const T* cp = reinterpret_cast<const T*>(qa.cos) + (long long)tc * 64;
        const T* sp = reinterpret_cast<const T*>(qa.sin) + (long long)tc * 64;
        // (32 bytes per request where the run allows: every row of a warp is a different line, so the number of requests is
        // what the load/store unit pays for — and the TMA issue and the staging stores queue behind it)
        if constexpr (kSplit == 1) {
#pragma unroll
          for (int j = 0; j < 8; j += 2) {
            ptx::ld_global_nc_v8(reinterpret_cast<const uint4*>(cp) + j, cs[j], cs[j + 1]);
            ptx::ld_global_nc_v8(reinterpret_cast<const uint4*>(sp) + j, sn[j], sn[j + 1]);
          }
        } else if constexpr (G == 16) {
          ptx::ld_global_nc_v8(cp + z * G, cs[0], cs[1]);
          ptx::ld_global_nc_v8(sp + z * G, sn[0], sn[1]);
          ptx::ld_global_nc_v8(cp + 32 + z * G, cs[2], cs[3]);
          ptx::ld_global_nc_v8(sp + 32 + z * G, sn[2], sn[3]);
        } else {
#pragma unroll
          for (int j = 0; j < G / 8; ++j) {
            cs[j] = ptx::ld_global_nc_v4(reinterpret_cast<const uint4*>(cp + z * G) + j);
            sn[j] = ptx::ld_global_nc_v4(reinterpret_cast<const uint4*>(sp + z * G) + j);
            cs[G / 8 + j] = ptx::ld_global_nc_v4(reinterpret_cast<const uint4*>(cp + 32 + z * G) + j);
            sn[G / 8 + j] = ptx::ld_global_nc_v4(reinterpret_cast<const uint4*>(sp + 32 + z * G) + j);
          }
        }
      }

      ptx::mbar_wait(ptx::smem_u32(mm2_done), 0);
      __syncwarp();
      ptx::tc_fence_after();
      if (issuer && g == 0) tl_mark(tl, 6);  // MM2 retired

      const uint32_t acc2 = lane_base + m * 256;
      // nope slice -> swizzled A-operand tiles (kb = c0/64).  The TMEM load of the next 32 columns is in flight while the
      // current 32 are dequantised (tcgen05.wait::ld covers every outstanding load, so exactly one is kept outstanding).
      auto nope_chunk = [&](const uint32_t (&r)[32], int c0) {
        const uint32_t tile = ptx::smem_u32(smem + kQNopeOff + (m * 2 + (c0 >> 6)) * kTile16K);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t w[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            float f0, f1;
            deq2(r[8 * j + 2 * u], r[8 * j + 2 * u + 1], c0 + 8 * j + 2 * u, f0, f1);
            w[u] = Act<T>::pack(f0, f1);
          }
          ptx::st_shared_v4(tile + ptx::sw128_offset(row, ((c0 & 63) >> 3) + j), w[0], w[1], w[2], w[3]);
        }
      };
      if constexpr (kSplit == 1) {
      float x[64];
      {
        uint32_t ra[32], rb[32];
        ptx::tmem_ld_32x32b_x32(acc2 + 0, ra);
        ptx::tmem_ld_wait();
        ptx::tmem_ld_32x32b_x32(acc2 + 32, rb);
        nope_chunk(ra, 0);
        ptx::tmem_ld_wait();
        ptx::tmem_ld_32x32b_x32(acc2 + 64, ra);
        nope_chunk(rb, 32);
        ptx::tmem_ld_wait();
        ptx::tmem_ld_32x32b_x32(acc2 + 96, rb);
        nope_chunk(ra, 64);
        ptx::tmem_ld_wait();
        ptx::tmem_ld_32x32b_x32(acc2 + 128, ra);  // rope slice: columns 128..191
        nope_chunk(rb, 96);
        // q_nope rows written and acc2[m] columns 0..127 read: the first absorb unit of this token tile may start
        ptx::fence_proxy_async_smem();
        ptx::tmem_ld_wait();
        ptx::tmem_ld_32x32b_x32(acc2 + 160, rb);
        ptx::tc_fence_before();
        ptx::mbar_arrive(ptx::smem_u32(qn_ready + m));
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          deq2(ra[j], ra[j + 1], 128 + j, x[j], x[j + 1]);
          x[j] = Act<T>::round(x[j]);
          x[j + 1] = Act<T>::round(x[j + 1]);
        }
        ptx::tmem_ld_wait();
        // acc2[m] fully read: the absorb buffer that overlaps the rope columns is free too
        ptx::tc_fence_before();
        ptx::mbar_arrive(ptx::smem_u32(rope_done + m));
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          deq2(rb[j], rb[j + 1], 160 + j, x[32 + j], x[33 + j]);
          x[32 + j] = Act<T>::round(x[32 + j]);
          x[33 + j] = Act<T>::round(x[33 + j]);
        }
      }

      stage_acquire();
#pragma unroll
      for (int gg = 0; gg < 4; ++gg) {
        float c_lo[8], s_lo[8], c_hi[8], s_hi[8];
        unpack8<T>(cs[gg], c_lo);
        unpack8<T>(sn[gg], s_lo);
        unpack8<T>(cs[gg + 4], c_hi);
        unpack8<T>(sn[gg + 4], s_hi);
        uint32_t wl[4], wh[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int i0 = gg * 8 + 2 * u, i1 = i0 + 1;
          // out[i] = x[i]*cos[i] + (-x[i+32])*sin[i];  out[i+32] = x[i+32]*cos[i+32] + x[i]*sin[i+32]   (test_mla:57-67)
          const float a0 = __fadd_rn(__fmul_rn(x[i0], c_lo[2 * u]), __fmul_rn(-x[i0 + 32], s_lo[2 * u]));
          const float a1 = __fadd_rn(__fmul_rn(x[i1], c_lo[2 * u + 1]), __fmul_rn(-x[i1 + 32], s_lo[2 * u + 1]));
          const float b0 = __fadd_rn(__fmul_rn(x[i0 + 32], c_hi[2 * u]), __fmul_rn(x[i0], s_hi[2 * u]));
          const float b1 = __fadd_rn(__fmul_rn(x[i1 + 32], c_hi[2 * u + 1]), __fmul_rn(x[i1], s_hi[2 * u + 1]));
          wl[u] = Act<T>::pack(a0, a1);
          wh[u] = Act<T>::pack(b0, b1);
        }
        ptx::st_shared_v4(st_tile + ptx::sw128_offset(row, gg), wl[0], wl[1], wl[2], wl[3]);
        ptx::st_shared_v4(st_tile + ptx::sw128_offset(row, gg + 4), wh[0], wh[1], wh[2], wh[3]);
      }
      stage_publish();
      if (issuer) {
        ptx::tma_store_2d(tm_out1, st_tile, head * MLAPO_ROPE, tok0 + m * kBlockM);
        ptx::tma_store_commit();
      }
      } else {
        // split forms: the rope slice of this CTA is columns [128, 128 + G) (pairs' low halves) and [128 + G, 128 + 2 G) (their
        // partners); its 2 x G results go to q_out1 with plain 16-byte stores (two short runs per row: no staging tile needed)
        float xl[G], xh[G];
        {
          uint32_t ra[32], rb[32];
          ptx::tmem_ld_32x32b_x32(acc2 + 0, ra);
          ptx::tmem_ld_wait();
          ptx::tmem_ld_32x32b_x32(acc2 + 32, rb);
          nope_chunk(ra, 0);
          ptx::tmem_ld_wait();
          ptx::tmem_ld_32x32b_x32(acc2 + 64, ra);
          nope_chunk(rb, 32);
          ptx::tmem_ld_wait();
          ptx::tmem_ld_32x32b_x32(acc2 + 96, rb);
          nope_chunk(ra, 64);
          ptx::tmem_ld_wait();
          uint32_t rr[2 * G];
          if constexpr (G == 16) ptx::tmem_ld_32x32b_x32(acc2 + 128, rr);
          else ptx::tmem_ld_32x32b_x16(acc2 + 128, rr);
          nope_chunk(rb, 96);
          ptx::fence_proxy_async_smem();
          ptx::tmem_ld_wait();
          // q_nope rows written and acc2[m] read completely: both absorb buffers of this token tile are free
          ptx::tc_fence_before();
          ptx::mbar_arrive(ptx::smem_u32(qn_ready + m));
          ptx::mbar_arrive(ptx::smem_u32(rope_done + m));
#pragma unroll
          for (int j = 0; j < G; j += 2) {
            deq2(rr[j], rr[j + 1], 128 + j, xl[j], xl[j + 1]);
            deq2(rr[G + j], rr[G + j + 1], 128 + G + j, xh[j], xh[j + 1]);
            xl[j] = Act<T>::round(xl[j]);
            xl[j + 1] = Act<T>::round(xl[j + 1]);
            xh[j] = Act<T>::round(xh[j]);
            xh[j + 1] = Act<T>::round(xh[j + 1]);
          }
        }
        if (t < qa.m_total) {
          T* const o_lo = reinterpret_cast<T*>(qa.q_out1) + ((long long)t * qa.num_heads + head) * MLAPO_ROPE + z * G;
          uint32_t wlo[G / 2], whi[G / 2];
#pragma unroll
          for (int v = 0; v < G / 8; ++v) {
            float c_lo[8], s_lo[8], c_hi[8], s_hi[8];
            unpack8<T>(cs[v], c_lo);
            unpack8<T>(sn[v], s_lo);
            unpack8<T>(cs[G / 8 + v], c_hi);
            unpack8<T>(sn[G / 8 + v], s_hi);
            uint32_t* const wl = wlo + 4 * v;
            uint32_t* const wh = whi + 4 * v;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int i0 = v * 8 + 2 * u, i1 = i0 + 1;
              // out[i] = x[i]*cos[i] + (-x[i+32])*sin[i];  out[i+32] = x[i+32]*cos[i+32] + x[i]*sin[i+32]   (test_mla:57-67)
              const float a0 = __fadd_rn(__fmul_rn(xl[i0], c_lo[2 * u]), __fmul_rn(-xh[i0], s_lo[2 * u]));
              const float a1 = __fadd_rn(__fmul_rn(xl[i1], c_lo[2 * u + 1]), __fmul_rn(-xh[i1], s_lo[2 * u + 1]));
              const float b0 = __fadd_rn(__fmul_rn(xh[i0], c_hi[2 * u]), __fmul_rn(xl[i0], s_hi[2 * u]));
              const float b1 = __fadd_rn(__fmul_rn(xh[i1], c_hi[2 * u + 1]), __fmul_rn(xl[i1], s_hi[2 * u + 1]));
              wl[u] = Act<T>::pack(a0, a1);
              wh[u] = Act<T>::pack(b0, b1);
            }
          }
          if constexpr (G == 16) {
            ptx::st_global_v8(o_lo, wlo[0], wlo[1], wlo[2], wlo[3], wlo[4], wlo[5], wlo[6], wlo[7]);
            ptx::st_global_v8(o_lo + 32, whi[0], whi[1], whi[2], whi[3], whi[4], whi[5], whi[6], whi[7]);
          } else {
            ptx::st_global_v4(o_lo, wlo[0], wlo[1], wlo[2], wlo[3]);
            ptx::st_global_v4(o_lo + 32, whi[0], whi[1], whi[2], whi[3]);
          }
        }
      }
      if (issuer && g == 0) tl_mark(tl, 7);  // epi2 done

      // absorb drains
      float qs = 1.f;
      if (kInt8Out) qs = Act<T>::to_f(reinterpret_cast<const T*>(qa.q_nope_scale)[head]);
#pragma unroll 1
      for (int c = 0; c < CH; ++c) {   // local chunk index; its output columns start at 128 * cg
        const int cg = z * CH + c;
        const int ub = m * 2 + (c & 1);
        ptx::mbar_wait(ptx::smem_u32(unit_full + ub), (c >> 1) & 1);
        __syncwarp();
        ptx::tc_fence_after();
        if (issuer && g == 0 && c == 0) tl_mark(tl, 8);  // first absorb unit ready
        if (issuer && g == 0 && c == CH - 1) tl_mark(tl, 9);
        const uint32_t acc3 = lane_base + m * 256 + (c & 1) * 128;
        if (kInt8Out) {
          // S9: int8( fp16( act(acc) * q_nope_scale[h] ) )   (golden test_mla:86-92, 466-470); 128 cols = one 128-byte row
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
          if (c + 2 < CH) ptx::mbar_arrive(ptx::smem_u32(buf_empty + ub));
          stage_acquire();
#pragma unroll
          for (int j = 0; j < 8; ++j) ptx::st_shared_v4(st_tile + ptx::sw128_offset(row, j), w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
          stage_publish();
          if (issuer) {
            ptx::tma_store_2d(tm_out0, st_tile, head * MLAPO_KV_LORA + cg * 128, tok0 + m * kBlockM);
            ptx::tma_store_commit();
          }
        } else {
#pragma unroll 1
          for (int half = 0; half < 2; ++half) {  // 64 act columns = one 128-byte row per token
            uint32_t w[32];
#pragma unroll
            for (int c0 = 0; c0 < 64; c0 += 32) {
              uint32_t r[32];
              ptx::tmem_ld_32x32b_x32(acc3 + half * 64 + c0, r);
              ptx::tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 16; ++j) w[(c0 >> 1) + j] = Act<T>::pack(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
            }
            if (half == 1) {
              ptx::tc_fence_before();
              if (c + 2 < CH) ptx::mbar_arrive(ptx::smem_u32(buf_empty + ub));
            }
            // unit_full(c) implies every MMA of chunk c-1 has retired, so Wuk slot c-1 is dead: its g-th 16 KB tile
            // becomes this group's second staging buffer (first half -> own tile, second half -> dead slot).
            uint32_t dst = st_tile;
            if (c >= 1 && half == 1) dst = ptx::smem_u32(smem + kQWukOff + (c - 1) * 32768 + g * kTile16K);
            if (c >= 2 || (c == 1 && half == 1)) stage_acquire_pp();  // the newest outstanding store reads the other buffer
            else stage_acquire();
#pragma unroll
            for (int j = 0; j < 8; ++j) ptx::st_shared_v4(dst + ptx::sw128_offset(row, j), w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
            stage_publish();
            if (issuer) {
              ptx::tma_store_2d(tm_out0, dst, head * MLAPO_KV_LORA + cg * 128 + half * 64, tok0 + m * kBlockM);
              ptx::tma_store_commit();
            }
          }
        }
      }
      if (issuer) {
        if (g == 0) tl_mark(tl, 10);
        ptx::tma_store_wait_read<0>();
        if (g == 0) tl_mark(tl, 11);
      }
    }
  }

}

template <class T, bool kInt8Out, int kSplit>
__global__ void __launch_bounds__(kQThreads, 1)
q_fused_kernel(const __grid_constant__ CUtensorMap tm_a1, const __grid_constant__ CUtensorMap tm_wuq,
               const __grid_constant__ CUtensorMap tm_wuq_n, const __grid_constant__ CUtensorMap tm_wuq_r,
               const __grid_constant__ CUtensorMap tm_wuk, const __grid_constant__ CUtensorMap tm_out0,
               const __grid_constant__ CUtensorMap tm_out1, QArgs qa, long long* tl) {
  if (threadIdx.x == 0) { tl_mark_global(tl, 0); tl_mark(tl, 1); }
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  QCtx c;
  c.smem = smem_raw;
  if ((ptx::smem_u32(smem_raw) & 1023u) != 0) __trap();  // the 128-byte swizzle needs 1024-byte aligned tiles
  c.bars = reinterpret_cast<uint64_t*>(c.smem + kQBarOff);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(c.bars + 26);
  c.s_cst = reinterpret_cast<int2*>(c.smem + kQBarOff + 256);
  c.set_work(blockIdx.x, blockIdx.y, blockIdx.z, qa);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {  // barriers + the first W_UQ k-blocks before anything else: they depend on nothing
    ptx::prefetch_tmap(kSplit == 1 ? &tm_wuq : &tm_wuq_n);
    q_init_barriers(c.bars);
    ptx::fence_mbar_init();
    q_issue_weights<kSplit>(c, &tm_wuq, &tm_wuq_n, &tm_wuq_r);
    ptx::prefetch_tmap(&tm_a1);
    ptx::prefetch_tmap(&tm_wuk);
    ptx::prefetch_tmap(&tm_out0);
    ptx::prefetch_tmap(&tm_out1);
  }
  if (warp == 1) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  c.tmem = *tmem_slot;
  if (threadIdx.x == 0) tl_mark(tl, 2);  // prologue done
  if (MLAPO_TL_ON && threadIdx.x == 0 && tl) {  // which SM ran this CTA (timeline analysis only)
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    tl[(blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z)) * 16 + 13] = smid;
  }

  q_roles<T, kInt8Out, kSplit>(c, &tm_a1, &tm_wuq, &tm_wuq_n, &tm_wuq_r, &tm_wuk, &tm_out0, &tm_out1, qa, warp, lane, tl);

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(c.tmem, 512);
  if (threadIdx.x == 32) { tl_mark(tl, 12); tl_mark_global(tl, 15); }
}


}  // namespace mlapo
