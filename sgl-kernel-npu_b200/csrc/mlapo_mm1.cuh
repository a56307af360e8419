// k2 MM1: a0 . Wdqkv^T as tcgen05 kind::i8: split-K partial sums (decode) or the dequantised rows y0 straight from the epilogue (prefill).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "../../include/mla_preprocess_b200.h"
#include "mlapo_numerics.cuh"

namespace mlapo {

// ============================================================================================ k2: MM1 (W_DQKV), split-K
// acc[MT*128 tokens x 128] (int32, TMEM) = a0[tokens x Kslice] (int8, K-major) . Wdqkv[128 rows x Kslice]^T (int8, K-major).
// grid = (n_tiles=17, splits, token super-tiles of MT*128).  One CTA streams its weight tile once for all MT token tiles.
// 10 warps: w0 = TMA producer, w1 = tcgen05.mma issuer (+TMEM alloc), w2..w9 = two epilogue groups (TMEM lane quarter = warp%4).
// Weight (B) tiles do not depend on the upstream kernel, so they are requested before griddepcontrol.wait (PDL).
constexpr int kGemmThreads = 320;
constexpr int kBlockM = 128;
constexpr int kBlockKBytes = 128;  // one 128B swizzle atom along K
constexpr int kUmmaKBytes = 32;    // K per tcgen05.mma for 8-bit (32 elems) and 16-bit (16 elems) operands
constexpr int kTile16K = 16384;    // [128 rows x 128 B] swizzled tile
constexpr int kMm1BlockN = 128;   // decode: 17 weight tiles x up to 8 K-splits fill one wave
constexpr int kMm1BlockNWide = 256;  // prefill (no split-K): 256 x 256 CTA tiles halve the activation re-reads per weight byte

// (Seven stages for the one-token-tile form — its whole weight slice requested before the wait — were tried: C1 17.75 -> 17.45 us,
// but C5 24.2 -> 26.0 us: at 230 KB the CTA leaves no room for the norm kernel's CTAs to start under it.)
template <int MT, int BN>
__host__ __device__ constexpr int mm1_stages() { return (MT * kTile16K + BN * kBlockKBytes) > 49152 ? 3 : 4; }
template <int MT, int BN, bool kDequant = false>
constexpr int mm1_smem_bytes() {
  return mm1_stages<MT, BN>() * (MT * kTile16K + BN * kBlockKBytes) + 256 + (kDequant ? BN * 8 : 0) + 1024;
}

// The dequantising epilogue's inputs (prefill: no split-K, so the accumulators are complete when the MMAs retire):
// y0 = act_rne(fp32(acc + bias0[n]) * descale0[n]) (per-tensor) or act_rne(fp32(acc) * descale0[n] * tokScale0[m]) (per-token)
// — PpMatmulW8a8Aiv, mla_preprocess_mix_bf16.hpp:2204-2330, which the reference also runs tile by tile behind the matmul.
struct Mm1DequantArgs {
  const int32_t* bias0;     // [2112] or null (per-token mode)
  const void* descale0;     // f32 | int64 [2112]
  const float* tok_scale0;  // [N] or null
  int descale_is_i64;
  int tap_acc;              // debug: also store the raw int32 accumulators through the partial-sum tensor map
};

// Where one CTA's MM1 work item lives: its k-range, its tiles, its pipeline in shared memory and its TMEM columns.
template <int MT, int BN>
struct Mm1Ctx {
  static constexpr int STAGES = mm1_stages<MT, BN>();
  static constexpr int A_BYTES = MT * kTile16K;
  static constexpr int B_BYTES = BN * kBlockKBytes;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int TMEM_COLS = MT * BN;   // 128, 256 or 512
  static constexpr int GRANULES = BN / 32;    // 32 int32 columns = one 128-byte row per token
  static constexpr int NUM_BARS = 2 * STAGES + 1;
  static_assert(TMEM_COLS <= 512 && (TMEM_COLS & (TMEM_COLS - 1)) == 0, "TMEM allocation");
  static_assert(GRANULES * kTile16K <= STAGES * STAGE_BYTES, "one token tile of epilogue staging must fit in the drained pipeline");
  uint8_t* smem;        // STAGES x [A: MT x 16 KB | B: BN x 128 B], 1024-byte aligned
  uint64_t* bars;       // full[STAGES], empty[STAGES], done
  uint32_t tmem_acc;    // first TMEM column of the accumulators
  int n0, split, m0, kb0, nkb, mt_valid, m_total;

  __device__ void set_work(int n_tile, int split_, int splits, int m_super, int m_total_, int kb_total) {
    n0 = n_tile * BN;
    split = split_;
    m0 = m_super * (MT * kBlockM);
    m_total = m_total_;
    kb0 = (int)(((long long)split_ * kb_total) / splits);
    nkb = (int)(((long long)(split_ + 1) * kb_total) / splits) - kb0;
    // token tiles that hold at least one valid row (the others are neither loaded nor multiplied)
    const int tiles = (m_total_ - m0 + kBlockM - 1) / kBlockM;
    mt_valid = tiles < MT ? tiles : MT;
  }
};

template <int MT, int BN>
__device__ __forceinline__ void mm1_init_barriers(uint64_t* bars) {  // one thread
  for (int i = 0; i < Mm1Ctx<MT, BN>::NUM_BARS; ++i) ptx::mbar_init(ptx::smem_u32(bars + i), 1);
}

// One thread: arm the first pipeline stages and request their weight tiles (which depend on nothing upstream).
template <int MT, int BN>
__device__ __forceinline__ void mm1_issue_weights(const Mm1Ctx<MT, BN>& c, const CUtensorMap* tm_b) {
  using C = Mm1Ctx<MT, BN>;
  const uint32_t stage_tx = (uint32_t)(c.mt_valid * kTile16K + C::B_BYTES);
  const int pre = c.nkb < C::STAGES ? c.nkb : C::STAGES;
  for (int i = 0; i < pre; ++i) {
    const uint32_t fb = ptx::smem_u32(c.bars + i);
    ptx::mbar_arrive_expect_tx(fb, stage_tx);
    ptx::tma_load_2d(ptx::smem_u32(c.smem + i * C::STAGE_BYTES + C::A_BYTES), tm_b, fb, (c.kb0 + i) * kBlockKBytes, c.n0, ptx::kEvictFirst);
  }
}

// Warp roles of MM1: warp 0 = TMA producer, warp 1 = tcgen05.mma issuer, warps 2..9 = epilogue (TMEM lane quarter = warp%4);
// other warps fall through.  The caller has already issued mm1_issue_weights(); under programmatic dependent launch the producer
// waits for the upstream kernel before it touches a0.
template <int MT, int BN, bool kDequant = false, class T = __nv_bfloat16>
__device__ __forceinline__ void mm1_roles(const Mm1Ctx<MT, BN>& c, const CUtensorMap* tm_a, const CUtensorMap* tm_b,
                                          const CUtensorMap* tm_p, int warp, int lane, long long* tl,
                                          const CUtensorMap* tm_y = nullptr, const int2* s_cst = nullptr,
                                          const Mm1DequantArgs* dq = nullptr) {
  using C = Mm1Ctx<MT, BN>;
  constexpr int STAGES = C::STAGES, A_BYTES = C::A_BYTES, B_BYTES = C::B_BYTES, STAGE_BYTES = C::STAGE_BYTES, GRANULES = C::GRANULES;
  uint8_t* const smem = c.smem;
  uint64_t* const full_bar = c.bars;
  uint64_t* const empty_bar = c.bars + STAGES;
  uint64_t* const done_bar = c.bars + 2 * STAGES;
  const uint32_t tmem_acc = c.tmem_acc;
  const int n0 = c.n0, split = c.split, m0 = c.m0, kb0 = c.kb0, nkb = c.nkb, mt_valid = c.mt_valid;
  (void)B_BYTES;
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      const uint32_t stage_tx = (uint32_t)(mt_valid * kTile16K + B_BYTES);
      const int pre = nkb < STAGES ? nkb : STAGES;
      ptx::griddep_wait();  // a0 is produced by the quantise kernel (the weight tiles were requested at kernel start)
      tl_mark(tl, 3);       // upstream complete
      for (int i = 0; i < pre; ++i) {
        const uint32_t fb = ptx::smem_u32(full_bar + i);
        for (int m = 0; m < mt_valid; ++m)
          ptx::tma_load_2d(ptx::smem_u32(smem + i * STAGE_BYTES + m * kTile16K), tm_a, fb, (kb0 + i) * kBlockKBytes,
                           m0 + m * kBlockM, ptx::kEvictLast);
      }
      uint32_t stage = 0, phase = 1;  // stages 0..pre-1 are in their first use; the next wrap waits for phase 0 of empty
      for (int i = pre; i < nkb; ++i) {
        ptx::mbar_wait(ptx::smem_u32(empty_bar + stage), phase ^ 1);
        const uint32_t fb = ptx::smem_u32(full_bar + stage);
        ptx::mbar_arrive_expect_tx(fb, stage_tx);
        ptx::tma_load_2d(ptx::smem_u32(smem + stage * STAGE_BYTES + A_BYTES), tm_b, fb, (kb0 + i) * kBlockKBytes, n0, ptx::kEvictFirst);
        for (int m = 0; m < mt_valid; ++m)
          ptx::tma_load_2d(ptx::smem_u32(smem + stage * STAGE_BYTES + m * kTile16K), tm_a, fb, (kb0 + i) * kBlockKBytes,
                           m0 + m * kBlockM, ptx::kEvictLast);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (the converged warp; one elected lane issues)
    constexpr uint32_t idesc = ptx::make_idesc(ptx::kCFmtS32, ptx::kFmtS8, ptx::kFmtS8, 0, 0, kBlockM, BN);
    uint32_t stage = 0, phase = 0;
    for (int i = 0; i < nkb; ++i) {
      ptx::mbar_wait(ptx::smem_u32(full_bar + stage), phase);
      __syncwarp();
      ptx::tc_fence_after();
      if (lane == 0) {
        if (i == 0) tl_mark(tl, 4);             // first stage landed
        if (i == nkb - 1) tl_mark(tl, 5);       // last stage landed
      }
      {
        const uint32_t a_base = ptx::smem_u32(smem + stage * STAGE_BYTES);
        const uint64_t da0 = ptx::make_smem_desc(a_base, 1024, 0);
        const uint64_t db0 = ptx::make_smem_desc(a_base + A_BYTES, 1024, 0);
        for (int m = 0; m < mt_valid; ++m) {
#pragma unroll
          for (int k = 0; k < kBlockKBytes / kUmmaKBytes; ++k)
            ptx::umma_i8_elect(tmem_acc + m * BN, ptx::desc_advance(da0, m * kTile16K + k * kUmmaKBytes), ptx::desc_advance(db0, k * kUmmaKBytes),
                               idesc, (i > 0 || k > 0) ? 1u : 0u);
        }
        ptx::umma_commit_elect(ptx::smem_u32(empty_bar + stage));  // frees the smem slot when these MMAs retire
      }
      __syncwarp();
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
    ptx::umma_commit_elect(ptx::smem_u32(done_bar));
    __syncwarp();
  } else if (warp < 10) {
    // ------------------------------------------------------------------ epilogue: TMEM -> swizzled smem -> TMA store
    // Two groups of four warps (a warp reads the TMEM lane quarter warp % 4).  With two valid token tiles each group drains one;
    // with one, each drains half of its column granules.  Every granule tile is handed to the TMA store engine as soon as its 128
    // rows are written, so the stores of one granule run under the TMEM reads of the next.
    const int quarter = warp & 3, group = (warp - 2) >> 2;
    const int row = quarter * 32 + lane;
    const uint32_t t_base = tmem_acc + ((uint32_t)(quarter * 32) << 16);
    const bool issuer = threadIdx.x == 64 + group * 128;
    const int bar_id = 1 + group;
    constexpr int kGroupBytes = STAGES * STAGE_BYTES / 2 / kTile16K * kTile16K;  // each group stages in its half of the drained pipeline
    constexpr int TPG = kGroupBytes / kTile16K;                                    // staging tiles per group
    uint8_t* const stg = smem + group * kGroupBytes;
    ptx::mbar_wait(ptx::smem_u32(done_bar), 0);   // all MMAs retired => every pipeline stage has been consumed
    __syncwarp();
    ptx::tc_fence_after();
    if (issuer && group == 0) tl_mark(tl, 6);  // MMAs retired
    __syncwarp();                                 // tcgen05.ld below is warp-aligned: reconverge after the one-lane branch
    const bool two = mt_valid > 1;
    const int m_lo = two ? group : 0, m_hi = two ? group + 1 : 1;
    int used = 0;  // staging tiles this group has handed to the store engine so far
    auto tile_acquire = [&]() -> uint8_t* {   // the tile's previous store (TPG tiles ago) has been read out of shared memory
      if (used >= TPG) {
        if (issuer) ptx::tma_store_wait_read<TPG - 1>();
        ptx::named_bar_sync(bar_id, 128);
      }
      return stg + (used % TPG) * kTile16K;
    };
    bool raw = true;  // store the int32 accumulators (always in the split-K form; only as a debug tap in the dequantising form)
    if constexpr (kDequant) raw = dq->tap_acc != 0;
    if (raw) {
      const int g_lo = two ? 0 : group * (GRANULES / 2), g_hi = two ? GRANULES : (group + 1) * (GRANULES / 2);
      for (int m = m_lo; m < m_hi; ++m) {
#pragma unroll 1
        for (int g = g_lo; g < g_hi; ++g) {
          uint32_t r[32];
          ptx::tmem_ld_32x32b_x32(t_base + m * BN + g * 32, r);
          uint8_t* const tile_p = tile_acquire();
          ptx::tmem_ld_wait();
          const uint32_t tile = ptx::smem_u32(tile_p);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            ptx::st_shared_v4(tile + ptx::sw128_offset(row, j), r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
          ptx::fence_proxy_async_smem();
          ptx::named_bar_sync(bar_id, 128);
          if (issuer) {
            if (n0 + g * 32 < MLAPO_MM1_OUT) ptx::tma_store_3d(tm_p, tile, n0 + g * 32, m0 + m * kBlockM, split);
            ptx::tma_store_commit();
          }
          ++used;
        }
      }
    }
    if constexpr (kDequant) {
      // ---- dequantise in the epilogue: TMEM -> (+bias) * descale -> act -> swizzled smem -> TMA store into y0 [N, 2112]
      constexpr int G64 = BN / 64;  // 64 act columns = one 128-byte row per token
      const int g_lo = two ? 0 : group * (G64 / 2), g_hi = two ? G64 : (group + 1) * (G64 / 2);
      for (int m = m_lo; m < m_hi; ++m) {
        const int t = m0 + m * kBlockM + row;
        const float ts = dq->tok_scale0 ? dq->tok_scale0[t < c.m_total ? t : c.m_total - 1] : 1.f;
#pragma unroll 1
        for (int g = g_lo; g < g_hi; ++g) {
          uint32_t ra[32], rb[32];
          ptx::tmem_ld_32x32b_x32(t_base + m * BN + g * 64, ra);
          ptx::tmem_ld_32x32b_x32(t_base + m * BN + g * 64 + 32, rb);
          uint8_t* const tile_p = tile_acquire();
          ptx::tmem_ld_wait();
          uint32_t w[32];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            // one expression for both modes: per-tensor carries a token scale of 1 (x * 1.0f is exact), per-token a bias of 0
            const int4 ca = *reinterpret_cast<const int4*>(s_cst + g * 64 + 2 * j);
            const int4 cb = *reinterpret_cast<const int4*>(s_cst + g * 64 + 32 + 2 * j);
            const float a0 = __fmul_rn(__fmul_rn(__int2float_rn((int)ra[2 * j] + ca.x), __int_as_float(ca.y)), ts);
            const float a1 = __fmul_rn(__fmul_rn(__int2float_rn((int)ra[2 * j + 1] + ca.z), __int_as_float(ca.w)), ts);
            const float b0 = __fmul_rn(__fmul_rn(__int2float_rn((int)rb[2 * j] + cb.x), __int_as_float(cb.y)), ts);
            const float b1 = __fmul_rn(__fmul_rn(__int2float_rn((int)rb[2 * j + 1] + cb.z), __int_as_float(cb.w)), ts);
            w[j] = Act<T>::pack(a0, a1);
            w[16 + j] = Act<T>::pack(b0, b1);
          }
          const uint32_t tile = ptx::smem_u32(tile_p);
#pragma unroll
          for (int j = 0; j < 8; ++j) ptx::st_shared_v4(tile + ptx::sw128_offset(row, j), w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
          ptx::fence_proxy_async_smem();
          ptx::named_bar_sync(bar_id, 128);
          if (issuer) {
            if (n0 + g * 64 < MLAPO_MM1_OUT) ptx::tma_store_2d(tm_y, tile, n0 + g * 64, m0 + m * kBlockM);
            ptx::tma_store_commit();
          }
          ++used;
        }
      }
    }
    if (issuer) {
      if (group == 0) tl_mark(tl, 7);
      ptx::tma_store_wait_read<0>();
      if (group == 0) tl_mark(tl, 8);
    }
  }

}

// The same kernel with the dequantising epilogue (no split-K: grid.y == 1): writes y0 act [N, 2112] instead of int32 sums.
template <int MT, int BN, class T>
__global__ void __launch_bounds__(kGemmThreads, 1)
mm1_dequant_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                   const __grid_constant__ CUtensorMap tm_p, const __grid_constant__ CUtensorMap tm_y, Mm1DequantArgs dq,
                   int m_total, int kb_total, long long* tl) {
  using C = Mm1Ctx<MT, BN>;
  if (threadIdx.x == 0) { tl_mark_global(tl, 0); tl_mark(tl, 1); }
  extern __shared__ uint8_t smem_raw[];
  C c;
  c.smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  c.bars = reinterpret_cast<uint64_t*>(c.smem + C::STAGES * C::STAGE_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(c.bars + C::NUM_BARS + 1);
  int2* s_cst = reinterpret_cast<int2*>(c.smem + C::STAGES * C::STAGE_BYTES + 256);  // per column {bias, descale bits}
  c.set_work(blockIdx.x, 0, 1, blockIdx.z, m_total, kb_total);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {  // barriers + the first weight tiles before anything else: they depend on nothing
    ptx::prefetch_tmap(&tm_b);
    mm1_init_barriers<MT, BN>(c.bars);
    ptx::fence_mbar_init();
    mm1_issue_weights<MT, BN>(c, &tm_b);
    ptx::prefetch_tmap(&tm_a);
    ptx::prefetch_tmap(&tm_y);
    ptx::prefetch_tmap(&tm_p);
  }
  if (warp == 1) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), C::TMEM_COLS);
    ptx::tmem_relinquish();
  }
  for (int i = threadIdx.x; i < BN; i += kGemmThreads) {  // weight-like parameters: complete long before the call
    const int col = c.n0 + i;
    int2 v = make_int2(0, 0);
    if (col < MLAPO_MM1_OUT) v = make_int2(dq.bias0 ? dq.bias0[col] : 0, __float_as_int(load_descale(dq.descale0, dq.descale_is_i64, col)));
    s_cst[i] = v;
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  c.tmem_acc = *tmem_slot;
  ptx::griddep_launch_dependents();
  if (threadIdx.x == 0) tl_mark(tl, 2);  // prologue done

  mm1_roles<MT, BN, true, T>(c, &tm_a, &tm_b, &tm_p, warp, lane, tl, &tm_y, s_cst, &dq);

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(c.tmem_acc, C::TMEM_COLS);
  if (threadIdx.x == 32) { tl_mark(tl, 9); tl_mark_global(tl, 15); }
}

template <int MT, int BN>
__global__ void __launch_bounds__(kGemmThreads, 1)
mm1_splitk_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                  const __grid_constant__ CUtensorMap tm_p, int m_total, int kb_total, long long* tl) {
  using C = Mm1Ctx<MT, BN>;
  if (threadIdx.x == 0) { tl_mark_global(tl, 0); tl_mark(tl, 1); }
  extern __shared__ uint8_t smem_raw[];
  C c;
  c.smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  c.bars = reinterpret_cast<uint64_t*>(c.smem + C::STAGES * C::STAGE_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(c.bars + C::NUM_BARS + 1);
  c.set_work(blockIdx.x, blockIdx.y, gridDim.y, blockIdx.z, m_total, kb_total);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {  // barriers + the first weight tiles before anything else: they depend on nothing
    ptx::prefetch_tmap(&tm_b);
    mm1_init_barriers<MT, BN>(c.bars);
    ptx::fence_mbar_init();
    mm1_issue_weights<MT, BN>(c, &tm_b);
    ptx::prefetch_tmap(&tm_a);
    ptx::prefetch_tmap(&tm_p);
  }
  if (warp == 1) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), C::TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  c.tmem_acc = *tmem_slot;
  ptx::griddep_launch_dependents();
  if (threadIdx.x == 0) tl_mark(tl, 2);  // prologue done

  mm1_roles<MT, BN>(c, &tm_a, &tm_b, &tm_p, warp, lane, tl);

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(c.tmem_acc, C::TMEM_COLS);
  if (threadIdx.x == 32) { tl_mark(tl, 9); tl_mark_global(tl, 15); }
}


}  // namespace mlapo
