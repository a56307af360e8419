// k5 the whole chain as one persistent cooperative kernel for decode-sized calls (opt-in).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <type_traits>

#include "../../include/mla_preprocess_b200.h"
#include "mlapo_quant_norm.cuh"
#include "mlapo_mm1.cuh"
#include "mlapo_q.cuh"

namespace mlapo {

// ============================================================================================ k5: decode — the whole chain in one persistent kernel
// For decode-sized calls (<= 256 tokens, every phase fits one wave) the four phases run inside ONE cooperative kernel with
// one CTA per SM (10 warps, 224 KB smem, 512 TMEM columns):
//   quantise hidden -> | grid barrier | -> MM1 split-K -> | grid barrier | -> norm / RoPE / cache scatter -> | grid barrier | -> q path
// Shared memory is one region that is the MM1 pipeline first and the q-path pipeline afterwards, so the TMA producer runs
// ahead of the barriers: W_DQKV tiles are requested before the quantise phase and the first four W_UQ k-blocks as soon as
// the MM1 pipeline has drained (they land while the norm phase runs).  Intermediates (a0, partial sums, a1) stay in L2.
struct DecodeArgs {
  const void* hidden;          // act [N, D]
  int8_t* a0;                  // int8 [N, D]
  const void* quant_scale0;    // act [1]   (per-tensor)
  const int8_t* quant_offset0; // int8 [1]
  float* tok_scale0;           // f32 [N]   (per-token) or null
  int hidden_dim;
  int n_tokens;
  int splits0;                 // MM1 K-splits; 17 * splits0 CTAs take part in MM1
  int q_tok_tiles;             // token tiles of the q grid; num_heads * q_tok_tiles CTAs take part in the q path
  NormArgs norm;
  QArgs q;
  long long* tl_mm1;           // debug timelines or null
  long long* tl_q;
  int stop_after;              // debug: leave after this phase (0 = run everything)
};
constexpr int kMm1NTiles = (MLAPO_MM1_OUT + kMm1BlockN - 1) / kMm1BlockN;  // 17
constexpr int kDecodeMm1BarOff = kQBarOff + 256;  // MM1's barriers share bytes with the q path's per-column constants (disjoint phases)
constexpr int kDecodeSmemBytes = kQSmemBytes;
constexpr int kDecodeStaticSmem = 256;            // reduction scratch of the quant / norm phases (static __shared__)
static_assert(kDecodeSmemBytes + kDecodeStaticSmem <= 232448, "persistent decode smem");

// Grid barrier: three arrival counters, one per barrier of the kernel, each on its own 128-byte line.  A barrier costs one
// returning atomic (release) and a poll of the same word (relaxed loads, one acquire fence at the end): two L2 round trips.
// The counters clean up after themselves — the last CTA to arrive at barrier k zeroes the counter of barrier k-1, which every CTA
// has left by then — so they are valid from module load on and from launch to launch; one set per device is enough because
// cooperative launches of this kernel on one device cannot overlap (each needs every SM).
// (The first version kept {arrivals, generation} on one line and bracketed them with three full membar.gl: 4-5 us per barrier.)
constexpr int kGridBarriers = 3;
__device__ unsigned g_grid_ctr[kGridBarriers * 32];

__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// All CTAs of the (co-resident) grid.  Global writes made before it by any thread — through the generic proxy or by
// completed TMA stores — are visible afterwards to generic loads and to TMA loads of every CTA.
__device__ __forceinline__ void grid_sync(int k, unsigned nblocks) {
  fence_proxy_async_all();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned* ctr = g_grid_ctr + k * 32;
    unsigned old;
    asm volatile("atom.add.release.gpu.global.u32 %0, [%1], 1;" : "=r"(old) : "l"(ctr) : "memory");
    if (old == nblocks - 1) {
      unsigned* prev = g_grid_ctr + ((k + kGridBarriers - 1) % kGridBarriers) * 32;
      asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(prev), "r"(0u) : "memory");
    } else {
      unsigned v;
      unsigned n = 0;
      do {
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
        if (++n > (1u << 24)) ptx::wait_timeout(blockIdx.x, 0xFFFFu, (unsigned)k, v);  // a missing participant becomes a launch failure, not a hang
      } while (v != nblocks);
    }
    asm volatile("fence.acq_rel.gpu;" ::: "memory");
  }
  __syncthreads();
  fence_proxy_async_all();
}

template <class T, bool kInt8Out, int MT>
__global__ void __launch_bounds__(kQThreads, 1)
mla_decode_persistent_kernel(const __grid_constant__ CUtensorMap tm_a0, const __grid_constant__ CUtensorMap tm_wdqkv,
                             const __grid_constant__ CUtensorMap tm_part, const __grid_constant__ CUtensorMap tm_a1,
                             const __grid_constant__ CUtensorMap tm_wuq, const __grid_constant__ CUtensorMap tm_wuk,
                             const __grid_constant__ CUtensorMap tm_out0, const __grid_constant__ CUtensorMap tm_out1, DecodeArgs da) {
  using M1 = Mm1Ctx<MT, kMm1BlockN>;
  static_assert(M1::STAGES * M1::STAGE_BYTES <= kQBarOff, "MM1 pipeline must fit under the q-path barrier block");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cta = blockIdx.x;
  const unsigned nblocks = gridDim.x;
  if (threadIdx.x == 0) { tl_mark_global(da.tl_mm1, 0); tl_mark(da.tl_mm1, 1); }

  QCtx qc;
  qc.smem = smem;
  qc.bars = reinterpret_cast<uint64_t*>(smem + kQBarOff);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(qc.bars + 26);
  qc.s_cst = reinterpret_cast<int2*>(smem + kQBarOff + 256);
  M1 mc;
  mc.smem = smem;
  mc.bars = reinterpret_cast<uint64_t*>(smem + kDecodeMm1BarOff);

  const bool has_mm1 = cta < kMm1NTiles * da.splits0;
  const bool has_q = cta < da.q.num_heads * da.q_tok_tiles;
  if (has_mm1) mc.set_work(cta % kMm1NTiles, cta / kMm1NTiles, da.splits0, 0, da.n_tokens, da.hidden_dim / kBlockKBytes);
  if (has_q) qc.set_work(cta % da.q.num_heads, cta / da.q.num_heads, da.q);

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_a0);
    ptx::prefetch_tmap(&tm_wdqkv);
    ptx::prefetch_tmap(&tm_part);
    ptx::prefetch_tmap(&tm_a1);
    ptx::prefetch_tmap(&tm_wuq);
    ptx::prefetch_tmap(&tm_wuk);
    ptx::prefetch_tmap(&tm_out0);
    ptx::prefetch_tmap(&tm_out1);
  }
  if (warp == 1) {
    if (lane == 0) {
      mm1_init_barriers<MT, kMm1BlockN>(mc.bars);
      q_init_barriers(qc.bars);
      ptx::fence_mbar_init();
    }
    __syncwarp();
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  mc.tmem_acc = qc.tmem = *tmem_slot;

  // W_DQKV tiles of this CTA's MM1 work item: requested now, consumed after the first grid barrier
  if (warp == 0 && lane == 0 && has_mm1) mm1_issue_weights<MT, kMm1BlockN>(mc, &tm_wdqkv);

  // ---- phase 1: quantise hiddenState -> a0
  if (da.tok_scale0) {
    for (int row = cta; row < da.n_tokens; row += (int)nblocks)
      quant_row_per_token<T>(reinterpret_cast<const T*>(da.hidden), da.a0, da.tok_scale0, da.hidden_dim, row);
  } else {
    QuantScale qs;
    qs.load(reinterpret_cast<const T*>(da.quant_scale0), da.quant_offset0);
    quant_vectors<T, 4>(reinterpret_cast<const T*>(da.hidden), da.a0, qs, (long long)da.n_tokens * da.hidden_dim / 8,
                        (long long)cta * blockDim.x + threadIdx.x, (long long)nblocks * blockDim.x);
  }
  if (threadIdx.x == 0) tl_mark(da.tl_mm1, 10);  // quant done, entering barrier 1
  grid_sync(0, nblocks);

  // ---- phase 2: MM1 split-K partial sums
  if (threadIdx.x == 0) tl_mark(da.tl_mm1, 2);
  if (has_mm1 && da.stop_after != 1) mm1_roles<MT, kMm1BlockN, false>(mc, &tm_a0, &tm_wdqkv, &tm_part, warp, lane, da.tl_mm1);
  if (threadIdx.x == 64 && has_mm1) ptx::tma_store_wait<0>();  // partial tiles fully written, not just read out of smem
  ptx::tc_fence_before();
  __syncthreads();                                              // the MM1 pipeline bytes are free again
  ptx::tc_fence_after();
  if (threadIdx.x == 32) tl_mark(da.tl_mm1, 9);
  // first W_UQ k-blocks of this CTA's head: requested now, they land while the norm phase runs
  if (warp == 0 && lane == 0 && has_q && (da.stop_after == 0 || da.stop_after > 3)) q_issue_weights(qc, &tm_wuq);
  grid_sync(1, nblocks);
  if (threadIdx.x == 0) tl_mark(da.tl_mm1, 11);  // barrier 2 passed
  const bool run3 = da.stop_after == 0 || da.stop_after >= 3;

  // ---- phase 3: norms, quant, RoPE, cache scatter (one token at a time per CTA)
  for (int t = cta; run3 && t < da.n_tokens; t += (int)nblocks) {
    norm_token<T, 8>(da.norm, t, threadIdx.x);
    __syncthreads();  // the reduction scratch is re-used by the next token
  }
  if (threadIdx.x == 0) tl_mark(da.tl_mm1, 12);  // norm done, entering barrier 3
  grid_sync(2, nblocks);

  // ---- phase 4: per-head q path
  if (threadIdx.x == 0) { tl_mark_global(da.tl_q, 0); tl_mark(da.tl_q, 1); tl_mark(da.tl_q, 2); }
  if (has_q && da.stop_after == 0) q_roles<T, kInt8Out, false>(qc, &tm_a1, &tm_wuq, &tm_wuk, &tm_out0, &tm_out1, da.q, warp, lane, da.tl_q);

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(qc.tmem, 512);
  if (threadIdx.x == 32) { tl_mark(da.tl_q, 12); tl_mark_global(da.tl_q, 15); tl_mark_global(da.tl_mm1, 15); }
}


}  // namespace mlapo
