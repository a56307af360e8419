// mla_preprocess for B200 (sm_100a): hand-written CUDA kernels + the C-ABI launcher.
//
// Path (reference phases P1..P6, csrc/mla_preprocess/op_kernel/mla_preprocess_mix_bf16.hpp:2773-2916):
//   k1 quant_hidden        hidden[N,D] act -> a0 int8                       (Quant, bf16.hpp:283-534)
//   k2 mm1_splitk          a0 . Wdqkv^T -> split-K int32 partial sums       (PpMatmulW8a8Aic, bf16.hpp:1515-1975)
//   k3 norm_rope_cache     sum partials, +bias, *descale -> act y0;
//                          q_a RMSNorm(+beta)+quant -> a1 int8              (RmsNormQuant, bf16.hpp:536-840)
//                          latent RMSNorm(512), RoPE(64), paged-cache scatter (RmsNormAndRopeConvergence1, :2499-2686)
//   k4 q_fused (per head)  a1 . Wuq[h]^T, dequant, RoPE -> q_out1; nope slice stays in smem and is absorbed:
//                          q_nope . Wuk[h] -> q_out0 (+ int8 re-quant)      (PpMatmulW8a8Aiv, RopeFp16, PpMatmulEinSum, EinSumQuant)
//   k5 decode_persistent   k1..k4 as phases of ONE cooperative kernel with grid barriers, for calls of <= 256 tokens
// As separate kernels (prefill sizes) k1..k4 are chained with programmatic dependent launch: each requests its weight
// tiles, then griddepcontrol.wait.
// The contractions are tcgen05.mma (kind::i8 with int32 accumulators, kind::f16 for the bf16/fp16 absorb) fed by TMA
// into 128B-swizzled shared memory with accumulators in TMEM.  Arithmetic follows the reference's torch golden
// (tests/python/sgl_kernel_npu/test_mla_preprocess.py:57-107, 407-484) operation by operation (see oracle/).
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <type_traits>

#include "../../include/mla_preprocess_b200.h"
#include "sm100_ptx.cuh"

namespace mlapo {

// ============================================================================================ numerics
template <class T> struct Act;
template <> struct Act<__nv_bfloat16> {
  static __device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_f(float f) { return __float2bfloat16_rn(f); }
  static __device__ __forceinline__ float lo(uint32_t w) { return __uint_as_float(w << 16); }
  static __device__ __forceinline__ float hi(uint32_t w) { return __uint_as_float(w & 0xFFFF0000u); }
  static __device__ __forceinline__ uint32_t pack(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  }
  static __device__ __forceinline__ float round(float f) { return __bfloat162float(__float2bfloat16_rn(f)); }
};
template <> struct Act<__half> {
  static __device__ __forceinline__ float to_f(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_f(float f) { return __float2half_rn(f); }
  static __device__ __forceinline__ float lo(uint32_t w) { return __half2float(__ushort_as_half((unsigned short)(w & 0xFFFF))); }
  static __device__ __forceinline__ float hi(uint32_t w) { return __half2float(__ushort_as_half((unsigned short)(w >> 16))); }
  static __device__ __forceinline__ uint32_t pack(float a, float b) {
    __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
  }
  static __device__ __forceinline__ float round(float f) { return __half2float(__float2half_rn(f)); }
};

// fp32 -> fp16 (RNE) -> clamp [-128,127] -> round-half-even -> int8   (golden: test_mla:81-83)
__device__ __forceinline__ int sat_i8(float f) {
  float h = __half2float(__float2half_rn(f));
  h = fminf(fmaxf(h, -128.f), 127.f);
  return __float2int_rn(h);
}
__device__ __forceinline__ uint32_t pack4_i8(int a, int b, int c, int d) {
  return (uint32_t)(a & 0xFF) | ((uint32_t)(b & 0xFF) << 8) | ((uint32_t)(c & 0xFF) << 16) | ((uint32_t)(d & 0xFF) << 24);
}
template <class T>
__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = Act<T>::lo(v.x); f[1] = Act<T>::hi(v.x); f[2] = Act<T>::lo(v.y); f[3] = Act<T>::hi(v.y);
  f[4] = Act<T>::lo(v.z); f[5] = Act<T>::hi(v.z); f[6] = Act<T>::lo(v.w); f[7] = Act<T>::hi(v.w);
}
__device__ __forceinline__ float load_descale(const void* p, int is_i64, int i) {
  if (is_i64) return __int_as_float((int)(reinterpret_cast<const long long*>(p)[i] & 0xFFFFFFFFll));
  return reinterpret_cast<const float*>(p)[i];
}

// Debug timeline (flag bit3): per-CTA clock64 marks, 16 slots per CTA, written only when the pointer is non-null.
__device__ __forceinline__ void tl_mark(long long* tl, int slot) {
  if (tl) {
    const int cta = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
    tl[cta * 16 + slot] = clock64();
  }
}
__device__ __forceinline__ void tl_mark_global(long long* tl, int slot) {
  if (tl) {
    const int cta = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    tl[cta * 16 + slot] = (long long)t;
  }
}

// ============================================================================================ k0: NZ -> K-major repack
// FRACTAL_NZ [1, K/32, n_out, 32]: element (n,k) at ((k/32)*n_out + n)*32 + k%32  ->  [n_out, K] row-major.
__global__ void repack_nz_kernel(const int8_t* __restrict__ nz, int8_t* __restrict__ km, int k, int n_out) {
  const long long chunks = (long long)n_out * (k / 32) * 2;  // 16-byte chunks
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < chunks; i += (long long)gridDim.x * blockDim.x) {
    const int half = (int)(i & 1);
    const long long j = i >> 1;
    const int kb = (int)(j % (k / 32));
    const int n = (int)(j / (k / 32));
    const uint4 v = *reinterpret_cast<const uint4*>(nz + ((long long)kb * n_out + n) * 32 + half * 16);
    *reinterpret_cast<uint4*>(km + (long long)n * k + kb * 32 + half * 16) = v;
  }
}

// ============================================================================================ k1: quantise hiddenState
// x / s, correctly rounded, without the generic IEEE-division sequence: x and s both carry <= 11 significant bits
// (bf16/fp16), so the exact quotient lies >= 2^-36 (relative) away from every fp32 rounding boundary, while
// q1 = fma(fma(-q0, s, x), r, q0) with r = RN(1/s), q0 = RN(x*r) carries a relative error of ~2^-46 before its
// final rounding: q1 == RN(x / s).  Guarded to the range where neither r nor the products leave the normal range.
struct QuantScale {
  float s, zp, rcp;
  bool fast;
  template <class T>
  __device__ __forceinline__ void load(const T* scale, const int8_t* offset) {
    s = Act<T>::to_f(scale[0]);
    zp = (float)offset[0];
    rcp = __frcp_rn(s);
    fast = fabsf(s) > 1e-30f && fabsf(s) < 1e30f;
  }
};

// Per-tensor quantisation of the 16-byte vectors first, first+stride, ...; U independent loads in flight per thread.
template <class T, int U>
__device__ __forceinline__ void quant_vectors(const T* __restrict__ x, int8_t* __restrict__ q, const QuantScale& qs, long long n_vec8,
                                              long long first, long long stride) {
  for (long long i0 = first; i0 < n_vec8; i0 += stride * U) {
    uint4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u)  // all loads of the batch are issued before any is consumed
      if (i0 + u * stride < n_vec8) v[u] = ptx::ld_global_nc_v4(reinterpret_cast<const uint4*>(x) + i0 + u * stride);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i >= n_vec8) break;
      float f[8];
      unpack8<T>(v[u], f);
      int r[8];
      if (qs.fast) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float q0 = __fmul_rn(f[j], qs.rcp);
          const float q1 = __fmaf_rn(__fmaf_rn(-q0, qs.s, f[j]), qs.rcp, q0);
          r[j] = sat_i8(__fadd_rn(q1, qs.zp));
        }
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = sat_i8(__fadd_rn(__fdiv_rn(f[j], qs.s), qs.zp));
      }
      ptx::st_global_v2(q + i * 8, pack4_i8(r[0], r[1], r[2], r[3]), pack4_i8(r[4], r[5], r[6], r[7]));
    }
  }
}

template <class T, int U>
__global__ void quant_hidden_kernel(const T* __restrict__ x, int8_t* __restrict__ q, const T* __restrict__ scale,
                                    const int8_t* __restrict__ offset, long long n_vec8, long long* tl) {
  if (threadIdx.x == 0 && blockIdx.x < 4096) tl_mark_global(tl, 0);
  ptx::griddep_launch_dependents();
  QuantScale qs;
  qs.load(scale, offset);
  ptx::griddep_wait();  // a0 may still be read by the previous call's MM1 until its successors have completed
  quant_vectors<T, U>(x, q, qs, n_vec8, blockIdx.x * (long long)blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x);
  if (threadIdx.x == 0 && blockIdx.x < 4096) tl_mark_global(tl, 15);
}

// per-token symmetric (UNPINNED restatement of bf16.hpp:437-475): one row, executed by the whole CTA (any block size)
template <class T>
__device__ __forceinline__ void quant_row_per_token(const T* __restrict__ x, int8_t* __restrict__ q, float* __restrict__ tok_scale, int d,
                                                    int row) {
  __shared__ float s_red[32];
  const uint4* xr = reinterpret_cast<const uint4*>(x + (long long)row * d);
  float amax = 0.f;
  for (int i = threadIdx.x; i < d / 8; i += blockDim.x) {
    float f[8];
    unpack8<T>(xr[i], f);
#pragma unroll
    for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fabsf(f[j]));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = amax;
  __syncthreads();
  amax = s_red[0];
  for (int w = 1; w < (int)(blockDim.x >> 5); ++w) amax = fmaxf(amax, s_red[w]);
  const float s = __fdiv_rn(amax, 127.f);
  const float inv = __fdiv_rn(1.f, s);
  if (threadIdx.x == 0) tok_scale[row] = s;
  for (int i = threadIdx.x; i < d / 8; i += blockDim.x) {
    float f[8];
    unpack8<T>(xr[i], f);
    int r[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = sat_i8(__fmul_rn(f[j], inv));
    ptx::st_global_v2(q + (long long)row * d + i * 8, pack4_i8(r[0], r[1], r[2], r[3]), pack4_i8(r[4], r[5], r[6], r[7]));
  }
  __syncthreads();  // s_red is re-used by the next row
}

template <class T>
__global__ void __launch_bounds__(256) quant_hidden_per_token_kernel(const T* __restrict__ x, int8_t* __restrict__ q,
                                                                      float* __restrict__ tok_scale, int d) {
  ptx::griddep_launch_dependents();
  ptx::griddep_wait();
  quant_row_per_token<T>(x, q, tok_scale, d, blockIdx.x);
}

// ============================================================================================ k2: MM1 (W_DQKV), split-K
// acc[MT*128 tokens x 128] (int32, TMEM) = a0[tokens x Kslice] (int8, K-major) . Wdqkv[128 rows x Kslice]^T (int8, K-major).
// grid = (n_tiles=17, splits, token super-tiles of MT*128).  One CTA streams its weight tile once for all MT token tiles.
// 6 warps: w0 = TMA producer, w1 = tcgen05.mma issuer (+TMEM alloc), w2..w5 = epilogue (TMEM lane quarter = warp%4).
// Weight (B) tiles do not depend on the upstream kernel, so they are requested before griddepcontrol.wait (PDL).
constexpr int kGemmThreads = 192;
constexpr int kBlockM = 128;
constexpr int kBlockKBytes = 128;  // one 128B swizzle atom along K
constexpr int kUmmaKBytes = 32;    // K per tcgen05.mma for 8-bit (32 elems) and 16-bit (16 elems) operands
constexpr int kTile16K = 16384;    // [128 rows x 128 B] swizzled tile
constexpr int kMm1BlockN = 128;   // decode: 17 weight tiles x up to 8 K-splits fill one wave
constexpr int kMm1BlockNWide = 256;  // prefill (no split-K): 256 x 256 CTA tiles halve the activation re-reads per weight byte

template <int MT, int BN>
__host__ __device__ constexpr int mm1_stages() { return (MT * kTile16K + BN * kBlockKBytes) > 49152 ? 3 : 4; }
template <int MT, int BN>
constexpr int mm1_smem_bytes() {
  return mm1_stages<MT, BN>() * (MT * kTile16K + BN * kBlockKBytes) + 256 + 1024;
}

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

// Warp roles of MM1: warp 0 = TMA producer, warp 1 = tcgen05.mma issuer, warps 2..5 = epilogue (TMEM lane quarter = warp%4);
// other warps fall through.  The caller has already issued mm1_issue_weights().  kStandalone: own kernel under PDL, the
// producer waits for the upstream kernel before it touches a0; otherwise the caller guarantees a0 is complete.
template <int MT, int BN, bool kStandalone>
__device__ __forceinline__ void mm1_roles(const Mm1Ctx<MT, BN>& c, const CUtensorMap* tm_a, const CUtensorMap* tm_b,
                                          const CUtensorMap* tm_p, int warp, int lane, long long* tl) {
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
      if (kStandalone) {
        ptx::griddep_wait();  // a0 is produced by the quantise kernel (the weight tiles were requested at kernel start)
        tl_mark(tl, 3);       // upstream complete
      }
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
    // ------------------------------------------------------------------ MMA issuer (one thread)
    constexpr uint32_t idesc = ptx::make_idesc(ptx::kCFmtS32, ptx::kFmtS8, ptx::kFmtS8, 0, 0, kBlockM, BN);
    uint32_t stage = 0, phase = 0;
    for (int i = 0; i < nkb; ++i) {
      ptx::mbar_wait(ptx::smem_u32(full_bar + stage), phase);
      __syncwarp();
      ptx::tc_fence_after();
      if (lane == 0) {
        if (i == 0) tl_mark(tl, 4);             // first stage landed
        if (i == nkb - 1) tl_mark(tl, 5);       // last stage landed
        const uint32_t a_base = ptx::smem_u32(smem + stage * STAGE_BYTES);
        const uint32_t b_base = a_base + A_BYTES;
        for (int m = 0; m < mt_valid; ++m) {
#pragma unroll
          for (int k = 0; k < kBlockKBytes / kUmmaKBytes; ++k) {
            const uint64_t da = ptx::make_smem_desc(a_base + m * kTile16K + k * kUmmaKBytes, 1024, 0);
            const uint64_t db = ptx::make_smem_desc(b_base + k * kUmmaKBytes, 1024, 0);
            ptx::umma_i8(tmem_acc + m * BN, da, db, idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
        }
        ptx::umma_commit(ptx::smem_u32(empty_bar + stage));  // frees the smem slot when these MMAs retire
      }
      __syncwarp();
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
    if (lane == 0) ptx::umma_commit(ptx::smem_u32(done_bar));
    __syncwarp();
  } else if (warp < 6) {
    // ------------------------------------------------------------------ epilogue: TMEM -> swizzled smem -> TMA store
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t t_base = tmem_acc + ((uint32_t)(quarter * 32) << 16);
    const bool issuer = threadIdx.x == 64;
    ptx::mbar_wait(ptx::smem_u32(done_bar), 0);   // all MMAs retired => every pipeline stage has been consumed
    __syncwarp();
    ptx::tc_fence_after();
    if (issuer) tl_mark(tl, 6);  // MMAs retired
    __syncwarp();                                 // tcgen05.ld below is warp-aligned: reconverge after the one-lane branch
    for (int m = 0; m < mt_valid; ++m) {
      if (m > 0) {  // the staging tiles are re-used per token tile: the previous tile's stores must have been read
        if (issuer) ptx::tma_store_wait_read<0>();
        ptx::named_bar_sync(1, 128);
      }
#pragma unroll 1
      for (int g = 0; g < GRANULES; ++g) {
        uint32_t r[32];
        ptx::tmem_ld_32x32b_x32(t_base + m * BN + g * 32, r);
        ptx::tmem_ld_wait();
        const uint32_t tile = ptx::smem_u32(smem + g * kTile16K);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          ptx::st_shared_v4(tile + ptx::sw128_offset(row, j), r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
      }
      ptx::fence_proxy_async_smem();
      ptx::named_bar_sync(1, 128);
      if (issuer) {
        for (int g = 0; g < GRANULES; ++g)
          if (n0 + g * 32 < MLAPO_MM1_OUT)
            ptx::tma_store_3d(tm_p, ptx::smem_u32(smem + g * kTile16K), n0 + g * 32, m0 + m * kBlockM, split);
        ptx::tma_store_commit();
      }
    }
    if (issuer) { tl_mark(tl, 7); ptx::tma_store_wait_read<0>(); tl_mark(tl, 8); }
  }

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

  mm1_roles<MT, BN, true>(c, &tm_a, &tm_b, &tm_p, warp, lane, tl);

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(c.tmem_acc, C::TMEM_COLS);
  if (threadIdx.x == 32) { tl_mark(tl, 9); tl_mark_global(tl, 15); }
}

// ============================================================================================ k3: norms, quant, rope, cache
struct NormArgs {
  const int32_t* partial;  // [splits, N, 2112]
  int splits;
  int n_tokens;
  const int32_t* bias0;    // null in per-token mode
  const void* descale0;
  int descale_is_i64;
  const float* tok_scale0;  // per-token: s0[N]; else null
  float* tok_scale1;        // per-token: s1[N] out
  const void* gamma1; const void* beta1; const void* gamma2;
  const void* quant_scale1; const int8_t* quant_offset1;
  const void* cos; const void* sin;
  const int32_t* slots;
  const void* ctkv_scale;
  int8_t* a1;               // [N,1536]
  void* y0_tap;             // act [N,2112] or null
  void* kv_cache; void* kv_cache_rope;
  int cache_mode;
  long long cache_slots;    // num_blocks * block_size
  long long* tl;            // debug timeline or null
};

constexpr int kNormThreads = 288;  // warps 0-1: k_nope (512), warp 2 lanes 0-7: k_pe (64), warps 3-8: q_a (1536)

__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One token row, executed by every thread of the CTA (threads beyond the first 288 only take part in the barriers).
template <class T, int kMaxSplits>
__device__ __forceinline__ void norm_token(const NormArgs& a, const int t, const int tid) {
  __shared__ double s_sum[9];
  __shared__ float s_max[9];
  const int warp = tid >> 5, lane = tid & 31;
  const bool is_nope = warp < 2;
  const bool is_pe = warp == 2;
  const bool is_q = warp >= 3 && warp < 9;
  const bool active = is_nope || (is_pe && lane < 8) || is_q;
  int col = 0;  // first of this thread's 8 columns in [0, 2112)
  if (is_nope) col = tid * 8;
  else if (is_pe) col = 512 + (lane & 7) * 8;
  else if (is_q) col = 576 + (tid - 96) * 8;
  // ---- sum split-K partials (+bias), dequantise, round to act  (S2 of the contract)
  // All global loads are issued before anything is consumed (the kernel is latency-, not bandwidth-bound).
  float x[8];
  uint4 g_vec = make_uint4(0, 0, 0, 0), b_vec = make_uint4(0, 0, 0, 0);  // gamma (+beta) or cos/sin of this thread's 8 columns
  {
    int4 pu[kMaxSplits], pv[kMaxSplits];
#pragma unroll
    for (int s = 0; s < kMaxSplits; ++s) {
      pu[s] = make_int4(0, 0, 0, 0);
      pv[s] = make_int4(0, 0, 0, 0);
      if (active && s < a.splits) {
        const int4* p = reinterpret_cast<const int4*>(a.partial + ((long long)s * a.n_tokens + t) * MLAPO_MM1_OUT + col);
        pu[s] = p[0];
        pv[s] = p[1];
      }
    }
    float d[8];
    int bias[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (a.descale_is_i64) {
#pragma unroll
      for (int j = 0; j < 8; ++j) d[j] = load_descale(a.descale0, 1, col + j);
    } else {
      const float4 d0 = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(a.descale0) + col);
      const float4 d1 = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(a.descale0) + col + 4);
      d[0] = d0.x; d[1] = d0.y; d[2] = d0.z; d[3] = d0.w; d[4] = d1.x; d[5] = d1.y; d[6] = d1.z; d[7] = d1.w;
    }
    if (a.bias0) {
      const int4 b0 = *reinterpret_cast<const int4*>(a.bias0 + col), b1 = *reinterpret_cast<const int4*>(a.bias0 + col + 4);
      bias[0] = b0.x; bias[1] = b0.y; bias[2] = b0.z; bias[3] = b0.w; bias[4] = b1.x; bias[5] = b1.y; bias[6] = b1.z; bias[7] = b1.w;
    }
    if (is_nope) {
      g_vec = reinterpret_cast<const uint4*>(a.gamma2)[tid];
    } else if (is_pe) {
      g_vec = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(a.cos) + (long long)t * 64)[lane & 7];
      b_vec = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(a.sin) + (long long)t * 64)[lane & 7];
    } else if (is_q) {
      g_vec = reinterpret_cast<const uint4*>(a.gamma1)[tid - 96];
      b_vec = reinterpret_cast<const uint4*>(a.beta1)[tid - 96];
    }
    const float ts0 = a.tok_scale0 ? a.tok_scale0[t] : 1.f;
    int acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int s = 0; s < kMaxSplits; ++s) {
      acc[0] += pu[s].x; acc[1] += pu[s].y; acc[2] += pu[s].z; acc[3] += pu[s].w;
      acc[4] += pv[s].x; acc[5] += pv[s].y; acc[6] += pv[s].z; acc[7] += pv[s].w;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float y;
      if (a.tok_scale0) y = __fmul_rn(__fmul_rn(__int2float_rn(acc[j]), d[j]), ts0);
      else y = __fmul_rn(__int2float_rn(acc[j] + bias[j]), d[j]);
      x[j] = Act<T>::round(y);
    }
    if (a.y0_tap && active) {
      T* yo = reinterpret_cast<T*>(a.y0_tap) + (long long)t * MLAPO_MM1_OUT + col;
      ptx::st_global_v4(yo, Act<T>::pack(x[0], x[1]), Act<T>::pack(x[2], x[3]), Act<T>::pack(x[4], x[5]), Act<T>::pack(x[6], x[7]));
    }
  }

  // ---- row sums of squares, accumulated in fp64 (order-free), rounded once to fp32
  double ss = 0.0;
  if (active && !is_pe) {
#pragma unroll
    for (int j = 0; j < 8; ++j) ss += (double)x[j] * (double)x[j];
  }
  ss = warp_sum_f64(ss);
  if (lane == 0 && warp < 9) s_sum[warp] = ss;
  __syncthreads();
  const double tot_nope = s_sum[0] + s_sum[1];
  const double tot_q = ((s_sum[3] + s_sum[4]) + (s_sum[5] + s_sum[6])) + (s_sum[7] + s_sum[8]);

  const int slot = a.slots[t];
  const bool slot_ok = slot >= 0 && (long long)slot < a.cache_slots;
  const int blk = slot >> 7, r128 = slot & 127;  // NZ modes: block size 128

  if (is_nope) {
    // S8: k_nope = act(x * rsqrt(mean+eps) * gamma2)   (golden test_mla:70-74, 433)
    const float var = __fdiv_rn(__double2float_rn(tot_nope), 512.f);
    const float rs = __fdiv_rn(1.f, __fsqrt_rn(__fadd_rn(var, 1e-6f)));
    float g[8];
    unpack8<T>(g_vec, g);
    float y[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) y[j] = __fmul_rn(__fmul_rn(x[j], rs), g[j]);
    if (slot_ok) {
      if (a.cache_mode == MLAPO_CACHE_INT8_NZ) {
        const float cs = Act<T>::to_f(reinterpret_cast<const T*>(a.ctkv_scale)[0]);
        int q[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) q[j] = sat_i8(__fdiv_rn(y[j], cs));
        int8_t* dst = reinterpret_cast<int8_t*>(a.kv_cache) + (long long)blk * (128 * 512) + (col >> 5) * (128 * 32) + r128 * 32 + (col & 31);
        ptx::st_global_v2(dst, pack4_i8(q[0], q[1], q[2], q[3]), pack4_i8(q[4], q[5], q[6], q[7]));
      } else {
        T* base = reinterpret_cast<T*>(a.kv_cache);
        T* dst = a.cache_mode == MLAPO_CACHE_NZ
                     ? base + (long long)blk * (128 * 512) + (col >> 4) * (128 * 16) + r128 * 16 + (col & 15)
                     : base + (long long)slot * 512 + col;
        ptx::st_global_v4(dst, Act<T>::pack(y[0], y[1]), Act<T>::pack(y[2], y[3]), Act<T>::pack(y[4], y[5]), Act<T>::pack(y[6], y[7]));
      }
    }
  } else if (is_pe) {
    // S7: k_pe = rope(x[512:576])   lanes l and l^4 hold partner halves
    float xp[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) xp[j] = __shfl_xor_sync(0xffffffffu, x[j], 4);
    if (lane < 8) {
      const int c = lane * 8;  // column within the 64 rope lanes
      float cs[8], sn[8];
      unpack8<T>(g_vec, cs);
      unpack8<T>(b_vec, sn);
      float y[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float rot = lane < 4 ? -xp[j] : xp[j];
        y[j] = __fadd_rn(__fmul_rn(x[j], cs[j]), __fmul_rn(rot, sn[j]));
      }
      if (slot_ok) {
        T* base = reinterpret_cast<T*>(a.kv_cache_rope);
        T* dst = a.cache_mode == MLAPO_CACHE_KROPE_CTKV
                     ? base + (long long)slot * 64 + c
                     : base + (long long)blk * (128 * 64) + (c >> 4) * (128 * 16) + r128 * 16 + (c & 15);
        ptx::st_global_v4(dst, Act<T>::pack(y[0], y[1]), Act<T>::pack(y[2], y[3]), Act<T>::pack(y[4], y[5]), Act<T>::pack(y[6], y[7]));
      }
    }
  }

  // S4: q_a RMSNorm(+beta) + quant -> a1   (golden test_mla:432, 436-438); needs all warps for the per-token max
  float yq[8];
  float amax = 0.f;
  if (is_q) {
    const float var = __fdiv_rn(__double2float_rn(tot_q), 1536.f);
    const float rs = __fdiv_rn(1.f, __fsqrt_rn(__fadd_rn(var, 1e-6f)));
    float g[8], b[8];
    unpack8<T>(g_vec, g);
    unpack8<T>(b_vec, b);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      yq[j] = __fadd_rn(__fmul_rn(__fmul_rn(x[j], rs), g[j]), b[j]);
      amax = fmaxf(amax, fabsf(yq[j]));
    }
  }
  if (a.tok_scale1) {  // per-token symmetric (UNPINNED; bf16.hpp:739-777)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if (lane == 0 && warp < 9) s_max[warp] = amax;
    __syncthreads();
    float m = s_max[3];
#pragma unroll
    for (int w = 4; w < 9; ++w) m = fmaxf(m, s_max[w]);
    const float s1 = __fdiv_rn(m, 127.f);
    const float inv = __fdiv_rn(1.f, s1);
    if (tid == 96) a.tok_scale1[t] = s1;
    if (is_q) {
      int q[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) q[j] = sat_i8(__fmul_rn(yq[j], inv));
      ptx::st_global_v2(a.a1 + (long long)t * MLAPO_Q_LORA + (col - 576), pack4_i8(q[0], q[1], q[2], q[3]), pack4_i8(q[4], q[5], q[6], q[7]));
    }
  } else if (is_q) {
    const float s = Act<T>::to_f(reinterpret_cast<const T*>(a.quant_scale1)[0]);
    const float zp = (float)a.quant_offset1[0];
    int q[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) q[j] = sat_i8(__fadd_rn(__fdiv_rn(yq[j], s), zp));
    ptx::st_global_v2(a.a1 + (long long)t * MLAPO_Q_LORA + (col - 576), pack4_i8(q[0], q[1], q[2], q[3]), pack4_i8(q[4], q[5], q[6], q[7]));
  }
}

template <class T, int kMaxSplits>
__global__ void __launch_bounds__(kNormThreads, kMaxSplits == 1 ? 4 : 2) norm_rope_cache_kernel(NormArgs a) {
  const int tid = threadIdx.x;
  if (tid == 0 && blockIdx.x < 4096) tl_mark_global(a.tl, 0);
  ptx::griddep_launch_dependents();
  ptx::griddep_wait();  // partial sums come from the MM1 kernel
  if (tid == 0 && blockIdx.x < 4096) tl_mark_global(a.tl, 14);
  norm_token<T, kMaxSplits>(a, blockIdx.x, tid);
  if (tid == 0 && blockIdx.x < 4096) tl_mark_global(a.tl, 15);
}

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
constexpr int kQSmemBytes = kQBarOff + 256 + MLAPO_Q_HEAD * 8 + 1024;
static_assert(kQSmemBytes <= 232448, "q_fused smem");
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
};

// Where one CTA's q-path work item lives.
struct QCtx {
  uint8_t* smem;     // 224 KB pipeline / overlay region, 1024-byte aligned
  uint64_t* bars;    // full[4] empty[4] mm2_done wuk_full[4] qn_ready[2] unit_full[4] buf_empty[4] rope_done[2]
  int2* s_cst;       // per column {bias, descale bits}
  uint32_t tmem;     // base of a 512-column TMEM allocation
  int head, tok0, num_mt, load_mt;
  static constexpr int NUM_BARS = 25;

  __device__ void set_work(int head_, int tok_tile, const QArgs& qa) {
    head = head_;
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

// One thread: arm the four MM2 stages and request their W_UQ k-blocks (which depend on nothing upstream).
__device__ __forceinline__ void q_issue_weights(const QCtx& c, const CUtensorMap* tm_wuq) {
  const uint32_t stage_tx = (uint32_t)(c.load_mt * kTile16K + MLAPO_Q_HEAD * kBlockKBytes);
  for (int kb = 0; kb < kQStages; ++kb) {
    const uint32_t fb = ptx::smem_u32(c.bars + kb);
    ptx::mbar_arrive_expect_tx(fb, stage_tx);
    ptx::tma_load_2d(ptx::smem_u32(c.smem + kb * kQStageBytes + 2 * kTile16K), tm_wuq, fb, kb * kBlockKBytes, c.head * MLAPO_Q_HEAD,
                     ptx::kEvictFirst);
  }
}

// Warp roles of the q path: warp 0 = TMA producer, warp 1 = MMA issuer, warps 2..5 / 6..9 = epilogue groups of token tile 0 / 1.
// The caller has already issued q_issue_weights().  kStandalone: own kernel under PDL (waits for the upstream kernel
// before touching a1 / the per-token scales); otherwise the caller guarantees they are complete.
template <class T, bool kInt8Out, bool kStandalone>
__device__ __forceinline__ void q_roles(const QCtx& c, const CUtensorMap* tm_a1, const CUtensorMap* tm_wuq, const CUtensorMap* tm_wuk,
                                        const CUtensorMap* tm_out0, const CUtensorMap* tm_out1, const QArgs& qa, int warp, int lane,
                                        long long* tl) {
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
  const int head = c.head, tok0 = c.tok0, num_mt = c.num_mt, load_mt = c.load_mt;
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      const uint32_t stage_tx = (uint32_t)(load_mt * kTile16K + MLAPO_Q_HEAD * kBlockKBytes);
      auto load_b = [&](int kb) {
        const int s = kb & 3;
        ptx::tma_load_2d(ptx::smem_u32(smem + s * kQStageBytes + 2 * kTile16K), tm_wuq, ptx::smem_u32(full_bar + s),
                         kb * kBlockKBytes, head * MLAPO_Q_HEAD, ptx::kEvictFirst);
      };
      auto load_a = [&](int kb) {  // one box = all token tiles of the stage (box rows = 128 * load_mt)
        const int s = kb & 3;
        ptx::tma_load_2d(ptx::smem_u32(smem + s * kQStageBytes), tm_a1, ptx::smem_u32(full_bar + s), kb * kBlockKBytes, tok0,
                         ptx::kEvictLast);
      };
      if (kStandalone) {
        ptx::griddep_wait();  // a1 is produced by the norm kernel (the first W_UQ k-blocks were requested at kernel start)
        tl_mark(tl, 3);
      }
      for (int kb = 0; kb < kQStages; ++kb) load_a(kb);
      for (int kb = kQStages; kb < kQKBlocks; ++kb) {
        const int s = kb & 3;
        ptx::mbar_wait(ptx::smem_u32(empty_bar + s), ((kb >> 2) & 1) ^ 1);
        ptx::mbar_arrive_expect_tx(ptx::smem_u32(full_bar + s), stage_tx);
        load_b(kb);
        load_a(kb);
      }
      // Wuk chunk c -> bytes [32K*c, 32K*c+32K): needs the last use (3rd release, parity 0) of the stages it overlaps
      auto load_wuk = [&](int c) {
        const uint32_t fb = ptx::smem_u32(wuk_full + c);
        ptx::mbar_arrive_expect_tx(fb, 2 * kTile16K);
        for (int at = 0; at < 2; ++at)
          ptx::tma_load_2d(ptx::smem_u32(smem + kQWukOff + c * 32768 + at * kTile16K), tm_wuk, fb, c * 128 + at * 64,
                           head * MLAPO_Q_NOPE, ptx::kEvictFirst);
      };
      ptx::mbar_wait(ptx::smem_u32(empty_bar + 0), 0);
      load_wuk(0);                                   // [0, 32K)    in stage 0
      ptx::mbar_wait(ptx::smem_u32(empty_bar + 1), 0);
      load_wuk(1);                                   // [32K, 64K)  in stages 0, 1
      load_wuk(2);                                   // [64K, 96K)  in stage 1
      ptx::mbar_wait(ptx::smem_u32(empty_bar + 2), 0);
      load_wuk(3);                                   // [96K, 128K) in stages 1, 2
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc2 = ptx::make_idesc(ptx::kCFmtS32, ptx::kFmtS8, ptx::kFmtS8, 0, 0, kBlockM, MLAPO_Q_HEAD);
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
        const uint32_t a_base = ptx::smem_u32(smem + s * kQStageBytes);
        const uint32_t b_base = a_base + 2 * kTile16K;
        for (int m = 0; m < num_mt; ++m) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint64_t da = ptx::make_smem_desc(a_base + m * kTile16K + k * kUmmaKBytes, 1024, 0);
            const uint64_t db = ptx::make_smem_desc(b_base + k * kUmmaKBytes, 1024, 0);
            ptx::umma_i8(tmem + m * 256, da, db, idesc2, (kb > 0 || k > 0) ? 1u : 0u);
          }
        }
        ptx::umma_commit(ptx::smem_u32(empty_bar + s));
      }
      __syncwarp();
    }
    if (lane == 0) ptx::umma_commit(ptx::smem_u32(mm2_done));
    __syncwarp();
    // absorb units (c, m): token tile m ping-pongs its own two TMEM buffers over the four Wuk chunks
    for (int c = 0; c < 4; ++c) {
      for (int m = 0; m < num_mt; ++m) {
        const int ub = m * 2 + (c & 1);
        if (c == 0) ptx::mbar_wait(ptx::smem_u32(qn_ready + m), 0);   // buffer 0 = acc2[m] columns 0..127 (nope, drained)
        if (c == 1) ptx::mbar_wait(ptx::smem_u32(rope_done + m), 0);  // buffer 1 overlaps acc2[m] columns 128..191 (rope)
        if (c >= 2) ptx::mbar_wait(ptx::smem_u32(buf_empty + ub), 0);
        if (m == 0) ptx::mbar_wait(ptx::smem_u32(wuk_full + c), 0);
        __syncwarp();
        ptx::tc_fence_after();
        if (lane == 0) {
          const uint32_t b_base = ptx::smem_u32(smem + kQWukOff + c * 32768);
#pragma unroll
          for (int k = 0; k < 8; ++k) {  // 8 x UMMA_K(16) = 128
            const uint32_t a_addr = ptx::smem_u32(smem + kQNopeOff + (m * 2 + (k >> 2)) * kTile16K) + (k & 3) * kUmmaKBytes;
            const uint64_t da = ptx::make_smem_desc(a_addr, 1024, 0);
            const uint64_t db = ptx::make_smem_desc(b_base + k * 2048, 1024, kTile16K);  // MN-major: SBO 8 k-rows, LBO next 64-col atom
            ptx::umma_f16(tmem + m * 256 + (c & 1) * 128, da, db, idesc3, k > 0 ? 1u : 0u);
          }
          ptx::umma_commit(ptx::smem_u32(unit_full + ub));
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
      for (int c = et; c < MLAPO_Q_HEAD; c += 128) {
        const float dsc = load_descale(qa.descale1, qa.descale_is_i64, head * MLAPO_Q_HEAD + c);
        s_cst[c] = make_int2(qa.bias1 ? qa.bias1[head * MLAPO_Q_HEAD + c] : 0, __float_as_int(dsc));
      }
    }
    if (kStandalone) ptx::griddep_wait();  // tok_scale1 comes from the norm kernel (and keeps the PDL completion chain transitive)
    ptx::named_bar_sync(3, 256);

    if (g < num_mt) {
      const int m = g;
      const bool per_token = qa.tok_scale1 != nullptr;
      const int t = tok0 + m * kBlockM + row;
      const int tc = t < qa.m_total ? t : qa.m_total - 1;  // clamped row for loads; TMA stores clip rows >= N
      const float ts = per_token ? qa.tok_scale1[tc] : 1.f;
      auto deq = [&](uint32_t acc, int c) -> float {
        const int2 cst = s_cst[c];
        return per_token ? __fmul_rn(__fmul_rn(__int2float_rn((int)acc), __int_as_float(cst.y)), ts)
                         : __fmul_rn(__int2float_rn((int)acc + cst.x), __int_as_float(cst.y));
      };
      // cos/sin rows of this token (issued before the MM2 wait; consumed after the nope slice)
      uint4 cs[8], sn[8];
      {
        const uint4* cp = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(qa.cos) + (long long)tc * 64);
        const uint4* sp = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(qa.sin) + (long long)tc * 64);
#pragma unroll
        for (int j = 0; j < 8; ++j) { cs[j] = ptx::ld_global_nc_v4(cp + j); sn[j] = ptx::ld_global_nc_v4(sp + j); }
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
          for (int u = 0; u < 4; ++u) w[u] = Act<T>::pack(deq(r[8 * j + 2 * u], c0 + 8 * j + 2 * u), deq(r[8 * j + 2 * u + 1], c0 + 8 * j + 2 * u + 1));
          ptx::st_shared_v4(tile + ptx::sw128_offset(row, ((c0 & 63) >> 3) + j), w[0], w[1], w[2], w[3]);
        }
      };
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
        for (int j = 0; j < 32; ++j) x[j] = Act<T>::round(deq(ra[j], 128 + j));
        ptx::tmem_ld_wait();
        // acc2[m] fully read: the absorb buffer that overlaps the rope columns is free too
        ptx::tc_fence_before();
        ptx::mbar_arrive(ptx::smem_u32(rope_done + m));
#pragma unroll
        for (int j = 0; j < 32; ++j) x[32 + j] = Act<T>::round(deq(rb[j], 160 + j));
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
      if (issuer && g == 0) tl_mark(tl, 7);  // epi2 done

      // absorb drains
      float qs = 1.f;
      if (kInt8Out) qs = Act<T>::to_f(reinterpret_cast<const T*>(qa.q_nope_scale)[head]);
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        const int ub = m * 2 + (c & 1);
        ptx::mbar_wait(ptx::smem_u32(unit_full + ub), (c >> 1) & 1);
        __syncwarp();
        ptx::tc_fence_after();
        if (issuer && g == 0 && c == 0) tl_mark(tl, 8);  // first absorb unit ready
        if (issuer && g == 0 && c == 3) tl_mark(tl, 9);
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
          if (c < 2) ptx::mbar_arrive(ptx::smem_u32(buf_empty + ub));
          stage_acquire();
#pragma unroll
          for (int j = 0; j < 8; ++j) ptx::st_shared_v4(st_tile + ptx::sw128_offset(row, j), w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
          stage_publish();
          if (issuer) {
            ptx::tma_store_2d(tm_out0, st_tile, head * MLAPO_KV_LORA + c * 128, tok0 + m * kBlockM);
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
              if (c < 2) ptx::mbar_arrive(ptx::smem_u32(buf_empty + ub));
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
              ptx::tma_store_2d(tm_out0, dst, head * MLAPO_KV_LORA + c * 128 + half * 64, tok0 + m * kBlockM);
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

template <class T, bool kInt8Out>
__global__ void __launch_bounds__(kQThreads, 1)
q_fused_kernel(const __grid_constant__ CUtensorMap tm_a1, const __grid_constant__ CUtensorMap tm_wuq,
               const __grid_constant__ CUtensorMap tm_wuk, const __grid_constant__ CUtensorMap tm_out0,
               const __grid_constant__ CUtensorMap tm_out1, QArgs qa, long long* tl) {
  if (threadIdx.x == 0) { tl_mark_global(tl, 0); tl_mark(tl, 1); }
  extern __shared__ uint8_t smem_raw[];
  QCtx c;
  c.smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  c.bars = reinterpret_cast<uint64_t*>(c.smem + kQBarOff);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(c.bars + 26);
  c.s_cst = reinterpret_cast<int2*>(c.smem + kQBarOff + 256);
  c.set_work(blockIdx.x, blockIdx.y, qa);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {  // barriers + the first W_UQ k-blocks before anything else: they depend on nothing
    ptx::prefetch_tmap(&tm_wuq);
    q_init_barriers(c.bars);
    ptx::fence_mbar_init();
    q_issue_weights(c, &tm_wuq);
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
  ptx::griddep_launch_dependents();
  if (threadIdx.x == 0) tl_mark(tl, 2);  // prologue done

  q_roles<T, kInt8Out, true>(c, &tm_a1, &tm_wuq, &tm_wuk, &tm_out0, &tm_out1, qa, warp, lane, tl);

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(c.tmem, 512);
  if (threadIdx.x == 32) { tl_mark(tl, 12); tl_mark_global(tl, 15); }
}

// ============================================================================================ k6: q path, streaming form (prefill)
// Same arithmetic as k4, re-shaped for calls with many more (head, 256-token) units than SMs: persistent CTAs walk a list
// of units and the phases of consecutive units overlap —  MM2 of unit i+1 (TMA + int8 MMAs) runs while unit i is being
// absorbed, drained and stored.  That needs the MM2 pipeline, the q_nope tile, the W_UK chunks and the output staging to
// live in DISJOINT shared memory, and separate issuing threads for the two MMA streams and the two TMA streams:
//   w0  TMA: a1 + W_UQ k-blocks (2 stages x 56 KB)          w10 TMA: W_UK 64-column chunks (3 slots x 16 KB)
//   w1  MMA: MM2 (kind::i8, acc2[m] = TMEM [256m, 256m+192)) w11 MMA: absorb (kind::f16, N=64, buffers TMEM [192,256), [448,512))
//   w2..w5 / w6..w9: epilogue group of token tile 0 / 1 (dequant, RoPE -> q_out1, q_nope -> smem, drains -> q_out0)
// The q_nope smem tile is single (32 KB): token tile 0 is absorbed first, then tile 1 re-uses the bytes, so every W_UK
// chunk is fetched twice per unit (L2 hits).  Units are ordered head-major: a CTA's next unit is the same or the next head.
constexpr int kSThreads = 384;
constexpr int kSStages = 2;
constexpr int kSWukSlots = 3;
constexpr int kSNopeOff = kSStages * kQStageBytes;              // 114688: 2 x 16 KB (k-block 0 / 1 of one token tile)
constexpr int kSWukOff = kSNopeOff + 2 * kTile16K;              // 147456: 3 x 16 KB
constexpr int kSStageOff = kSWukOff + kSWukSlots * kTile16K;    // 196608: 2 x 16 KB, one per epilogue group
constexpr int kSBarOff = kSStageOff + 2 * kTile16K;             // 229376
constexpr int kSSmemBytes = kSBarOff + 256 + MLAPO_Q_HEAD * 8 + 1024;
static_assert(kSSmemBytes <= 232448, "q_stream smem");

template <class T, bool kInt8Out>
__global__ void __launch_bounds__(kSThreads, 1)
q_stream_kernel(const __grid_constant__ CUtensorMap tm_a1, const __grid_constant__ CUtensorMap tm_wuq,
                const __grid_constant__ CUtensorMap tm_wuk, const __grid_constant__ CUtensorMap tm_out0,
                const __grid_constant__ CUtensorMap tm_out1, QArgs qa, int total_units, long long* tl) {
  if (threadIdx.x == 0) { tl_mark_global(tl, 0); tl_mark(tl, 1); }
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kSBarOff);
  uint64_t* full_bar = bars;           // [2]  TMA -> MM2
  uint64_t* empty_bar = bars + 2;      // [2]  MM2 -> TMA
  uint64_t* mm2_done = bars + 4;       // [1]  MM2 -> epilogue groups
  uint64_t* acc2_free = bars + 5;      // [1]  both epilogue groups (256 threads) -> MM2: acc2 of the previous unit drained
  uint64_t* qn_ready = bars + 6;       // [2]  epilogue group m -> absorb: q_nope tile of token tile m is in smem
  uint64_t* qn_free = bars + 8;        // [1]  absorb -> epilogue: the q_nope tile has been consumed (twice per unit)
  uint64_t* wuk_full = bars + 9;       // [3]  TMA -> absorb
  uint64_t* wuk_empty = bars + 12;     // [3]  absorb -> TMA
  uint64_t* unit_full = bars + 15;     // [2]  absorb -> draining epilogue group
  uint64_t* abuf_empty = bars + 17;    // [2]  draining epilogue group (128 threads) -> absorb
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 20);
  int2* s_cst = reinterpret_cast<int2*>(smem + kSBarOff + 256);  // per column {bias, descale bits} of the current head

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // units are walked head-major: the CTAs running at any moment share ~5 heads' weights (always L2-hot) and a1 (12.6 MB)
  const int n_st = (qa.m_total + 255) / 256;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_a1);
    ptx::prefetch_tmap(&tm_wuq);
    ptx::prefetch_tmap(&tm_wuk);
    ptx::prefetch_tmap(&tm_out0);
    ptx::prefetch_tmap(&tm_out1);
    for (int i = 0; i < 19; ++i) {
      const uint32_t cnt = i == 5 ? 256u : (i == 6 || i == 7 || i == 17 || i == 18) ? 128u : 1u;
      ptx::mbar_init(ptx::smem_u32(bars + i), cnt);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  ptx::griddep_launch_dependents();

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA: a1 + W_UQ k-blocks
    if (lane == 0) {
      ptx::griddep_wait();  // a1 is produced by the norm kernel
      constexpr uint32_t stage_tx = 2 * kTile16K + MLAPO_Q_HEAD * kBlockKBytes;
      uint32_t gk = 0;
      for (int u = blockIdx.x; u < total_units; u += gridDim.x) {
        const int head = u / n_st, tok0 = (u % n_st) * 256;
        for (int kb = 0; kb < kQKBlocks; ++kb, ++gk) {
          const uint32_t s = gk & 1;
          if (gk >= kSStages) ptx::mbar_wait(ptx::smem_u32(empty_bar + s), ((gk >> 1) & 1) ^ 1);
          const uint32_t fb = ptx::smem_u32(full_bar + s);
          ptx::mbar_arrive_expect_tx(fb, stage_tx);
          ptx::tma_load_2d(ptx::smem_u32(smem + s * kQStageBytes + 2 * kTile16K), &tm_wuq, fb, kb * kBlockKBytes, head * MLAPO_Q_HEAD,
                           ptx::kEvictLast);
          ptx::tma_load_2d(ptx::smem_u32(smem + s * kQStageBytes), &tm_a1, fb, kb * kBlockKBytes, tok0, ptx::kEvictLast);
        }
      }
    }
  } else if (warp == 10) {
    // ------------------------------------------------------------------ TMA: W_UK chunks (each twice per unit: tile 0, then tile 1)
    if (lane == 0) {
      uint32_t gc = 0;
      for (int u = blockIdx.x; u < total_units; u += gridDim.x) {
        const int head = u / n_st;
        for (int j = 0; j < 16; ++j, ++gc) {
          const uint32_t w = gc % kSWukSlots, use = gc / kSWukSlots;
          if (use >= 1) ptx::mbar_wait(ptx::smem_u32(wuk_empty + w), (use & 1) ^ 1);
          const uint32_t fb = ptx::smem_u32(wuk_full + w);
          ptx::mbar_arrive_expect_tx(fb, kTile16K);
          ptx::tma_load_2d(ptx::smem_u32(smem + kSWukOff + w * kTile16K), &tm_wuk, fb, (j & 7) * 64, head * MLAPO_Q_NOPE, ptx::kEvictLast);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA: MM2
    constexpr uint32_t idesc2 = ptx::make_idesc(ptx::kCFmtS32, ptx::kFmtS8, ptx::kFmtS8, 0, 0, kBlockM, MLAPO_Q_HEAD);
    uint32_t gk = 0, it = 0;
    for (int u = blockIdx.x; u < total_units; u += gridDim.x, ++it) {
      if (it > 0) ptx::mbar_wait(ptx::smem_u32(acc2_free), (it - 1) & 1);  // the previous unit's accumulators have been read
      for (int kb = 0; kb < kQKBlocks; ++kb, ++gk) {
        const uint32_t s = gk & 1;
        ptx::mbar_wait(ptx::smem_u32(full_bar + s), (gk >> 1) & 1);
        __syncwarp();
        ptx::tc_fence_after();
        if (lane == 0) {
          const uint32_t a_base = ptx::smem_u32(smem + s * kQStageBytes);
          const uint32_t b_base = a_base + 2 * kTile16K;
#pragma unroll
          for (int m = 0; m < 2; ++m) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t da = ptx::make_smem_desc(a_base + m * kTile16K + k * kUmmaKBytes, 1024, 0);
              const uint64_t db = ptx::make_smem_desc(b_base + k * kUmmaKBytes, 1024, 0);
              ptx::umma_i8(tmem + m * 256, da, db, idesc2, (kb > 0 || k > 0) ? 1u : 0u);
            }
          }
          ptx::umma_commit(ptx::smem_u32(empty_bar + s));
        }
        __syncwarp();
      }
      if (lane == 0) ptx::umma_commit(ptx::smem_u32(mm2_done));
      __syncwarp();
    }
  } else if (warp == 11) {
    // ------------------------------------------------------------------ MMA: absorb, 64 output columns per unit of work
    constexpr uint32_t kAFmt = std::is_same<T, __nv_bfloat16>::value ? ptx::kFmtBF16 : ptx::kFmtF16;
    constexpr uint32_t idesc3 = ptx::make_idesc(ptx::kCFmtF32, kAFmt, kAFmt, 0, 1, kBlockM, 64);
    uint32_t gc = 0, it = 0;
    for (int u = blockIdx.x; u < total_units; u += gridDim.x, ++it) {
      for (int m = 0; m < 2; ++m) {
        ptx::mbar_wait(ptx::smem_u32(qn_ready + m), it & 1);
        for (int c = 0; c < 8; ++c, ++gc) {
          const uint32_t w = gc % kSWukSlots, b = gc & 1;
          ptx::mbar_wait(ptx::smem_u32(wuk_full + w), (gc / kSWukSlots) & 1);
          if (gc >= 2) ptx::mbar_wait(ptx::smem_u32(abuf_empty + b), ((gc >> 1) - 1) & 1);
          __syncwarp();
          ptx::tc_fence_after();
          if (lane == 0) {
            const uint32_t b_base = ptx::smem_u32(smem + kSWukOff + w * kTile16K);
            const uint32_t d_col = tmem + (b ? 448u : 192u);
#pragma unroll
            for (int k = 0; k < 8; ++k) {  // 8 x UMMA_K(16) = 128
              const uint32_t a_addr = ptx::smem_u32(smem + kSNopeOff + (k >> 2) * kTile16K) + (k & 3) * kUmmaKBytes;
              const uint64_t da = ptx::make_smem_desc(a_addr, 1024, 0);
              const uint64_t db = ptx::make_smem_desc(b_base + k * 2048, 1024, 0);  // MN-major, one 64-column atom
              ptx::umma_f16(d_col, da, db, idesc3, k > 0 ? 1u : 0u);
            }
            ptx::umma_commit(ptx::smem_u32(wuk_empty + w));
            ptx::umma_commit(ptx::smem_u32(unit_full + b));
          }
          __syncwarp();
        }
        if (lane == 0) ptx::umma_commit(ptx::smem_u32(qn_free));  // every MMA that reads this q_nope tile has retired
        __syncwarp();
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue groups (thread = token row of tile g)
    const int g = (warp - 2) >> 2;
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int et = threadIdx.x - 64 - g * 128;
    const bool issuer = et == 0;
    const int bar_id = 1 + g;
    const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
    const uint32_t st_tile = ptx::smem_u32(smem + kSStageOff + g * kTile16K);
    const bool per_token = qa.tok_scale1 != nullptr;
    auto stage_acquire = [&]() {
      if (issuer) ptx::tma_store_wait_read<0>();
      ptx::named_bar_sync(bar_id, 128);
    };
    auto stage_publish = [&]() {
      ptx::fence_proxy_async_smem();
      ptx::named_bar_sync(bar_id, 128);
    };
    ptx::griddep_wait();  // per-token scales come from the norm kernel

    uint32_t it = 0;
    for (int u = blockIdx.x; u < total_units; u += gridDim.x, ++it) {
      const int head = u / n_st, tok0 = (u % n_st) * 256;
      const int t = tok0 + g * kBlockM + row;
      const int tc = t < qa.m_total ? t : qa.m_total - 1;  // clamped row for loads; TMA stores clip rows >= N
      // per-column constants of this head.  Both groups meet here first: nobody may still be dequantising the previous
      // unit with the old constants (free in steady state: the next MM2 needs far longer than the other group's drains).
      ptx::named_bar_sync(3, 256);
      if (g == 0) {
        for (int c = et; c < MLAPO_Q_HEAD; c += 128) {
          const float dsc = load_descale(qa.descale1, qa.descale_is_i64, head * MLAPO_Q_HEAD + c);
          s_cst[c] = make_int2(qa.bias1 ? qa.bias1[head * MLAPO_Q_HEAD + c] : 0, __float_as_int(dsc));
        }
      }
      const float ts = per_token ? qa.tok_scale1[tc] : 1.f;
      uint4 cs[8], sn[8];
      {
        const uint4* cp = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(qa.cos) + (long long)tc * 64);
        const uint4* sp = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(qa.sin) + (long long)tc * 64);
#pragma unroll
        for (int j = 0; j < 8; ++j) { cs[j] = ptx::ld_global_nc_v4(cp + j); sn[j] = ptx::ld_global_nc_v4(sp + j); }
      }
      ptx::named_bar_sync(3, 256);  // constants visible to both groups
      auto deq = [&](uint32_t acc, int c) -> float {
        const int2 cst = s_cst[c];
        return per_token ? __fmul_rn(__fmul_rn(__int2float_rn((int)acc), __int_as_float(cst.y)), ts)
                         : __fmul_rn(__int2float_rn((int)acc + cst.x), __int_as_float(cst.y));
      };

      ptx::mbar_wait(ptx::smem_u32(mm2_done), it & 1);
      __syncwarp();
      ptx::tc_fence_after();
      if (issuer && g == 0 && it < 4) tl_mark(tl, 2 + 3 * it);   // debug timeline: MM2 of unit `it` retired
      const uint32_t acc2 = lane_base + g * 256;

      // rope slice first (columns 128..191): it needs no shared q_nope bytes, so tile 1 does it while tile 0 is absorbed
      float x[64];
      {
        uint32_t ra[32], rb[32];
        ptx::tmem_ld_32x32b_x32(acc2 + 128, ra);
        ptx::tmem_ld_32x32b_x32(acc2 + 160, rb);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) x[j] = Act<T>::round(deq(ra[j], 128 + j));
#pragma unroll
        for (int j = 0; j < 32; ++j) x[32 + j] = Act<T>::round(deq(rb[j], 160 + j));
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
        for (int v = 0; v < 4; ++v) {
          const int i0 = gg * 8 + 2 * v, i1 = i0 + 1;
          // out[i] = x[i]*cos[i] + (-x[i+32])*sin[i];  out[i+32] = x[i+32]*cos[i+32] + x[i]*sin[i+32]   (test_mla:57-67)
          const float a0 = __fadd_rn(__fmul_rn(x[i0], c_lo[2 * v]), __fmul_rn(-x[i0 + 32], s_lo[2 * v]));
          const float a1 = __fadd_rn(__fmul_rn(x[i1], c_lo[2 * v + 1]), __fmul_rn(-x[i1 + 32], s_lo[2 * v + 1]));
          const float b0 = __fadd_rn(__fmul_rn(x[i0 + 32], c_hi[2 * v]), __fmul_rn(x[i0], s_hi[2 * v]));
          const float b1 = __fadd_rn(__fmul_rn(x[i1 + 32], c_hi[2 * v + 1]), __fmul_rn(x[i1], s_hi[2 * v + 1]));
          wl[v] = Act<T>::pack(a0, a1);
          wh[v] = Act<T>::pack(b0, b1);
        }
        ptx::st_shared_v4(st_tile + ptx::sw128_offset(row, gg), wl[0], wl[1], wl[2], wl[3]);
        ptx::st_shared_v4(st_tile + ptx::sw128_offset(row, gg + 4), wh[0], wh[1], wh[2], wh[3]);
      }
      stage_publish();
      if (issuer) {
        ptx::tma_store_2d(&tm_out1, st_tile, head * MLAPO_ROPE, tok0 + g * kBlockM);
        ptx::tma_store_commit();
      }

      // nope slice: dequantised and packed in registers first, so that acc2[g] is released to the next unit's MM2 without
      // waiting for the shared q_nope tile
      uint32_t wn[64];
      {
        auto nope_chunk = [&](const uint32_t (&r)[32], int c0) {
#pragma unroll
          for (int j = 0; j < 16; ++j) wn[(c0 >> 1) + j] = Act<T>::pack(deq(r[2 * j], c0 + 2 * j), deq(r[2 * j + 1], c0 + 2 * j + 1));
        };
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
        nope_chunk(rb, 96);
      }
      ptx::tc_fence_before();
      ptx::mbar_arrive(ptx::smem_u32(acc2_free));      // acc2[g] fully read (both groups arrive: 256)
      if (issuer && g == 1 && it < 4) tl_mark(tl, 3 + 3 * it);   // debug timeline: accumulators of unit `it` released
      // the shared q_nope tile is free once the absorb of the previous token tile has retired
      // (qn_free completes twice per unit: after tile 0 and after tile 1; tile 0 waits for the previous unit's second one)
      if (g == 1) ptx::mbar_wait(ptx::smem_u32(qn_free), 0);
      else if (it > 0) ptx::mbar_wait(ptx::smem_u32(qn_free), 1);
      __syncwarp();
#pragma unroll
      for (int j = 0; j < 16; ++j)  // 16-byte chunk j of the 256-byte row: k-block j/8, chunk j%8
        ptx::st_shared_v4(ptx::smem_u32(smem + kSNopeOff + (j >> 3) * kTile16K) + ptx::sw128_offset(row, j & 7), wn[4 * j], wn[4 * j + 1],
                          wn[4 * j + 2], wn[4 * j + 3]);
      ptx::fence_proxy_async_smem();
      ptx::mbar_arrive(ptx::smem_u32(qn_ready + g));   // q_nope rows of tile g written

      // drains: this group's 8 absorb units, in the global order the absorb warp issues them
      float qs = 1.f;
      if (kInt8Out) qs = Act<T>::to_f(reinterpret_cast<const T*>(qa.q_nope_scale)[head]);
#pragma unroll 1
      for (int cp = 0; cp < 4; ++cp) {
        uint32_t w8[32];  // int8 output: two 64-column units fill one 128-byte row per token
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int c = cp * 2 + half;
          const uint32_t gd = (it * 2 + g) * 8 + c, b = gd & 1;
          ptx::mbar_wait(ptx::smem_u32(unit_full + b), (gd >> 1) & 1);
          __syncwarp();
          ptx::tc_fence_after();
          uint32_t ra[32], rb[32];
          const uint32_t acc3 = lane_base + (b ? 448u : 192u);
          ptx::tmem_ld_32x32b_x32(acc3, ra);
          ptx::tmem_ld_32x32b_x32(acc3 + 32, rb);
          ptx::tmem_ld_wait();
          ptx::tc_fence_before();
          ptx::mbar_arrive(ptx::smem_u32(abuf_empty + b));
          if (kInt8Out) {
            // S9: int8( fp16( act(acc) * q_nope_scale[h] ) )   (golden test_mla:86-92, 466-470)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              int qa4[4], qb4[4];
#pragma unroll
              for (int v = 0; v < 4; ++v) {
                qa4[v] = sat_i8(__fmul_rn(Act<T>::round(__uint_as_float(ra[4 * j + v])), qs));
                qb4[v] = sat_i8(__fmul_rn(Act<T>::round(__uint_as_float(rb[4 * j + v])), qs));
              }
              w8[half * 16 + j] = pack4_i8(qa4[0], qa4[1], qa4[2], qa4[3]);
              w8[half * 16 + 8 + j] = pack4_i8(qb4[0], qb4[1], qb4[2], qb4[3]);
            }
          } else {
            uint32_t w[32];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              w[j] = Act<T>::pack(__uint_as_float(ra[2 * j]), __uint_as_float(ra[2 * j + 1]));
              w[16 + j] = Act<T>::pack(__uint_as_float(rb[2 * j]), __uint_as_float(rb[2 * j + 1]));
            }
            stage_acquire();
#pragma unroll
            for (int j = 0; j < 8; ++j) ptx::st_shared_v4(st_tile + ptx::sw128_offset(row, j), w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
            stage_publish();
            if (issuer) {
              ptx::tma_store_2d(&tm_out0, st_tile, head * MLAPO_KV_LORA + c * 64, tok0 + g * kBlockM);
              ptx::tma_store_commit();
            }
          }
        }
        if (kInt8Out) {
          stage_acquire();
#pragma unroll
          for (int j = 0; j < 8; ++j) ptx::st_shared_v4(st_tile + ptx::sw128_offset(row, j), w8[4 * j], w8[4 * j + 1], w8[4 * j + 2], w8[4 * j + 3]);
          stage_publish();
          if (issuer) {
            ptx::tma_store_2d(&tm_out0, st_tile, head * MLAPO_KV_LORA + cp * 128, tok0 + g * kBlockM);
            ptx::tma_store_commit();
          }
        }
      }
      if (issuer && g == 1 && it < 4) tl_mark(tl, 4 + 3 * it);   // debug timeline: unit `it` fully drained
    }
    if (issuer) ptx::tma_store_wait_read<0>();
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem, 512);
}

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

// Grid barrier state {arrivals, generation}.  Self-cleaning (arrivals return to 0), so it is valid from module load on;
// one per device, which is enough because cooperative launches of this kernel on one device are stream-ordered.
__device__ unsigned g_grid_sync_state[2];

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// All CTAs of the (co-resident) grid.  Global writes made before it by any thread — through the generic proxy or by
// completed TMA stores — are visible afterwards to generic loads and to TMA loads of every CTA.
__device__ __forceinline__ void grid_sync(unsigned nblocks) {
  fence_proxy_async_all();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned* st = g_grid_sync_state;
    const unsigned gen = ld_acquire_gpu(st + 1);
    __threadfence();
    if (atomicAdd(st, 1u) == nblocks - 1) {
      atomicExch(st, 0u);
      __threadfence();
      atomicAdd(st + 1, 1u);
    } else {
      const long long t0 = clock64();
      while (ld_acquire_gpu(st + 1) == gen) {
        __nanosleep(32);
        const long long dt = clock64() - t0;
        if (dt > 4000000000LL) ptx::wait_timeout(blockIdx.x, 0xFFFFu, gen, 0);  // a missing participant becomes a launch failure, not a hang
      }
    }
    __threadfence();
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
  grid_sync(nblocks);

  // ---- phase 2: MM1 split-K partial sums
  if (threadIdx.x == 0) tl_mark(da.tl_mm1, 2);
  if (da.stop_after == 1) {  // debug exit: the requested weight tiles must still land before the CTA leaves
    if (has_mm1 && warp == 1) for (int i = 0; i < (mc.nkb < M1::STAGES ? mc.nkb : M1::STAGES); ++i) (void)i;
  }
  if (has_mm1 && da.stop_after != 1) mm1_roles<MT, kMm1BlockN, false>(mc, &tm_a0, &tm_wdqkv, &tm_part, warp, lane, da.tl_mm1);
  if (threadIdx.x == 64 && has_mm1) ptx::tma_store_wait<0>();  // partial tiles fully written, not just read out of smem
  ptx::tc_fence_before();
  __syncthreads();                                              // the MM1 pipeline bytes are free again
  ptx::tc_fence_after();
  if (threadIdx.x == 32) tl_mark(da.tl_mm1, 9);
  // first W_UQ k-blocks of this CTA's head: requested now, they land while the norm phase runs
  if (warp == 0 && lane == 0 && has_q && (da.stop_after == 0 || da.stop_after > 3)) q_issue_weights(qc, &tm_wuq);
  grid_sync(nblocks);
  if (threadIdx.x == 0) tl_mark(da.tl_mm1, 11);  // barrier 2 passed
  const bool run3 = da.stop_after == 0 || da.stop_after >= 3;

  // ---- phase 3: norms, quant, RoPE, cache scatter (one token at a time per CTA)
  for (int t = cta; run3 && t < da.n_tokens; t += (int)nblocks) {
    norm_token<T, 8>(da.norm, t, threadIdx.x);
    __syncthreads();  // the reduction scratch is re-used by the next token
  }
  if (threadIdx.x == 0) tl_mark(da.tl_mm1, 12);  // norm done, entering barrier 3
  grid_sync(nblocks);

  // ---- phase 4: per-head q path
  if (threadIdx.x == 0) { tl_mark_global(da.tl_q, 0); tl_mark(da.tl_q, 1); tl_mark(da.tl_q, 2); }
  if (has_q && da.stop_after == 0) q_roles<T, kInt8Out, false>(qc, &tm_a1, &tm_wuq, &tm_wuk, &tm_out0, &tm_out1, da.q, warp, lane, da.tl_q);

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(qc.tmem, 512);
  if (threadIdx.x == 32) { tl_mark(da.tl_q, 12); tl_mark_global(da.tl_q, 15); tl_mark_global(da.tl_mm1, 15); }
}

// ============================================================================================ host side
thread_local char g_err[512] = "";
static int g_debug_flags = 0;  // bit0: sync + check after each kernel; bit1: write the y0 tap; bit2: time each kernel
static unsigned* g_wait_log_host = nullptr;  // host-mapped log of stuck barrier waits (debug flag bit11)
constexpr int kMaxKernels = 8;
static cudaEvent_t g_ev[kMaxKernels + 1];
static bool g_ev_ready = false;
static int g_ev_count = 0;  // events recorded by the last launch
static const char* g_kernel_names[kMaxKernels] = {"quant_hidden", "mm1_splitk", "norm_rope_cache", "q_fused", "", "", "", ""};
static void record_event(int i, cudaStream_t st) {
  if (!(g_debug_flags & 4)) return;
  if (!g_ev_ready) {
    for (int k = 0; k <= kMaxKernels; ++k) cudaEventCreate(&g_ev[k]);
    g_ev_ready = true;
  }
  cudaEventRecord(g_ev[i], st);
  g_ev_count = i + 1;
}

static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;
static std::once_flag g_encode_once;

static EncodeTiledFn get_encode() {
  std::call_once(g_encode_once, [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      g_encode = reinterpret_cast<EncodeTiledFn>(fn);
  });
  return g_encode;
}

// Row-major tensor [d2][rows][cols] (d2 = 1 for rank 2) of `elem_bytes` elements, box [1][box_rows][box_cols] with a
// 128-byte inner extent and 128B swizzle.
static int make_tmap(CUtensorMap* m, CUtensorMapDataType dt, int elem_bytes, const void* base, uint64_t rows, uint64_t cols,
                     uint64_t row_stride_bytes, uint32_t box_rows, uint32_t box_cols, uint64_t d2 = 1, uint64_t d2_stride_bytes = 0) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return fail(MLAPO_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not available");
  if (box_cols * elem_bytes != 128) return fail(MLAPO_ERR_INVALID_ARG, "tensor map inner box must be 128 bytes");
  const cuuint32_t rank = d2_stride_bytes ? 3 : 2;
  cuuint64_t gdim[3] = {cols, rows, d2};
  cuuint64_t gstr[2] = {row_stride_bytes, d2_stride_bytes};
  cuuint32_t box[3] = {box_cols, box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(m, dt, rank, const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(MLAPO_ERR_DRIVER, "cuTensorMapEncodeTiled failed (%d) rows=%llu cols=%llu", (int)r,
                                     (unsigned long long)rows, (unsigned long long)cols);
  return MLAPO_OK;
}

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static int num_sms() {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  return sms;
}

// Launch with the programmatic-dependent-launch attribute: the kernel may start while its predecessor drains;
// every kernel of this file calls griddepcontrol.wait before touching anything its predecessors produce or consume.
template <class... KArgs, class... Args>
static cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, bool pdl, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

static int choose_splits0(int n_tokens, int hidden_dim) {
  const int m_super = n_tokens > 0 ? (n_tokens + 2 * kBlockM - 1) / (2 * kBlockM) : 1;  // CTAs cover up to 256 tokens
  const int n_tiles = (MLAPO_MM1_OUT + kMm1BlockN - 1) / kMm1BlockN;                     // 17
  const int kb_total = hidden_dim / kBlockKBytes;
  int s = 148 / (n_tiles * m_super);  // one wave of CTAs
  if (s < 1) s = 1;
  if (s > 8) s = 8;
  if (s > kb_total) s = kb_total;
  return s;
}

static int check_params(const MlapoParams* p) {
  if (!p) return fail(MLAPO_ERR_INVALID_ARG, "params is null");
  if (p->struct_bytes != sizeof(MlapoParams))
    return fail(MLAPO_ERR_INVALID_ARG, "MlapoParams.struct_bytes=%u, library expects %zu", p->struct_bytes, sizeof(MlapoParams));
  if (p->num_tokens < 0 || p->num_heads <= 0) return fail(MLAPO_ERR_INVALID_ARG, "bad num_tokens/num_heads");
  if (p->hidden_dim <= 0 || p->hidden_dim % 128 != 0) return fail(MLAPO_ERR_INVALID_ARG, "hidden_dim must be a positive multiple of 128");
  if (p->act_dtype != MLAPO_ACT_BF16 && p->act_dtype != MLAPO_ACT_FP16) return fail(MLAPO_ERR_INVALID_ARG, "bad act_dtype");
  if (p->cache_mode < 1 || p->cache_mode > 3) return fail(MLAPO_ERR_INVALID_ARG, "Unsupported cache_mode value: %d", p->cache_mode);
  if (p->quant_mode < 0 || p->quant_mode > 1) return fail(MLAPO_ERR_INVALID_ARG, "Unsupported quant_mode value: %d", p->quant_mode);
  if (p->cache_mode != MLAPO_CACHE_KROPE_CTKV && p->cache_block_size != 128)
    return fail(MLAPO_ERR_INVALID_ARG, "NZ cache modes require cache_block_size == 128");
  if (p->num_cache_blocks <= 0 || p->cache_block_size <= 0) return fail(MLAPO_ERR_INVALID_ARG, "bad cache geometry");
  return MLAPO_OK;
}

static int compute_layout(const MlapoParams* p, MlapoWorkspaceLayout* l) {
  const size_t n = (size_t)p->num_tokens, d = (size_t)p->hidden_dim;
  const int splits = choose_splits0(p->num_tokens, p->hidden_dim);
  size_t off = 0;
  l->a0 = off;        off = align_up(off + n * d, 256);
  l->partial0 = off;  off = align_up(off + (size_t)splits * n * MLAPO_MM1_OUT * 4, 256);
  l->a1 = off;        off = align_up(off + n * MLAPO_Q_LORA, 256);
  l->tok_scale = off; off = align_up(off + 2 * n * 4, 256);
  l->y0_tap = off;
  if (g_debug_flags & 2) off += align_up(n * MLAPO_MM1_OUT * 2, 256);
  l->timeline = off;   // flag bit3: 4 kernels (mm1, q_fused, quant, norm) x 4096 CTAs x 16 marks
  if (g_debug_flags & 8) off += 4 * 4096 * 16 * 8;
  l->total_bytes = off;
  l->splits0 = splits;
  l->reserved = 0;
  return MLAPO_OK;
}

#define MLAPO_CHECK_LAUNCH(name)                                                                    \
  do {                                                                                              \
    if (e__ == cudaSuccess) e__ = cudaGetLastError();                                               \
    if (e__ == cudaSuccess && (g_debug_flags & 1)) e__ = cudaStreamSynchronize(st);                 \
    if (e__ != cudaSuccess) return fail(MLAPO_ERR_CUDA, "%s: %s", name, cudaGetErrorString(e__));   \
  } while (0)

template <class K>
static int set_smem_attr(K kern, int bytes, bool* done, const char* name) {
  if (*done) return MLAPO_OK;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess)
    return fail(MLAPO_ERR_CUDA, "cudaFuncSetAttribute(%s, %d): %s", name, bytes, cudaGetErrorString(cudaGetLastError()));
  *done = true;
  return MLAPO_OK;
}

// Token rows per q-path CTA: two token tiles per CTA stream each head's weights once per 256 tokens; when that leaves most
// SMs idle (few heads, e.g. a TP slice) one tile per CTA doubles the CTA count at the price of re-reading the head's
// weights from L2.
static int q_cta_tokens(int n, int h) { return (n > kBlockM && (long long)h * ((n + 255) / 256) >= 96) ? 256 : 128; }

// Decode-sized calls CAN run as one persistent cooperative kernel (k5) when every phase fits a single wave of CTAs.
// It is opt-in (flag bit5): measured on B200 at C2 it takes 52.4 us against 38.5 us for the four PDL-chained kernels —
// each of its three grid barriers costs 2-3 us and the norm phase handles a CTA's tokens one after the other
// (profiles/README.md, "persistent decode kernel") — so the chain stays the default until those two costs are fixed.
static bool decode_persistent_eligible(const MlapoParams* p, int splits0) {
  if (!(g_debug_flags & 32)) return false;  // bit5: use the persistent decode kernel where eligible
  const int n = p->num_tokens, h = p->num_heads;
  if (n < 1 || n > 256) return false;
  const int ct = q_cta_tokens(n, h);
  return kMm1NTiles * splits0 <= num_sms() && h * ((n + ct - 1) / ct) <= num_sms();
}

template <class T>
static int launch_impl(const MlapoParams* p, uint8_t* ws, const MlapoWorkspaceLayout& l, cudaStream_t st) {
  const int n = p->num_tokens, d = p->hidden_dim, h = p->num_heads;
  const bool per_token = p->quant_mode == MLAPO_QUANT_PER_TOKEN_SYMM;
  const bool int8_out = p->cache_mode == MLAPO_CACHE_INT8_NZ;
  const bool pdl = (g_debug_flags & 5) == 0;  // per-kernel sync/timing wants serialised kernels
  long long* tl1 = (g_debug_flags & 8) ? reinterpret_cast<long long*>(ws + l.timeline) : nullptr;
  long long* tl4 = tl1 ? tl1 + 4096 * 16 : nullptr;
  long long* tl0 = tl1 ? tl1 + 2 * 4096 * 16 : nullptr;
  long long* tl3 = tl1 ? tl1 + 3 * 4096 * 16 : nullptr;
  int8_t* a0 = reinterpret_cast<int8_t*>(ws + l.a0);
  int32_t* partial0 = reinterpret_cast<int32_t*>(ws + l.partial0);
  int8_t* a1 = reinterpret_cast<int8_t*>(ws + l.a1);
  float* tok_scale = reinterpret_cast<float*>(ws + l.tok_scale);
  const CUtensorMapDataType act_dt = std::is_same<T, __nv_bfloat16>::value ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  int rc;
  cudaError_t e__ = cudaSuccess;
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  const bool persistent = decode_persistent_eligible(p, l.splits0);

  // ---- tensor maps and argument blocks shared by both launch shapes
  const bool wide = !persistent && l.splits0 == 1 && n >= 1024;  // prefill: no split-K, 256-wide weight tiles
  const int bn = wide ? kMm1BlockNWide : kMm1BlockN;
  CUtensorMap tm_a0, tm_wdqkv, tm_part, tm_a1, tm_wuq, tm_wuk, tm_o0, tm_o1;
  if ((rc = make_tmap(&tm_a0, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a0, n, d, d, kBlockM, 128))) return rc;
  if ((rc = make_tmap(&tm_wdqkv, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, p->wdqkv_km, MLAPO_MM1_OUT, d, d, bn, 128))) return rc;
  if ((rc = make_tmap(&tm_part, CU_TENSOR_MAP_DATA_TYPE_INT32, 4, partial0, n, MLAPO_MM1_OUT, MLAPO_MM1_OUT * 4, kBlockM, 32, l.splits0,
                      (uint64_t)n * MLAPO_MM1_OUT * 4))) return rc;
  const int cta_tokens = q_cta_tokens(n, h);
  const int a1_box_rows = cta_tokens > kBlockM ? 2 * kBlockM : kBlockM;  // one box per stage covers every token tile of the CTA
  if ((rc = make_tmap(&tm_a1, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a1, n, MLAPO_Q_LORA, MLAPO_Q_LORA, a1_box_rows, 128))) return rc;
  if ((rc = make_tmap(&tm_wuq, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, p->wuq_km, (uint64_t)h * MLAPO_Q_HEAD, MLAPO_Q_LORA, MLAPO_Q_LORA,
                      MLAPO_Q_HEAD, 128))) return rc;
  if ((rc = make_tmap(&tm_wuk, act_dt, 2, p->wuk, (uint64_t)h * MLAPO_Q_NOPE, MLAPO_KV_LORA, MLAPO_KV_LORA * 2, 128, 64))) return rc;
  if (int8_out) {
    if ((rc = make_tmap(&tm_o0, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, p->q_out0, n, (uint64_t)h * MLAPO_KV_LORA, (uint64_t)h * MLAPO_KV_LORA,
                        kBlockM, 128))) return rc;
  } else {
    if ((rc = make_tmap(&tm_o0, act_dt, 2, p->q_out0, n, (uint64_t)h * MLAPO_KV_LORA, (uint64_t)h * MLAPO_KV_LORA * 2, kBlockM, 64))) return rc;
  }
  if ((rc = make_tmap(&tm_o1, act_dt, 2, p->q_out1, n, (uint64_t)h * MLAPO_ROPE, (uint64_t)h * MLAPO_ROPE * 2, kBlockM, 64))) return rc;

  NormArgs na{};
  na.partial = partial0; na.splits = l.splits0; na.n_tokens = n;
  na.bias0 = per_token ? nullptr : reinterpret_cast<const int32_t*>(p->bias0);
  na.descale0 = p->descale0; na.descale_is_i64 = p->descale_is_int64;
  na.tok_scale0 = per_token ? tok_scale : nullptr;
  na.tok_scale1 = per_token ? tok_scale + n : nullptr;
  na.gamma1 = p->gamma1; na.beta1 = p->beta1; na.gamma2 = p->gamma2;
  na.quant_scale1 = p->quant_scale1; na.quant_offset1 = reinterpret_cast<const int8_t*>(p->quant_offset1);
  na.cos = p->cos; na.sin = p->sin;
  na.slots = reinterpret_cast<const int32_t*>(p->slot_mapping);
  na.ctkv_scale = p->ctkv_scale;
  na.a1 = a1;
  na.y0_tap = (g_debug_flags & 2) ? ws + l.y0_tap : nullptr;
  na.kv_cache = p->kv_cache; na.kv_cache_rope = p->kv_cache_rope;
  na.cache_mode = p->cache_mode;
  na.cache_slots = (long long)p->num_cache_blocks * p->cache_block_size;
  na.tl = tl3;

  QArgs qa{};
  qa.descale1 = p->descale1; qa.descale_is_i64 = p->descale_is_int64;
  qa.bias1 = per_token ? nullptr : reinterpret_cast<const int32_t*>(p->bias1);
  qa.tok_scale1 = per_token ? tok_scale + n : nullptr;
  qa.cos = p->cos; qa.sin = p->sin; qa.q_nope_scale = p->q_nope_scale;
  qa.num_heads = h; qa.m_total = n; qa.cta_tokens = cta_tokens;
  const int q_tok_tiles = (n + cta_tokens - 1) / cta_tokens;

  if (persistent) {
    // ---- k5: one cooperative launch, one CTA per SM
    DecodeArgs da{};
    da.hidden = p->hidden_state; da.a0 = a0;
    da.quant_scale0 = p->quant_scale0; da.quant_offset0 = reinterpret_cast<const int8_t*>(p->quant_offset0);
    da.tok_scale0 = per_token ? tok_scale : nullptr;
    da.hidden_dim = d; da.n_tokens = n; da.splits0 = l.splits0; da.q_tok_tiles = q_tok_tiles;
    da.norm = na; da.q = qa; da.tl_mm1 = tl1; da.tl_q = tl4;
    da.stop_after = (g_debug_flags >> 8) & 7;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(num_sms());
    cfg.blockDim = dim3(kQThreads);
    cfg.dynamicSmemBytes = kDecodeSmemBytes;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeCooperative;  // the grid barriers need every CTA resident
    at[0].val.cooperative = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    record_event(0, st);
#define MLAPO_LAUNCH_DECODE(I8, MT)                                                                                        \
  do {                                                                                                                     \
    static bool done[64] = {};                                                                                             \
    if ((rc = set_smem_attr(mla_decode_persistent_kernel<T, I8, MT>, kDecodeSmemBytes, &done[dev], "decode_persistent"))) return rc; \
    e__ = cudaLaunchKernelEx(&cfg, mla_decode_persistent_kernel<T, I8, MT>, tm_a0, tm_wdqkv, tm_part, tm_a1, tm_wuq, tm_wuk, tm_o0,  \
                             tm_o1, da);                                                                                   \
  } while (0)
    if (n > kBlockM) {
      if (int8_out) MLAPO_LAUNCH_DECODE(true, 2); else MLAPO_LAUNCH_DECODE(false, 2);
    } else {
      if (int8_out) MLAPO_LAUNCH_DECODE(true, 1); else MLAPO_LAUNCH_DECODE(false, 1);
    }
#undef MLAPO_LAUNCH_DECODE
    MLAPO_CHECK_LAUNCH("decode_persistent");
    record_event(1, st);
    g_kernel_names[0] = "decode_persistent";
    return MLAPO_OK;
  }
  g_kernel_names[0] = "quant_hidden";

  // ---- k1: quantise hiddenState
  record_event(0, st);
  if (per_token) {
    e__ = launch_k(quant_hidden_per_token_kernel<T>, dim3(n), dim3(256), 0, st, pdl, reinterpret_cast<const T*>(p->hidden_state), a0,
                   tok_scale, d);
  } else {
    const long long nvec = (long long)n * d / 8;
    const long long one_wave = (long long)num_sms() * 8 * 256;  // threads of one full wave (8 blocks of 256 per SM)
    if (nvec <= one_wave) {  // one 16-byte vector per thread
      e__ = launch_k(quant_hidden_kernel<T, 1>, dim3((unsigned)((nvec + 255) / 256)), dim3(256), 0, st, pdl,
                     reinterpret_cast<const T*>(p->hidden_state), a0, reinterpret_cast<const T*>(p->quant_scale0),
                     reinterpret_cast<const int8_t*>(p->quant_offset0), nvec, tl0);
    } else {                 // prefill sizes: grid-stride, 4 independent loads in flight per thread
      e__ = launch_k(quant_hidden_kernel<T, 4>, dim3(num_sms() * 8), dim3(256), 0, st, pdl,
                     reinterpret_cast<const T*>(p->hidden_state), a0, reinterpret_cast<const T*>(p->quant_scale0),
                     reinterpret_cast<const int8_t*>(p->quant_offset0), nvec, tl0);
    }
  }
  MLAPO_CHECK_LAUNCH("quant_hidden");
  record_event(1, st);

  // ---- k2: MM1, split-K partial sums
  {
    const int n_tiles = (MLAPO_MM1_OUT + bn - 1) / bn;
    if (wide) {
      static bool done[64] = {};
      if ((rc = set_smem_attr(mm1_splitk_kernel<2, kMm1BlockNWide>, mm1_smem_bytes<2, kMm1BlockNWide>(), &done[dev], "mm1<2,256>"))) return rc;
      e__ = launch_k(mm1_splitk_kernel<2, kMm1BlockNWide>, dim3(n_tiles, 1, (n + 255) / 256), dim3(kGemmThreads),
                     mm1_smem_bytes<2, kMm1BlockNWide>(), st, pdl, tm_a0, tm_wdqkv, tm_part, n, d / kBlockKBytes, tl1);
    } else if (n > kBlockM) {
      static bool done[64] = {};
      if ((rc = set_smem_attr(mm1_splitk_kernel<2, kMm1BlockN>, mm1_smem_bytes<2, kMm1BlockN>(), &done[dev], "mm1<2,128>"))) return rc;
      e__ = launch_k(mm1_splitk_kernel<2, kMm1BlockN>, dim3(n_tiles, l.splits0, (n + 255) / 256), dim3(kGemmThreads),
                     mm1_smem_bytes<2, kMm1BlockN>(), st, pdl, tm_a0, tm_wdqkv, tm_part, n, d / kBlockKBytes, tl1);
    } else {
      static bool done[64] = {};
      if ((rc = set_smem_attr(mm1_splitk_kernel<1, kMm1BlockN>, mm1_smem_bytes<1, kMm1BlockN>(), &done[dev], "mm1<1,128>"))) return rc;
      e__ = launch_k(mm1_splitk_kernel<1, kMm1BlockN>, dim3(n_tiles, l.splits0, 1), dim3(kGemmThreads), mm1_smem_bytes<1, kMm1BlockN>(),
                     st, pdl, tm_a0, tm_wdqkv, tm_part, n, d / kBlockKBytes, tl1);
    }
    MLAPO_CHECK_LAUNCH("mm1_splitk");
  }
  record_event(2, st);

  // ---- k3: norms, quant, rope, cache scatter
  if (l.splits0 == 1) e__ = launch_k(norm_rope_cache_kernel<T, 1>, dim3(n), dim3(kNormThreads), 0, st, pdl, na);  // prefill: no split-K
  else e__ = launch_k(norm_rope_cache_kernel<T, 8>, dim3(n), dim3(kNormThreads), 0, st, pdl, na);
  MLAPO_CHECK_LAUNCH("norm_rope_cache");
  record_event(3, st);

  // ---- k4 / k6: fused MM2 + dequant + RoPE + W_UK absorb.  One CTA per (head, token super-tile) while those fit one wave;
  //      beyond that, persistent CTAs stream the units so that consecutive units overlap.
  const int q_units = h * ((n + 255) / 256);
  if (n > 256 && q_units > num_sms() && !(g_debug_flags & 64)) {  // bit6: keep k4 for A/B comparisons
    int rcx;
    CUtensorMap tm_a1s;  // the streaming kernel always fetches both token tiles of a unit in one box
    if ((rcx = make_tmap(&tm_a1s, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a1, n, MLAPO_Q_LORA, MLAPO_Q_LORA, 2 * kBlockM, 128))) return rcx;
    dim3 grid(num_sms() < q_units ? num_sms() : q_units);
    if (int8_out) {
      static bool done[64] = {};
      if ((rc = set_smem_attr(q_stream_kernel<T, true>, kSSmemBytes, &done[dev], "q_stream<int8>"))) return rc;
      e__ = launch_k(q_stream_kernel<T, true>, grid, dim3(kSThreads), kSSmemBytes, st, pdl, tm_a1s, tm_wuq, tm_wuk, tm_o0, tm_o1, qa, q_units, tl4);
    } else {
      static bool done[64] = {};
      if ((rc = set_smem_attr(q_stream_kernel<T, false>, kSSmemBytes, &done[dev], "q_stream"))) return rc;
      e__ = launch_k(q_stream_kernel<T, false>, grid, dim3(kSThreads), kSSmemBytes, st, pdl, tm_a1s, tm_wuq, tm_wuk, tm_o0, tm_o1, qa, q_units, tl4);
    }
    MLAPO_CHECK_LAUNCH("q_stream");
    g_kernel_names[3] = "q_stream";
  } else {
    g_kernel_names[3] = "q_fused";
    dim3 grid(h, q_tok_tiles);
    if (int8_out) {
      static bool done[64] = {};
      if ((rc = set_smem_attr(q_fused_kernel<T, true>, kQSmemBytes, &done[dev], "q_fused<int8>"))) return rc;
      e__ = launch_k(q_fused_kernel<T, true>, grid, dim3(kQThreads), kQSmemBytes, st, pdl, tm_a1, tm_wuq, tm_wuk, tm_o0, tm_o1, qa, tl4);
    } else {
      static bool done[64] = {};
      if ((rc = set_smem_attr(q_fused_kernel<T, false>, kQSmemBytes, &done[dev], "q_fused"))) return rc;
      e__ = launch_k(q_fused_kernel<T, false>, grid, dim3(kQThreads), kQSmemBytes, st, pdl, tm_a1, tm_wuq, tm_wuk, tm_o0, tm_o1, qa, tl4);
    }
    MLAPO_CHECK_LAUNCH("q_fused");
  }
  record_event(4, st);
  return MLAPO_OK;
}

}  // namespace mlapo

// ============================================================================================ C ABI
using namespace mlapo;

// Scratch for mla_preprocess_launch_static(): lives in the module's global memory on every device that loads the library.
// In host code this name is only a handle for cudaGetSymbolAddress(); it is not a device address.
static __device__ __align__(256) unsigned char g_static_scratch[MLAPO_STATIC_SCRATCH_BYTES];
static void* g_static_scratch_addr[64];  // device address of the array, per device ordinal, resolved on first use
static std::mutex g_static_scratch_mutex;

extern "C" {

const char* mlapo_version(void) { return "mla_preprocess_b200 0.1.0 (sm_100a; tcgen05+TMA; reference sgl-kernel-npu f425bdf1)"; }
const char* mlapo_last_error(void) { return g_err; }

int mlapo_set_debug_flags(int flags) {
  const int old = g_debug_flags;
  g_debug_flags = flags;
  if ((flags & 2048) && !g_wait_log_host) {  // bit11: register a host-mapped log for stuck barrier waits
    unsigned* dptr = nullptr;
    if (cudaHostAlloc(reinterpret_cast<void**>(&g_wait_log_host), 4096, cudaHostAllocMapped) == cudaSuccess &&
        cudaHostGetDevicePointer(reinterpret_cast<void**>(&dptr), g_wait_log_host, 0) == cudaSuccess) {
      memset(g_wait_log_host, 0, 4096);
      cudaMemcpyToSymbol(ptx::g_wait_log, &dptr, sizeof dptr);
    }
  }
  return old;
}
// Debug: copies the stuck-wait log {count, then {cta, thread, barrier, parity} x count}; returns the count.  Reads host
// memory only, so it still works after a trap.
int mlapo_debug_read_wait_log(unsigned* out, int cap_words) {
  if (!g_wait_log_host) return -1;
  for (int i = 0; i < cap_words && i < 1024; ++i) out[i] = g_wait_log_host[i];
  return (int)g_wait_log_host[0];
}
// Debug/profiling hook (flag bit2): per-kernel durations of the last mla_preprocess_launch on this thread's device.
int mlapo_get_kernel_times_ms(float* out_ms, int cap) {
  if (!g_ev_ready || g_ev_count < 2) return 0;
  if (cudaEventSynchronize(g_ev[g_ev_count - 1]) != cudaSuccess) return fail(MLAPO_ERR_CUDA, "event sync failed");
  int n = g_ev_count - 1;
  if (n > cap) n = cap;
  for (int i = 0; i < n; ++i) cudaEventElapsedTime(&out_ms[i], g_ev[i], g_ev[i + 1]);
  return n;
}
const char* mlapo_kernel_name(int i) { return (i >= 0 && i < kMaxKernels) ? g_kernel_names[i] : ""; }

int mlapo_get_workspace_layout(const MlapoParams* p, MlapoWorkspaceLayout* layout) {
  int rc = check_params(p);
  if (rc) return rc;
  if (!layout) return fail(MLAPO_ERR_INVALID_ARG, "layout is null");
  return compute_layout(p, layout);
}

int mlapo_get_workspace_size(const MlapoParams* p, size_t* bytes) {
  MlapoWorkspaceLayout l;
  int rc = mlapo_get_workspace_layout(p, &l);
  if (rc) return rc;
  if (!bytes) return fail(MLAPO_ERR_INVALID_ARG, "bytes is null");
  *bytes = l.total_bytes;
  return MLAPO_OK;
}

int mlapo_launches_per_call(const MlapoParams* p) {
  if (check_params(p)) return 0;
  return decode_persistent_eligible(p, choose_splits0(p->num_tokens, p->hidden_dim)) ? 1 : 4;
}

int mlapo_repack_nz_weight(const void* w_nz, void* w_km, int32_t k, int32_t n_out, void* stream) {
  if (!w_nz || !w_km) return fail(MLAPO_ERR_INVALID_ARG, "repack: null pointer");
  if (k <= 0 || k % 32 != 0 || n_out <= 0) return fail(MLAPO_ERR_INVALID_ARG, "repack: k must be a multiple of 32");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long chunks = (long long)n_out * (k / 32) * 2;
  int blocks = (int)((chunks + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  repack_nz_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const int8_t*>(w_nz), reinterpret_cast<int8_t*>(w_km), k, n_out);
  cudaError_t e__ = cudaSuccess;
  MLAPO_CHECK_LAUNCH("repack_nz");
  return MLAPO_OK;
}

// Operator call for hosts that cannot allocate (see the header): the workspace is the library's own fixed device array.
// Everything the kernels read from the workspace (a0, the split-K partial sums, a1, the per-token scales) is produced
// earlier in the same call for exactly the rows and columns that are read, so stale contents from the previous call are
// never observed and nothing is cleared here.  Calls that overlap on the GPU share the array and would race.
int mla_preprocess_launch_static(const MlapoParams* p, void* stream) {
  int rc = check_params(p);
  if (rc) return rc;
  if (p->num_tokens == 0) return MLAPO_OK;
  MlapoWorkspaceLayout l;
  compute_layout(p, &l);
  if (p->num_tokens > MLAPO_STATIC_MAX_TOKENS || p->hidden_dim > MLAPO_STATIC_MAX_HIDDEN || l.total_bytes > (size_t)MLAPO_STATIC_SCRATCH_BYTES)
    return fail(MLAPO_ERR_WORKSPACE,
                "mla_preprocess_launch_static: %d tokens x hidden %d need %zu workspace bytes; the built-in scratch holds %u bytes "
                "(at most %d tokens, hidden <= %d) - use mla_preprocess_launch() with your own workspace",
                p->num_tokens, p->hidden_dim, l.total_bytes, (unsigned)MLAPO_STATIC_SCRATCH_BYTES, MLAPO_STATIC_MAX_TOKENS,
                MLAPO_STATIC_MAX_HIDDEN);
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(MLAPO_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= 64) return fail(MLAPO_ERR_UNSUPPORTED, "device ordinal %d out of range for the built-in scratch table", dev);
  void* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
    if (!g_static_scratch_addr[dev]) {  // first call on this device: resolve the array's address in this device's module image
      void* addr = nullptr;
      e = cudaGetSymbolAddress(&addr, g_static_scratch);
      if (e != cudaSuccess || !addr)
        return fail(MLAPO_ERR_CUDA, "cudaGetSymbolAddress(built-in scratch) on device %d: %s", dev, cudaGetErrorString(e));
      g_static_scratch_addr[dev] = addr;
    }
    scratch = g_static_scratch_addr[dev];
  }
  return mla_preprocess_launch(p, scratch, (size_t)MLAPO_STATIC_SCRATCH_BYTES, stream);
}

int mla_preprocess_launch(const MlapoParams* p, void* workspace, size_t workspace_bytes, void* stream) {
  int rc = check_params(p);
  if (rc) return rc;
  if (p->num_tokens == 0) return MLAPO_OK;
  const void* req[] = {p->hidden_state, p->wdqkv_km, p->descale0, p->gamma1, p->beta1, p->wuq_km, p->descale1, p->gamma2,
                       p->cos, p->sin, p->wuk, p->slot_mapping, p->q_out0, p->q_out1, p->kv_cache, p->kv_cache_rope};
  for (size_t i = 0; i < sizeof(req) / sizeof(req[0]); ++i)
    if (!req[i]) return fail(MLAPO_ERR_INVALID_ARG, "required pointer #%zu is null", i);
  if (p->quant_mode == MLAPO_QUANT_PER_TENSOR_ASYMM &&
      (!p->quant_scale0 || !p->quant_offset0 || !p->quant_scale1 || !p->quant_offset1 || !p->bias0 || !p->bias1))
    return fail(MLAPO_ERR_INVALID_ARG, "per_tensor_quant_asymm needs quant_scale/offset 0/1 and bias0/1");
  if (p->cache_mode == MLAPO_CACHE_INT8_NZ && (!p->ctkv_scale || !p->q_nope_scale))
    return fail(MLAPO_ERR_INVALID_ARG, "int8_nzcache needs ctkv_scale and q_nope_scale");
  MlapoWorkspaceLayout l;
  compute_layout(p, &l);
  if (!workspace || (reinterpret_cast<uintptr_t>(workspace) & 255) != 0)
    return fail(MLAPO_ERR_WORKSPACE, "workspace must be non-null and 256-byte aligned");
  if (workspace_bytes < l.total_bytes)
    return fail(MLAPO_ERR_WORKSPACE, "workspace too small: %zu < %zu", workspace_bytes, l.total_bytes);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  if (p->act_dtype == MLAPO_ACT_BF16) return launch_impl<__nv_bfloat16>(p, ws, l, st);
  return launch_impl<__half>(p, ws, l, st);
}

}  // extern "C"
