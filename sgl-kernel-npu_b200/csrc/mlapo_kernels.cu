// mla_preprocess for B200 (sm_100a): hand-written CUDA kernels + the C-ABI launcher.
//
// Path (reference phases P1..P6, csrc/mla_preprocess/op_kernel/mla_preprocess_mix_bf16.hpp:2773-2916):
//   k1 quant_hidden        hidden[N,D] act -> a0 int8                       (Quant, bf16.hpp:283-534)
//   k2 gemm_i8<partial>    a0 . Wdqkv^T -> split-K int32 partial sums       (PpMatmulW8a8Aic, bf16.hpp:1515-1975)
//   k3 norm_rope_cache     sum partials, +bias, *descale -> act y0;
//                          q_a RMSNorm(+beta)+quant -> a1 int8              (RmsNormQuant, bf16.hpp:536-840)
//                          latent RMSNorm(512), RoPE(64), paged-cache scatter (RmsNormAndRopeConvergence1, :2499-2686)
//   k4 gemm_i8<q>          a1 . Wuq^T (per head), dequant, RoPE -> q_out1; nope slice -> ws (PpMatmulW8a8Aiv + RopeFp16)
//   k5 bmm_absorb          q_nope[:,h,:] . Wuk[h] -> q_out0 (+ int8 re-quant) (PpMatmulEinSum :1074-1513, EinSumQuant :842-1065)
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
template <class T>
__global__ void quant_hidden_kernel(const T* __restrict__ x, int8_t* __restrict__ q, const T* __restrict__ scale,
                                    const int8_t* __restrict__ offset, long long n_vec8) {
  const float s = Act<T>::to_f(scale[0]);
  const float zp = (float)offset[0];
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_vec8; i += (long long)gridDim.x * blockDim.x) {
    const uint4 v = ptx::ld_global_nc_v4(reinterpret_cast<const uint4*>(x) + i);
    float f[8];
    unpack8<T>(v, f);
    int r[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) r[j] = sat_i8(__fadd_rn(__fdiv_rn(f[j], s), zp));
    ptx::st_global_v2(q + i * 8, pack4_i8(r[0], r[1], r[2], r[3]), pack4_i8(r[4], r[5], r[6], r[7]));
  }
}

// per-token symmetric (UNPINNED restatement of bf16.hpp:437-475): one block per row
template <class T>
__global__ void __launch_bounds__(256) quant_hidden_per_token_kernel(const T* __restrict__ x, int8_t* __restrict__ q,
                                                                      float* __restrict__ tok_scale, int d) {
  __shared__ float s_red[8];
  const int row = blockIdx.x;
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
#pragma unroll
  for (int w = 1; w < 8; ++w) amax = fmaxf(amax, s_red[w]);
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
}

// ============================================================================================ k2/k4: int8 GEMM
// C[128 x BLOCK_N] (int32, TMEM) = A[128 tokens x K] (int8, K-major) . B[BLOCK_N rows x K]^T (int8, K-major).
// 6 warps: w0 = TMA producer, w1 = tcgen05.mma issuer (+TMEM alloc), w2..w5 = epilogue (TMEM lane quarter = warp%4).
constexpr int kGemmThreads = 192;
constexpr int kBlockM = 128;
constexpr int kBlockKBytes = 128;  // one 128B swizzle atom along K
constexpr int kUmmaKBytes = 32;    // K per tcgen05.mma for 8-bit (32 elems) and 16-bit (16 elems) operands

enum { kEpiPartial = 0, kEpiQ = 1 };

struct GemmEpiArgs {
  // kEpiPartial: int32 partial sums [splits, M, n_total]
  int32_t* partial;
  int n_total;
  // kEpiQ: dequant + RoPE + split
  const void* descale;   // [192*H]
  const int32_t* bias;   // [192*H] or null (per-token)
  const float* tok_scale;  // [M] or null
  const void* cos;       // act [M,64]
  const void* sin;
  void* q_nope;          // act [M, H, 128]
  void* q_out1;          // act [M, H, 64]
  int num_heads;
  int descale_is_i64;
};

template <int BLOCK_N>
constexpr int gemm_smem_bytes(int stages) {
  return stages * (kBlockM * kBlockKBytes + BLOCK_N * kBlockKBytes) + (2 * stages + 2) * 8 + 16 + 2 * BLOCK_N * 4 + 1024;
}

template <class T, int BLOCK_N, int STAGES, int EPI>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_i8_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, GemmEpiArgs ea,
               int m_total, int kb_total) {
  constexpr int A_BYTES = kBlockM * kBlockKBytes;
  constexpr int B_BYTES = BLOCK_N * kBlockKBytes;
  constexpr int TMEM_COLS = BLOCK_N <= 32 ? 32 : BLOCK_N <= 64 ? 64 : BLOCK_N <= 128 ? 128 : 256;
  static_assert(BLOCK_N % 16 == 0 && BLOCK_N <= 256, "UMMA N");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* s_a = smem;
  uint8_t* s_b = smem + STAGES * A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * (A_BYTES + B_BYTES));
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + STAGES;
  uint64_t* tmem_full_bar = bars + 2 * STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 2);
  float* s_descale = reinterpret_cast<float*>(tmem_slot + 4);
  int32_t* s_bias = reinterpret_cast<int32_t*>(s_descale + BLOCK_N);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n_tile = blockIdx.x;
  const int split = blockIdx.y;
  const int splits = gridDim.y;
  const int m0 = blockIdx.z * kBlockM;
  const int n0 = n_tile * BLOCK_N;
  const int kb0 = (int)(((long long)split * kb_total) / splits);
  const int kb1 = (int)(((long long)(split + 1) * kb_total) / splits);

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_a);
    ptx::prefetch_tmap(&tm_b);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < STAGES; ++i) {
        ptx::mbar_init(ptx::smem_u32(full_bar + i), 1);
        ptx::mbar_init(ptx::smem_u32(empty_bar + i), 1);
      }
      ptx::mbar_init(ptx::smem_u32(tmem_full_bar), 1);
      ptx::fence_mbar_init();
    }
    __syncwarp();
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(ptx::smem_u32(empty_bar + stage), phase ^ 1);
        const uint32_t fb = ptx::smem_u32(full_bar + stage);
        ptx::mbar_arrive_expect_tx(fb, A_BYTES + B_BYTES);
        ptx::tma_load_2d(ptx::smem_u32(s_a + stage * A_BYTES), &tm_a, fb, kb * kBlockKBytes, m0, ptx::kEvictLast);
        ptx::tma_load_2d(ptx::smem_u32(s_b + stage * B_BYTES), &tm_b, fb, kb * kBlockKBytes, n0, ptx::kEvictFirst);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one thread)
    constexpr uint32_t idesc = ptx::make_idesc(ptx::kCFmtS32, ptx::kFmtS8, ptx::kFmtS8, 0, 0, kBlockM, BLOCK_N);
    uint32_t stage = 0, phase = 0;
    for (int kb = kb0; kb < kb1; ++kb) {
      ptx::mbar_wait(ptx::smem_u32(full_bar + stage), phase);
      ptx::tc_fence_after();
      if (lane == 0) {
        const uint32_t a_base = ptx::smem_u32(s_a + stage * A_BYTES);
        const uint32_t b_base = ptx::smem_u32(s_b + stage * B_BYTES);
#pragma unroll
        for (int k = 0; k < kBlockKBytes / kUmmaKBytes; ++k) {
          const uint64_t da = ptx::make_smem_desc(a_base + k * kUmmaKBytes, 1024, 0);
          const uint64_t db = ptx::make_smem_desc(b_base + k * kUmmaKBytes, 1024, 0);
          ptx::umma_i8(tmem_acc, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
        }
        ptx::umma_commit(ptx::smem_u32(empty_bar + stage));  // frees the smem slot when these MMAs retire
      }
      __syncwarp();
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
    if (lane == 0) ptx::umma_commit(ptx::smem_u32(tmem_full_bar));
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ epilogue (4 warps, 128 threads)
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const int t = m0 + row;
    const uint32_t t_base = tmem_acc + ((uint32_t)(quarter * 32) << 16);
    const int et = threadIdx.x - 64;  // 0..127

    if (EPI == kEpiQ) {
      for (int c = et; c < BLOCK_N; c += 128) {
        s_descale[c] = load_descale(ea.descale, ea.descale_is_i64, n0 + c);
        s_bias[c] = ea.bias ? ea.bias[n0 + c] : 0;
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");  // epilogue warps only
    }

    ptx::mbar_wait(ptx::smem_u32(tmem_full_bar), 0);
    ptx::tc_fence_after();

    if (EPI == kEpiPartial) {
      int32_t* out = ea.partial + ((long long)split * m_total + t) * ea.n_total + n0;
#pragma unroll 1
      for (int c0 = 0; c0 < BLOCK_N; c0 += 32) {
        uint32_t r[32];
        ptx::tmem_ld_32x32b_x32(t_base + c0, r);
        ptx::tmem_ld_wait();
        if (t < m_total) {
#pragma unroll
          for (int j = 0; j < 32; j += 4)
            if (n0 + c0 + j < ea.n_total) ptx::st_global_v4(out + c0 + j, r[j], r[j + 1], r[j + 2], r[j + 3]);
        }
      }
    } else {
      // BLOCK_N == 192 = one head: [nope 128 | rope 64]
      const int head = n_tile;
      const float ts = ea.tok_scale ? ea.tok_scale[t < m_total ? t : 0] : 1.f;
      const bool per_token = ea.tok_scale != nullptr;
      auto deq = [&](uint32_t acc, int c) -> float {
        float y = per_token ? __fmul_rn(__fmul_rn(__int2float_rn((int)acc), s_descale[c]), ts)
                            : __fmul_rn(__int2float_rn((int)acc + s_bias[c]), s_descale[c]);
        return y;
      };
      T* qn = reinterpret_cast<T*>(ea.q_nope) + ((long long)t * ea.num_heads + head) * 128;
#pragma unroll 1
      for (int c0 = 0; c0 < 128; c0 += 32) {
        uint32_t r[32];
        ptx::tmem_ld_32x32b_x32(t_base + c0, r);
        ptx::tmem_ld_wait();
        if (t < m_total) {
#pragma unroll
          for (int j = 0; j < 32; j += 8) {
            uint32_t w[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) w[u] = Act<T>::pack(deq(r[j + 2 * u], c0 + j + 2 * u), deq(r[j + 2 * u + 1], c0 + j + 2 * u + 1));
            ptx::st_global_v4(qn + c0 + j, w[0], w[1], w[2], w[3]);
          }
        }
      }
      // rope slice: columns 128..191
      float x[64];
      {
        uint32_t r[32];
        ptx::tmem_ld_32x32b_x32(t_base + 128, r);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) x[j] = Act<T>::round(deq(r[j], 128 + j));
        ptx::tmem_ld_32x32b_x32(t_base + 160, r);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) x[32 + j] = Act<T>::round(deq(r[j], 160 + j));
      }
      if (t < m_total) {
        const uint4* cs = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(ea.cos) + (long long)t * 64);
        const uint4* sn = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(ea.sin) + (long long)t * 64);
        T* o = reinterpret_cast<T*>(ea.q_out1) + ((long long)t * ea.num_heads + head) * 64;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          float c_lo[8], s_lo[8], c_hi[8], s_hi[8];
          unpack8<T>(cs[g], c_lo);
          unpack8<T>(sn[g], s_lo);
          unpack8<T>(cs[g + 4], c_hi);
          unpack8<T>(sn[g + 4], s_hi);
          uint32_t wl[4], wh[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int i0 = g * 8 + 2 * u, i1 = i0 + 1;
            // out[i] = x[i]*cos[i] + (-x[i+32])*sin[i];  out[i+32] = x[i+32]*cos[i+32] + x[i]*sin[i+32]
            const float a0 = __fadd_rn(__fmul_rn(x[i0], c_lo[2 * u]), __fmul_rn(-x[i0 + 32], s_lo[2 * u]));
            const float a1 = __fadd_rn(__fmul_rn(x[i1], c_lo[2 * u + 1]), __fmul_rn(-x[i1 + 32], s_lo[2 * u + 1]));
            const float b0 = __fadd_rn(__fmul_rn(x[i0 + 32], c_hi[2 * u]), __fmul_rn(x[i0], s_hi[2 * u]));
            const float b1 = __fadd_rn(__fmul_rn(x[i1 + 32], c_hi[2 * u + 1]), __fmul_rn(x[i1], s_hi[2 * u + 1]));
            wl[u] = Act<T>::pack(a0, a1);
            wh[u] = Act<T>::pack(b0, b1);
          }
          ptx::st_global_v4(o + g * 8, wl[0], wl[1], wl[2], wl[3]);
          ptx::st_global_v4(o + 32 + g * 8, wh[0], wh[1], wh[2], wh[3]);
        }
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_acc, TMEM_COLS);
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
};

constexpr int kNormThreads = 288;  // warps 0-1: k_nope (512), warp 2 lanes 0-7: k_pe (64), warps 3-8: q_a (1536)

__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <class T>
__global__ void __launch_bounds__(kNormThreads) norm_rope_cache_kernel(NormArgs a) {
  __shared__ double s_sum[9];
  __shared__ float s_max[9];
  const int t = blockIdx.x;
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const bool is_nope = warp < 2;
  const bool is_pe = warp == 2;
  const bool active = is_nope || (is_pe && lane < 8) || warp >= 3;
  int col = 0;  // first of this thread's 8 columns in [0, 2112)
  if (is_nope) col = tid * 8;
  else if (is_pe) col = 512 + (lane & 7) * 8;
  else col = 576 + (tid - 96) * 8;

  // ---- sum split-K partials (+bias), dequantise, round to act  (S2 of the contract)
  float x[8];
  {
    int acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (active) {
      for (int s = 0; s < a.splits; ++s) {
        const int4* p = reinterpret_cast<const int4*>(a.partial + ((long long)s * a.n_tokens + t) * MLAPO_MM1_OUT + col);
        const int4 u = p[0], v = p[1];
        acc[0] += u.x; acc[1] += u.y; acc[2] += u.z; acc[3] += u.w;
        acc[4] += v.x; acc[5] += v.y; acc[6] += v.z; acc[7] += v.w;
      }
    }
    const float ts0 = a.tok_scale0 ? a.tok_scale0[t] : 1.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float d = load_descale(a.descale0, a.descale_is_i64, col + j);
      float y;
      if (a.tok_scale0) y = __fmul_rn(__fmul_rn(__int2float_rn(acc[j]), d), ts0);
      else y = __fmul_rn(__int2float_rn(acc[j] + a.bias0[col + j]), d);
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
  if (lane == 0) s_sum[warp] = ss;
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
    unpack8<T>(reinterpret_cast<const uint4*>(a.gamma2)[tid], g);
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
      unpack8<T>(reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(a.cos) + (long long)t * 64)[lane], cs);
      unpack8<T>(reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(a.sin) + (long long)t * 64)[lane], sn);
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
  if (warp >= 3) {
    const float var = __fdiv_rn(__double2float_rn(tot_q), 1536.f);
    const float rs = __fdiv_rn(1.f, __fsqrt_rn(__fadd_rn(var, 1e-6f)));
    float g[8], b[8];
    unpack8<T>(reinterpret_cast<const uint4*>(a.gamma1)[tid - 96], g);
    unpack8<T>(reinterpret_cast<const uint4*>(a.beta1)[tid - 96], b);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      yq[j] = __fadd_rn(__fmul_rn(__fmul_rn(x[j], rs), g[j]), b[j]);
      amax = fmaxf(amax, fabsf(yq[j]));
    }
  }
  if (a.tok_scale1) {  // per-token symmetric (UNPINNED; bf16.hpp:739-777)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if (lane == 0) s_max[warp] = amax;
    __syncthreads();
    float m = s_max[3];
#pragma unroll
    for (int w = 4; w < 9; ++w) m = fmaxf(m, s_max[w]);
    const float s1 = __fdiv_rn(m, 127.f);
    const float inv = __fdiv_rn(1.f, s1);
    if (tid == 96) a.tok_scale1[t] = s1;
    if (warp >= 3) {
      int q[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) q[j] = sat_i8(__fmul_rn(yq[j], inv));
      ptx::st_global_v2(a.a1 + (long long)t * MLAPO_Q_LORA + (col - 576), pack4_i8(q[0], q[1], q[2], q[3]), pack4_i8(q[4], q[5], q[6], q[7]));
    }
  } else if (warp >= 3) {
    const float s = Act<T>::to_f(reinterpret_cast<const T*>(a.quant_scale1)[0]);
    const float zp = (float)a.quant_offset1[0];
    int q[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) q[j] = sat_i8(__fadd_rn(__fdiv_rn(yq[j], s), zp));
    ptx::st_global_v2(a.a1 + (long long)t * MLAPO_Q_LORA + (col - 576), pack4_i8(q[0], q[1], q[2], q[3]), pack4_i8(q[4], q[5], q[6], q[7]));
  }
}

// ============================================================================================ k5: W_UK absorption
// q_out0[t, h, 0:512] = act( q_nope[t, h, 0:128] . wuk[h][128, 512] ), fp32 accumulate in TMEM.
// A = q_nope tile [128 tokens x 128] K-major (2 k-blocks of 64 elems = 128 B rows, SW128);
// B = wuk[h] is [K=128][N=512] with N contiguous = MN-major: 8 atoms of [128 k-rows x 64 n] (SW128), LBO = 16 KB, SBO = 1 KB.
struct BmmArgs {
  void* q_out0;              // act or int8 [M, H, 512]
  const void* q_nope_scale;  // act [H] (int8 out) or null
  int num_heads;
  int m_total;
};
constexpr int kBmmSmemBytes = 2 * 16384 + 8 * 16384 + 8 * 8 + 16 + 1024;

template <class T, bool kInt8Out>
__global__ void __launch_bounds__(kGemmThreads, 1)
bmm_absorb_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b, BmmArgs ba) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* s_a = smem;               // 2 x 16 KB
  uint8_t* s_b = smem + 2 * 16384;   // 8 x 16 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 10 * 16384);
  uint64_t* full_bar = bars;         // [2]: {A + B half 0}, {B half 1}
  uint64_t* tmem_full_bar = bars + 2;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int head = blockIdx.x;
  const int m0 = blockIdx.y * kBlockM;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tm_a);
    ptx::prefetch_tmap(&tm_b);
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < 4; ++i) ptx::mbar_init(ptx::smem_u32(bars + i), 1);
      ptx::fence_mbar_init();
    }
    __syncwarp();
    ptx::tmem_alloc(ptx::smem_u32(tmem_slot), 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_acc = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t fb0 = ptx::smem_u32(full_bar), fb1 = ptx::smem_u32(full_bar + 1);
      ptx::mbar_arrive_expect_tx(fb0, 2 * 16384 + 4 * 16384);
      for (int kb = 0; kb < 2; ++kb)
        ptx::tma_load_2d(ptx::smem_u32(s_a + kb * 16384), &tm_a, fb0, head * 128 + kb * 64, m0, ptx::kEvictFirst);
      for (int at = 0; at < 4; ++at)
        ptx::tma_load_2d(ptx::smem_u32(s_b + at * 16384), &tm_b, fb0, at * 64, head * 128, ptx::kEvictFirst);
      ptx::mbar_arrive_expect_tx(fb1, 4 * 16384);
      for (int at = 4; at < 8; ++at)
        ptx::tma_load_2d(ptx::smem_u32(s_b + at * 16384), &tm_b, fb1, at * 64, head * 128, ptx::kEvictFirst);
    }
  } else if (warp == 1) {
    constexpr uint32_t kAFmt = sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value ? ptx::kFmtBF16 : ptx::kFmtF16;
    constexpr uint32_t idesc = ptx::make_idesc(ptx::kCFmtF32, kAFmt, kAFmt, 0, 1, kBlockM, 256);
    for (int half = 0; half < 2; ++half) {
      ptx::mbar_wait(ptx::smem_u32(full_bar + half), 0);
      ptx::tc_fence_after();
      if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {  // 8 x UMMA_K(16) = 128
          const uint32_t a_addr = ptx::smem_u32(s_a + (k >> 2) * 16384) + (k & 3) * kUmmaKBytes;
          const uint32_t b_addr = ptx::smem_u32(s_b + half * 4 * 16384) + k * 2048;  // 16 k-rows x 128 B
          const uint64_t da = ptx::make_smem_desc(a_addr, 1024, 0);
          const uint64_t db = ptx::make_smem_desc(b_addr, 1024, 16384);
          ptx::umma_f16(tmem_acc + half * 256, da, db, idesc, k > 0 ? 1u : 0u);
        }
        ptx::umma_commit(ptx::smem_u32(tmem_full_bar + half));
      }
      __syncwarp();
    }
  } else {
    const int quarter = warp & 3;
    const int t = m0 + quarter * 32 + lane;
    const uint32_t t_base = tmem_acc + ((uint32_t)(quarter * 32) << 16);
    float qs = 1.f;
    if (kInt8Out) qs = Act<T>::to_f(reinterpret_cast<const T*>(ba.q_nope_scale)[head]);
    for (int half = 0; half < 2; ++half) {
      ptx::mbar_wait(ptx::smem_u32(tmem_full_bar + half), 0);
      ptx::tc_fence_after();
#pragma unroll 1
      for (int c0 = 0; c0 < 256; c0 += 32) {
        uint32_t r[32];
        ptx::tmem_ld_32x32b_x32(t_base + half * 256 + c0, r);
        ptx::tmem_ld_wait();
        if (t < ba.m_total) {
          const long long o = ((long long)t * ba.num_heads + head) * 512 + half * 256 + c0;
          if (kInt8Out) {
            // S9: int8( fp16( act(acc) * q_nope_scale[h] ) )   (golden test_mla:86-92, 466-470)
            uint32_t w[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              int q[4];
#pragma unroll
              for (int u = 0; u < 4; ++u) q[u] = sat_i8(__fmul_rn(Act<T>::round(__uint_as_float(r[4 * j + u])), qs));
              w[j] = pack4_i8(q[0], q[1], q[2], q[3]);
            }
            int8_t* dst = reinterpret_cast<int8_t*>(ba.q_out0) + o;
            ptx::st_global_v4(dst, w[0], w[1], w[2], w[3]);
            ptx::st_global_v4(dst + 16, w[4], w[5], w[6], w[7]);
          } else {
            T* dst = reinterpret_cast<T*>(ba.q_out0) + o;
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              ptx::st_global_v4(dst + j, Act<T>::pack(__uint_as_float(r[j]), __uint_as_float(r[j + 1])),
                                Act<T>::pack(__uint_as_float(r[j + 2]), __uint_as_float(r[j + 3])),
                                Act<T>::pack(__uint_as_float(r[j + 4]), __uint_as_float(r[j + 5])),
                                Act<T>::pack(__uint_as_float(r[j + 6]), __uint_as_float(r[j + 7])));
            }
          }
        }
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_acc, 512);
}

// ============================================================================================ host side
thread_local char g_err[512] = "";
static int g_debug_flags = 0;  // bit0: sync + check after each kernel; bit1: write the y0 tap; bit2: time each kernel
constexpr int kMaxKernels = 8;
static cudaEvent_t g_ev[kMaxKernels + 1];
static bool g_ev_ready = false;
static int g_ev_count = 0;  // events recorded by the last launch
static const char* g_kernel_names[kMaxKernels] = {"quant_hidden", "gemm1_i8_splitk", "norm_rope_cache", "gemm2_i8_rope",
                                                  "bmm_absorb", "", "", ""};
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

// 2-D row-major tensor [rows, cols] of `elem_bytes` elements, box [box_rows, box_cols], 128B swizzle.
static int make_tmap_2d(CUtensorMap* m, CUtensorMapDataType dt, int elem_bytes, const void* base, uint64_t rows, uint64_t cols,
                        uint64_t row_stride_bytes, uint32_t box_rows, uint32_t box_cols) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return fail(MLAPO_ERR_DRIVER, "cuTensorMapEncodeTiled entry point not available");
  if (box_cols * elem_bytes != 128) return fail(MLAPO_ERR_INVALID_ARG, "tensor map inner box must be 128 bytes");
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstr[1] = {row_stride_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, dt, 2, const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
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

constexpr int kMm1BlockN = 64;
constexpr int kMm1Stages = 4;
constexpr int kMm2Stages = 4;

static int choose_splits0(int n_tokens, int hidden_dim) {
  const int m_tiles = n_tokens > 0 ? (n_tokens + kBlockM - 1) / kBlockM : 1;
  const int n_tiles = (MLAPO_MM1_OUT + kMm1BlockN - 1) / kMm1BlockN;
  const int kb_total = hidden_dim / kBlockKBytes;
  int s = (2 * 148 + n_tiles * m_tiles / 2) / (n_tiles * m_tiles);
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
  const size_t n = (size_t)p->num_tokens, d = (size_t)p->hidden_dim, h = (size_t)p->num_heads;
  const int splits = choose_splits0(p->num_tokens, p->hidden_dim);
  size_t off = 0;
  l->a0 = off;        off = align_up(off + n * d, 256);
  l->partial0 = off;  off = align_up(off + (size_t)splits * n * MLAPO_MM1_OUT * 4, 256);
  l->a1 = off;        off = align_up(off + n * MLAPO_Q_LORA, 256);
  l->q_nope = off;    off = align_up(off + n * h * MLAPO_Q_NOPE * 2, 256);
  l->tok_scale = off; off = align_up(off + 2 * n * 4, 256);
  l->total_bytes = off + (g_debug_flags & 2 ? align_up(n * MLAPO_MM1_OUT * 2, 256) : 0);
  l->splits0 = splits;
  l->reserved = 0;
  return MLAPO_OK;
}

#define MLAPO_CHECK_LAUNCH(name)                                                                    \
  do {                                                                                              \
    cudaError_t e__ = cudaGetLastError();                                                           \
    if (e__ == cudaSuccess && (g_debug_flags & 1)) e__ = cudaStreamSynchronize(st);                 \
    if (e__ != cudaSuccess) return fail(MLAPO_ERR_CUDA, "%s: %s", name, cudaGetErrorString(e__));   \
  } while (0)

template <class T>
static int launch_impl(const MlapoParams* p, uint8_t* ws, const MlapoWorkspaceLayout& l, cudaStream_t st) {
  const int n = p->num_tokens, d = p->hidden_dim, h = p->num_heads;
  const bool per_token = p->quant_mode == MLAPO_QUANT_PER_TOKEN_SYMM;
  const int m_tiles = (n + kBlockM - 1) / kBlockM;
  int8_t* a0 = reinterpret_cast<int8_t*>(ws + l.a0);
  int32_t* partial0 = reinterpret_cast<int32_t*>(ws + l.partial0);
  int8_t* a1 = reinterpret_cast<int8_t*>(ws + l.a1);
  T* q_nope = reinterpret_cast<T*>(ws + l.q_nope);
  float* tok_scale = reinterpret_cast<float*>(ws + l.tok_scale);
  const CUtensorMapDataType act_dt = std::is_same<T, __nv_bfloat16>::value ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  int rc;
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;

  // ---- k1
  record_event(0, st);
  if (per_token) {
    quant_hidden_per_token_kernel<T><<<n, 256, 0, st>>>(reinterpret_cast<const T*>(p->hidden_state), a0, tok_scale, d);
  } else {
    const long long nvec = (long long)n * d / 8;
    int blocks = (int)((nvec + 255) / 256);
    if (blocks > num_sms() * 8) blocks = num_sms() * 8;
    quant_hidden_kernel<T><<<blocks, 256, 0, st>>>(reinterpret_cast<const T*>(p->hidden_state), a0,
                                                    reinterpret_cast<const T*>(p->quant_scale0),
                                                    reinterpret_cast<const int8_t*>(p->quant_offset0), nvec);
  }
  MLAPO_CHECK_LAUNCH("quant_hidden");
  record_event(1, st);

  // ---- k2: MM1, split-K partial sums
  {
    CUtensorMap tm_a, tm_b;
    if ((rc = make_tmap_2d(&tm_a, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a0, n, d, d, kBlockM, 128))) return rc;
    if ((rc = make_tmap_2d(&tm_b, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, p->wdqkv_km, MLAPO_MM1_OUT, d, d, kMm1BlockN, 128))) return rc;
    GemmEpiArgs ea{};
    ea.partial = partial0;
    ea.n_total = MLAPO_MM1_OUT;
    auto kern = gemm_i8_kernel<T, kMm1BlockN, kMm1Stages, kEpiPartial>;
    constexpr int smem = gemm_smem_bytes<kMm1BlockN>(kMm1Stages);
    static bool attr_set[64] = {};
    if (!attr_set[dev]) {
      if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess)
        return fail(MLAPO_ERR_CUDA, "cudaFuncSetAttribute(gemm1): %s", cudaGetErrorString(cudaGetLastError()));
      attr_set[dev] = true;
    }
    dim3 grid((MLAPO_MM1_OUT + kMm1BlockN - 1) / kMm1BlockN, l.splits0, m_tiles);
    kern<<<grid, kGemmThreads, smem, st>>>(tm_a, tm_b, ea, n, d / kBlockKBytes);
    MLAPO_CHECK_LAUNCH("gemm1_i8");
  }
  record_event(2, st);

  // ---- k3
  {
    NormArgs a{};
    a.partial = partial0; a.splits = l.splits0; a.n_tokens = n;
    a.bias0 = per_token ? nullptr : reinterpret_cast<const int32_t*>(p->bias0);
    a.descale0 = p->descale0; a.descale_is_i64 = p->descale_is_int64;
    a.tok_scale0 = per_token ? tok_scale : nullptr;
    a.tok_scale1 = per_token ? tok_scale + n : nullptr;
    a.gamma1 = p->gamma1; a.beta1 = p->beta1; a.gamma2 = p->gamma2;
    a.quant_scale1 = p->quant_scale1; a.quant_offset1 = reinterpret_cast<const int8_t*>(p->quant_offset1);
    a.cos = p->cos; a.sin = p->sin;
    a.slots = reinterpret_cast<const int32_t*>(p->slot_mapping);
    a.ctkv_scale = p->ctkv_scale;
    a.a1 = a1;
    a.y0_tap = (g_debug_flags & 2) ? ws + align_up(l.tok_scale + 2 * (size_t)n * 4, 256) : nullptr;
    a.kv_cache = p->kv_cache; a.kv_cache_rope = p->kv_cache_rope;
    a.cache_mode = p->cache_mode;
    a.cache_slots = (long long)p->num_cache_blocks * p->cache_block_size;
    norm_rope_cache_kernel<T><<<n, kNormThreads, 0, st>>>(a);
    MLAPO_CHECK_LAUNCH("norm_rope_cache");
  }
  record_event(3, st);

  // ---- k4: MM2 per head + dequant + RoPE
  {
    CUtensorMap tm_a, tm_b;
    if ((rc = make_tmap_2d(&tm_a, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a1, n, MLAPO_Q_LORA, MLAPO_Q_LORA, kBlockM, 128))) return rc;
    if ((rc = make_tmap_2d(&tm_b, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, p->wuq_km, (uint64_t)h * MLAPO_Q_HEAD, MLAPO_Q_LORA, MLAPO_Q_LORA, MLAPO_Q_HEAD, 128))) return rc;
    GemmEpiArgs ea{};
    ea.descale = p->descale1; ea.descale_is_i64 = p->descale_is_int64;
    ea.bias = per_token ? nullptr : reinterpret_cast<const int32_t*>(p->bias1);
    ea.tok_scale = per_token ? tok_scale + n : nullptr;
    ea.cos = p->cos; ea.sin = p->sin;
    ea.q_nope = q_nope; ea.q_out1 = p->q_out1; ea.num_heads = h;
    auto kern = gemm_i8_kernel<T, MLAPO_Q_HEAD, kMm2Stages, kEpiQ>;
    constexpr int smem = gemm_smem_bytes<MLAPO_Q_HEAD>(kMm2Stages);
    static bool attr_set[64] = {};
    if (!attr_set[dev]) {
      if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess)
        return fail(MLAPO_ERR_CUDA, "cudaFuncSetAttribute(gemm2): %s", cudaGetErrorString(cudaGetLastError()));
      attr_set[dev] = true;
    }
    dim3 grid(h, 1, m_tiles);
    kern<<<grid, kGemmThreads, smem, st>>>(tm_a, tm_b, ea, n, MLAPO_Q_LORA / kBlockKBytes);
    MLAPO_CHECK_LAUNCH("gemm2_i8_rope");
  }
  record_event(4, st);

  // ---- k5: absorb W_UK
  {
    CUtensorMap tm_a, tm_b;
    if ((rc = make_tmap_2d(&tm_a, act_dt, 2, q_nope, n, (uint64_t)h * 128, (uint64_t)h * 256, kBlockM, 64))) return rc;
    if ((rc = make_tmap_2d(&tm_b, act_dt, 2, p->wuk, (uint64_t)h * 128, 512, 1024, 128, 64))) return rc;
    BmmArgs ba{};
    ba.q_out0 = p->q_out0; ba.q_nope_scale = p->q_nope_scale; ba.num_heads = h; ba.m_total = n;
    dim3 grid(h, m_tiles);
    if (p->cache_mode == MLAPO_CACHE_INT8_NZ) {
      auto kern = bmm_absorb_kernel<T, true>;
      static bool attr_set[64] = {};
      if (!attr_set[dev]) { cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kBmmSmemBytes); attr_set[dev] = true; }
      kern<<<grid, kGemmThreads, kBmmSmemBytes, st>>>(tm_a, tm_b, ba);
    } else {
      auto kern = bmm_absorb_kernel<T, false>;
      static bool attr_set[64] = {};
      if (!attr_set[dev]) { cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kBmmSmemBytes); attr_set[dev] = true; }
      kern<<<grid, kGemmThreads, kBmmSmemBytes, st>>>(tm_a, tm_b, ba);
    }
    MLAPO_CHECK_LAUNCH("bmm_absorb");
  }
  record_event(5, st);
  return MLAPO_OK;
}

}  // namespace mlapo

// ============================================================================================ C ABI
using namespace mlapo;

extern "C" {

const char* mlapo_version(void) { return "mla_preprocess_b200 0.1.0 (sm_100a; tcgen05+TMA; reference sgl-kernel-npu f425bdf1)"; }
const char* mlapo_last_error(void) { return g_err; }

int mlapo_set_debug_flags(int flags) {
  const int old = g_debug_flags;
  g_debug_flags = flags;
  return old;
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
  (void)p;
  return 5;
}

int mlapo_repack_nz_weight(const void* w_nz, void* w_km, int32_t k, int32_t n_out, void* stream) {
  if (!w_nz || !w_km) return fail(MLAPO_ERR_INVALID_ARG, "repack: null pointer");
  if (k <= 0 || k % 32 != 0 || n_out <= 0) return fail(MLAPO_ERR_INVALID_ARG, "repack: k must be a multiple of 32");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long chunks = (long long)n_out * (k / 32) * 2;
  int blocks = (int)((chunks + 255) / 256);
  if (blocks > 148 * 16) blocks = 148 * 16;
  repack_nz_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const int8_t*>(w_nz), reinterpret_cast<int8_t*>(w_km), k, n_out);
  MLAPO_CHECK_LAUNCH("repack_nz");
  return MLAPO_OK;
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
