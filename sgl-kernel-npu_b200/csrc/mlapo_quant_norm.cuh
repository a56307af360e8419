// k0 weight repack, k1 quantise hiddenState, k3 norms / quant / RoPE / paged-cache scatter (memory-bound kernels: 128-bit accesses, warp-shuffle reductions).
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

// The same on PAIRS of adjacent vectors (prefill sizes): 32 bytes in and 16 bytes out per thread and step — half the memory
// instructions of quant_vectors for the same bytes.  n_vec8 must be even (hidden_dim is a multiple of 128).
template <class T, int U>
__device__ __forceinline__ void quant_vector_pairs(const T* __restrict__ x, int8_t* __restrict__ q, const QuantScale& qs, long long n_vec8,
                                                   long long first, long long stride) {
  const long long n_pair = n_vec8 >> 1;
  for (long long i0 = first; i0 < n_pair; i0 += stride * U) {
    uint4 v[U][2];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i0 + u * stride < n_pair) ptx::ld_global_nc_v8(reinterpret_cast<const uint4*>(x) + 2 * (i0 + u * stride), v[u][0], v[u][1]);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i >= n_pair) break;
      uint32_t w[4];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float f[8];
        unpack8<T>(v[u][h], f);
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
        w[2 * h] = pack4_i8(r[0], r[1], r[2], r[3]);
        w[2 * h + 1] = pack4_i8(r[4], r[5], r[6], r[7]);
      }
      ptx::st_global_v4(q + i * 16, w[0], w[1], w[2], w[3]);
    }
  }
}

// The call's three weight tensors (70 MB at 128 heads) fit the 126 MB L2.  The first kernel of the call asks for all of them
// before it waits for its predecessor, so HBM streams weights while quant / MM1 / norm — which move little data — run,
// and the weight loads of MM1 and of the q path find L2 lines (or fills already in flight) instead of starting at DRAM.
struct WeightPrefetch {
  const char* ptr[3];      // W_DQKV, W_UQ, W_UK in the order the call consumes them
  long long lines[3];      // 128-byte lines each
};
__device__ __forceinline__ void prefetch_weights_l2(const WeightPrefetch& wp) {
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (long long)gridDim.x * blockDim.x;
#pragma unroll
  for (int r = 0; r < 3; ++r)
    for (long long i = tid; i < wp.lines[r]; i += nth) asm volatile("prefetch.global.L2 [%0];" ::"l"(wp.ptr[r] + i * 128));
}

template <class T, int U>
__global__ void quant_hidden_kernel(const T* __restrict__ x, int8_t* __restrict__ q, const T* __restrict__ scale,
                                    const int8_t* __restrict__ offset, long long n_vec8, WeightPrefetch wp, long long* tl,
                                    int input_complete) {
  if (threadIdx.x == 0 && blockIdx.x < 4096) tl_mark_global(tl, 0);
  ptx::griddep_launch_dependents();
  prefetch_weights_l2(wp);
  // Default: every CTA waits for the previous kernel on the stream to COMPLETE before it reads hiddenState or writes a0.  That
  // is what makes the call safe behind any producer, including one that lets its dependents start before its last store
  // (programmatic dependent launch).
  //
  // MLAPO_FLAG_INPUT_COMPLETE (the caller's statement that nothing this call touches is still pending when the previous kernel
  // releases its dependents; see the header): no wait before the stores, this kernel runs under the tail of whatever precedes it.
  // When that is the previous call's q kernel, a0's last reader (that call's MM1) is complete by the time this grid can start,
  // because every q kernel signals its dependents only AFTER its own griddepcontrol.wait (norm complete => MM1 complete, the
  // chain being transitive); a kernel that does not use programmatic launch signals at its completion.  A wait at the END keeps
  // the chain transitive: this grid is not complete, and so MM1 does not pass its wait and write the split-K slabs, before the
  // predecessor is.  ONE CTA waiting is enough for that (the last one, which is dispatched last); CTAs that all waited would sit
  // on the few idle SMs and keep the rest of the grid from starting.
  if (!input_complete) ptx::griddep_wait();
  QuantScale qs;
  qs.load(scale, offset);
  if (threadIdx.x == 0 && blockIdx.x < 4096) tl_mark_global(tl, 14);
  if (U == 8) quant_vector_pairs<T, 4>(x, q, qs, n_vec8, blockIdx.x * (long long)blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x);
  else quant_vectors<T, U>(x, q, qs, n_vec8, blockIdx.x * (long long)blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x);
  if (threadIdx.x == 0 && blockIdx.x < 4096) tl_mark_global(tl, 15);
  if (input_complete && blockIdx.x == gridDim.x - 1) ptx::griddep_wait();
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
                                                                      float* __restrict__ tok_scale, int d, WeightPrefetch wp) {
  ptx::griddep_launch_dependents();
  prefetch_weights_l2(wp);
  ptx::griddep_wait();
  quant_row_per_token<T>(x, q, tok_scale, d, blockIdx.x);
}

// ============================================================================================ k3: norms, quant, rope, cache
struct NormArgs {
  const int32_t* partial;  // [splits, N, 2112]
  const void* y0_in;       // act [N, 2112]: the dequantised rows when MM1's epilogue produced them (kMaxSplits == 0), else unused
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
  int input_complete;       // MLAPO_FLAG_INPUT_COMPLETE: the per-call inputs (cos, sin, slots) may be read before the upstream wait
};

constexpr int kNormThreads = 288;  // warps 0-1: k_nope (512), warp 2 lanes 0-7: k_pe (64), warps 3-8: q_a (1536)

__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One token row, executed by every thread of the CTA (threads beyond the first 288 only take part in the barriers).
// kPdlWait: the kernel form; the weight-like parameters (descale, bias, gamma / beta), which are complete long before the call, are
// loaded BEFORE griddepcontrol.wait, i.e. while MM1 is still running, so that only the partial sums are waited for.  The per-call
// inputs (cos / sin, the slot) join them only under MLAPO_FLAG_INPUT_COMPLETE: this kernel can start as early as the call's first
// kernel (each kernel releases its dependents at once), so without the caller's word they might still be in the making.
template <class T, int kMaxSplits, bool kPdlWait = false>
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
  const bool early_inputs = !kPdlWait || a.input_complete;
  int slot = early_inputs ? a.slots[t] : 0;
  {
    float d[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    int bias[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if constexpr (kMaxSplits == 0) {
      // MM1's epilogue already dequantised: nothing to fetch for that
    } else if (a.descale_is_i64) {
#pragma unroll
      for (int j = 0; j < 8; ++j) d[j] = load_descale(a.descale0, 1, col + j);
    } else {
      const float4 d0 = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(a.descale0) + col);
      const float4 d1 = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(a.descale0) + col + 4);
      d[0] = d0.x; d[1] = d0.y; d[2] = d0.z; d[3] = d0.w; d[4] = d1.x; d[5] = d1.y; d[6] = d1.z; d[7] = d1.w;
    }
    if (kMaxSplits != 0 && a.bias0) {
      const int4 b0 = *reinterpret_cast<const int4*>(a.bias0 + col), b1 = *reinterpret_cast<const int4*>(a.bias0 + col + 4);
      bias[0] = b0.x; bias[1] = b0.y; bias[2] = b0.z; bias[3] = b0.w; bias[4] = b1.x; bias[5] = b1.y; bias[6] = b1.z; bias[7] = b1.w;
    }
    if (is_nope) {
      g_vec = reinterpret_cast<const uint4*>(a.gamma2)[tid];
    } else if (is_pe && early_inputs) {
      g_vec = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(a.cos) + (long long)t * 64)[lane & 7];
      b_vec = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(a.sin) + (long long)t * 64)[lane & 7];
    } else if (is_q) {
      g_vec = reinterpret_cast<const uint4*>(a.gamma1)[tid - 96];
      b_vec = reinterpret_cast<const uint4*>(a.beta1)[tid - 96];
    }
    if constexpr (kPdlWait) {
      ptx::griddep_wait();  // partial sums (and the per-token scales of a0) come from the kernels before
      if (tid == 0 && blockIdx.x < 4096) tl_mark_global(a.tl, 14);
      if (!early_inputs) {
        slot = a.slots[t];
        if (is_pe) {
          g_vec = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(a.cos) + (long long)t * 64)[lane & 7];
          b_vec = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(a.sin) + (long long)t * 64)[lane & 7];
        }
      }
    }
    if constexpr (kMaxSplits == 0) {
      uint4 v = make_uint4(0, 0, 0, 0);
      if (active) v = *reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(a.y0_in) + (long long)t * MLAPO_MM1_OUT + col);
      unpack8<T>(v, x);
    } else {
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
__global__ void __launch_bounds__(kNormThreads, kMaxSplits <= 1 ? 4 : 2) norm_rope_cache_kernel(NormArgs a) {
  const int tid = threadIdx.x;
  if (tid == 0 && blockIdx.x < 4096) tl_mark_global(a.tl, 0);
  ptx::griddep_launch_dependents();
  norm_token<T, kMaxSplits, true>(a, blockIdx.x, tid);  // waits for the MM1 kernel inside, after its own parameter loads
  if (tid == 0 && blockIdx.x < 4096) tl_mark_global(a.tl, 15);
}


// ---- prefill form of k3: one WARP per token row, behind the dequantising MM1 (the row is read as act y0 [N, 2112]).
// The CTA-per-token kernel above keeps four rows in flight per SM, which is right for decode (every SM has a row at once) and
// latency-bound for thousands of rows; here a warp owns a row end to end (no shared memory, no CTA barrier), 16+ rows are in
// flight per SM and the norm parameters are fetched once per warp.  Same arithmetic, operation by operation, as norm_token().
// Lane l holds k_nope vectors l, l+32 (8 columns each), k_pe vector l (l < 8) and q_a vectors l + 32 j, j < 6.
constexpr int kNormWarpThreads = 256;
template <class T>
__global__ void __launch_bounds__(kNormWarpThreads, 2) norm_rows_warp_kernel(NormArgs a) {
  const int lane = threadIdx.x & 31;
  const int gw = blockIdx.x * (kNormWarpThreads / 32) + (threadIdx.x >> 5), warps_total = gridDim.x * (kNormWarpThreads / 32);
  if (threadIdx.x == 0 && blockIdx.x < 4096) tl_mark_global(a.tl, 0);
  ptx::griddep_launch_dependents();
  uint4 g2[2];   // gamma2 stays in registers; gamma1 / beta1 (12 vectors per lane) are re-read per row from L1 (volatile asm loads,
                 // so that the compiler does not hoist them into 48 registers and spill)
#pragma unroll
  for (int j = 0; j < 2; ++j) g2[j] = reinterpret_cast<const uint4*>(a.gamma2)[lane + 32 * j];
  const uint4* const g1p = reinterpret_cast<const uint4*>(a.gamma1) + lane;
  const uint4* const b1p = reinterpret_cast<const uint4*>(a.beta1) + lane;
  const bool per_token = a.tok_scale1 != nullptr;
  const float qs1 = per_token ? 1.f : Act<T>::to_f(reinterpret_cast<const T*>(a.quant_scale1)[0]);
  const float zp1 = per_token ? 0.f : (float)a.quant_offset1[0];
  const float cs_kv = a.cache_mode == MLAPO_CACHE_INT8_NZ ? Act<T>::to_f(reinterpret_cast<const T*>(a.ctkv_scale)[0]) : 1.f;
  ptx::griddep_wait();  // y0 (and, per token, nothing else) comes from the MM1 kernel
  if (threadIdx.x == 0 && blockIdx.x < 4096) tl_mark_global(a.tl, 14);
  for (int t = gw; t < a.n_tokens; t += warps_total) {
    const T* row = reinterpret_cast<const T*>(a.y0_in) + (long long)t * MLAPO_MM1_OUT;
    uint4 xn[2], xq[6], xpe = make_uint4(0, 0, 0, 0), cs_v = xpe, sn_v = xpe;
#pragma unroll
    for (int j = 0; j < 2; ++j) xn[j] = *reinterpret_cast<const uint4*>(row + 8 * (lane + 32 * j));
#pragma unroll
    for (int j = 0; j < 6; ++j) xq[j] = *reinterpret_cast<const uint4*>(row + 576 + 8 * (lane + 32 * j));
    if (lane < 8) {
      xpe = *reinterpret_cast<const uint4*>(row + 512 + 8 * lane);
      cs_v = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(a.cos) + (long long)t * 64)[lane];
      sn_v = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(a.sin) + (long long)t * 64)[lane];
    }
    const int slot = a.slots[t];
    // ---- row sums of squares, accumulated in fp64 (order-free), rounded once to fp32
    double ssn = 0.0, ssq = 0.0;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      float f[8];
      unpack8<T>(xn[j], f);
#pragma unroll
      for (int e = 0; e < 8; ++e) ssn += (double)f[e] * (double)f[e];
    }
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      float f[8];
      unpack8<T>(xq[j], f);
#pragma unroll
      for (int e = 0; e < 8; ++e) ssq += (double)f[e] * (double)f[e];
    }
    ssn = warp_sum_f64(ssn);
    ssq = warp_sum_f64(ssq);
    const bool slot_ok = slot >= 0 && (long long)slot < a.cache_slots;
    const int blk = slot >> 7, r128 = slot & 127;  // NZ modes: block size 128
    // ---- S8: k_nope = act(x * rsqrt(mean+eps) * gamma2)
    {
      const float var = __fdiv_rn(__double2float_rn(ssn), 512.f);
      const float rs = __fdiv_rn(1.f, __fsqrt_rn(__fadd_rn(var, 1e-6f)));
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int col = 8 * (lane + 32 * j);
        float x[8], g[8], y[8];
        unpack8<T>(xn[j], x);
        unpack8<T>(g2[j], g);
#pragma unroll
        for (int e = 0; e < 8; ++e) y[e] = __fmul_rn(__fmul_rn(x[e], rs), g[e]);
        if (slot_ok) {
          if (a.cache_mode == MLAPO_CACHE_INT8_NZ) {
            int q[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) q[e] = sat_i8(__fdiv_rn(y[e], cs_kv));
            int8_t* dst = reinterpret_cast<int8_t*>(a.kv_cache) + (long long)blk * (128 * 512) + (col >> 5) * (128 * 32) + r128 * 32 + (col & 31);
            ptx::st_global_v2(dst, pack4_i8(q[0], q[1], q[2], q[3]), pack4_i8(q[4], q[5], q[6], q[7]));
          } else {
            T* base = reinterpret_cast<T*>(a.kv_cache);
            T* dst = a.cache_mode == MLAPO_CACHE_NZ ? base + (long long)blk * (128 * 512) + (col >> 4) * (128 * 16) + r128 * 16 + (col & 15)
                                                    : base + (long long)slot * 512 + col;
            ptx::st_global_v4(dst, Act<T>::pack(y[0], y[1]), Act<T>::pack(y[2], y[3]), Act<T>::pack(y[4], y[5]), Act<T>::pack(y[6], y[7]));
          }
        }
      }
    }
    // ---- S7: k_pe = rope(x[512:576]); lanes l and l^4 (l < 8) hold partner halves
    {
      float x[8], xp[8];
      unpack8<T>(xpe, x);
#pragma unroll
      for (int e = 0; e < 8; ++e) xp[e] = __shfl_xor_sync(0xffffffffu, x[e], 4);
      if (lane < 8 && slot_ok) {
        const int c = lane * 8;
        float cs[8], sn[8], y[8];
        unpack8<T>(cs_v, cs);
        unpack8<T>(sn_v, sn);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float rot = lane < 4 ? -xp[e] : xp[e];
          y[e] = __fadd_rn(__fmul_rn(x[e], cs[e]), __fmul_rn(rot, sn[e]));
        }
        T* base = reinterpret_cast<T*>(a.kv_cache_rope);
        T* dst = a.cache_mode == MLAPO_CACHE_KROPE_CTKV ? base + (long long)slot * 64 + c
                                                        : base + (long long)blk * (128 * 64) + (c >> 4) * (128 * 16) + r128 * 16 + (c & 15);
        ptx::st_global_v4(dst, Act<T>::pack(y[0], y[1]), Act<T>::pack(y[2], y[3]), Act<T>::pack(y[4], y[5]), Act<T>::pack(y[6], y[7]));
      }
    }
    // ---- S4: q_a RMSNorm(+beta) + quant -> a1
    {
      const float var = __fdiv_rn(__double2float_rn(ssq), 1536.f);
      const float rs = __fdiv_rn(1.f, __fsqrt_rn(__fadd_rn(var, 1e-6f)));
      float inv = 0.f;
      if (per_token) {  // UNPINNED mode: s1 = max|y| / 127 over the row
        float amax = 0.f;
#pragma unroll
        for (int j = 0; j < 6; ++j) {
          float x[8], g[8], b[8];
          unpack8<T>(xq[j], x);
          unpack8<T>(ptx::ld_global_nc_v4(g1p + 32 * j), g);
          unpack8<T>(ptx::ld_global_nc_v4(b1p + 32 * j), b);
#pragma unroll
          for (int e = 0; e < 8; ++e) amax = fmaxf(amax, fabsf(__fadd_rn(__fmul_rn(__fmul_rn(x[e], rs), g[e]), b[e])));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        const float s1 = __fdiv_rn(amax, 127.f);
        inv = __fdiv_rn(1.f, s1);
        if (lane == 0) a.tok_scale1[t] = s1;
      }
#pragma unroll
      for (int j = 0; j < 6; ++j) {
        float x[8], g[8], b[8];
        unpack8<T>(xq[j], x);
        unpack8<T>(ptx::ld_global_nc_v4(g1p + 32 * j), g);
        unpack8<T>(ptx::ld_global_nc_v4(b1p + 32 * j), b);
        int q[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float yq = __fadd_rn(__fmul_rn(__fmul_rn(x[e], rs), g[e]), b[e]);
          q[e] = per_token ? sat_i8(__fmul_rn(yq, inv)) : sat_i8(__fadd_rn(__fdiv_rn(yq, qs1), zp1));
        }
        ptx::st_global_v2(a.a1 + (long long)t * MLAPO_Q_LORA + 8 * (lane + 32 * j), pack4_i8(q[0], q[1], q[2], q[3]), pack4_i8(q[4], q[5], q[6], q[7]));
      }
    }
  }
  if (threadIdx.x == 0 && blockIdx.x < 4096) tl_mark_global(a.tl, 15);
}


}  // namespace mlapo
