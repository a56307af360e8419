// split_qkv_rmsnorm_rope: the non-MLA analogue of this prologue for GQA models — split a fused qkv row,
// RMSNorm every q and k head (weight, optional bias), rotate the first rope_dim dims, copy v.  Reference: the Triton kernel
// python/sgl_kernel_npu/sgl_kernel_npu/norm/split_qkv_rmsnorm_rope.py:9-372 (wrapper :374-439).  Contract restated from it:
// everything in fp32, ONE rounding at the store;  n = x * (1/sqrt(mean(x^2) + eps)) [* w [+ b]];  neox style:
// out[:h] = n[:h]*cos[:h] - n[h:r]*sin[:h], out[h:r] = n[h:r]*cos[h:r] + n[:h]*sin[h:r]  (h = rope_dim/2, r = rope_dim);
// interleaved style: pairs (2i, 2i+1) with sin[i], cos[i];  dims >= rope_dim pass through normalised.
// One warp per (token, head); the head vector sits in shared memory so that RoPE partners are a plain read.
#pragma once
#include "mlapo_numerics.cuh"

namespace mlapo {

struct SplitQkvArgs {
  const void* qkv;       // act [batch, q_hidden + 2 kv_hidden]
  const void* sin;       // act [batch, rope_dim]
  const void* cos;
  void* q_out;           // act [batch, q_hidden]
  void* k_out;           // act [batch, kv_hidden]
  void* v_out;
  const void* q_weight;  // act [head_dim] or null (no norm)
  const void* k_weight;
  const void* q_bias;    // act [head_dim] or null
  const void* k_bias;
  long long batch;
  int q_hidden, kv_hidden, head_dim, rope_dim;
  float eps;
  int has_norm, neox;
};

constexpr int kSplitWarps = 8;
constexpr int kSplitMaxHeadDim = 512;

template <class T>
__global__ void __launch_bounds__(kSplitWarps * 32) split_qkv_rmsnorm_rope_kernel(SplitQkvArgs a) {
  __shared__ float s_x[kSplitWarps][kSplitMaxHeadDim];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = a.q_hidden / a.head_dim, nk = a.kv_hidden / a.head_dim;
  const int heads = nq + 2 * nk;
  const long long total = a.batch * heads;
  const int hd = a.head_dim, rd = a.rope_dim, half = a.rope_dim / 2;
  float* x = s_x[warp];
  for (long long item = (long long)blockIdx.x * kSplitWarps + warp; item < total; item += (long long)gridDim.x * kSplitWarps) {
    const long long t = item / heads;
    const int h = (int)(item % heads);
    const T* src = reinterpret_cast<const T*>(a.qkv) + t * (a.q_hidden + 2LL * a.kv_hidden) + (long long)h * hd;
    T* dst;
    const T *w = nullptr, *b = nullptr;
    if (h < nq) { dst = reinterpret_cast<T*>(a.q_out) + t * a.q_hidden + (long long)h * hd; w = (const T*)a.q_weight; b = (const T*)a.q_bias; }
    else if (h < nq + nk) { dst = reinterpret_cast<T*>(a.k_out) + t * a.kv_hidden + (long long)(h - nq) * hd; w = (const T*)a.k_weight; b = (const T*)a.k_bias; }
    else {  // v: plain copy
      dst = reinterpret_cast<T*>(a.v_out) + t * a.kv_hidden + (long long)(h - nq - nk) * hd;
      for (int e = lane; e < hd; e += 32) dst[e] = src[e];
      continue;
    }
    float ss = 0.f;
    for (int e = lane; e < hd; e += 32) {
      const float v = Act<T>::to_f(src[e]);
      x[e] = v;
      ss = __fadd_rn(ss, __fmul_rn(v, v));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss = __fadd_rn(ss, __shfl_xor_sync(0xffffffffu, ss, o));
    if (a.has_norm) {
      const float rstd = __fdiv_rn(1.f, __fsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)hd), a.eps)));
      for (int e = lane; e < hd; e += 32) {
        float n = __fmul_rn(__fmul_rn(x[e], rstd), Act<T>::to_f(w[e]));
        if (b) n = __fadd_rn(n, Act<T>::to_f(b[e]));
        x[e] = n;
      }
    }
    __syncwarp();
    const T* sn = reinterpret_cast<const T*>(a.sin) + t * rd;
    const T* cs = reinterpret_cast<const T*>(a.cos) + t * rd;
    for (int e = lane; e < hd; e += 32) {
      float o = x[e];
      if (e < rd) {
        float rot, s, c;
        if (a.neox) {
          rot = e < half ? -x[e + half] : x[e - half];
          s = Act<T>::to_f(sn[e]);
          c = Act<T>::to_f(cs[e]);
        } else {
          rot = (e & 1) ? x[e - 1] : -x[e + 1];
          s = Act<T>::to_f(sn[e >> 1]);
          c = Act<T>::to_f(cs[e >> 1]);
        }
        o = __fadd_rn(__fmul_rn(rot, s), __fmul_rn(o, c));
      }
      dst[e] = Act<T>::from_f(o);
    }
    __syncwarp();
  }
}

}  // namespace mlapo
