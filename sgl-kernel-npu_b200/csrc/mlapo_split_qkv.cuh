// split_qkv_rmsnorm_rope: the non-MLA analogue of this prologue for GQA models — split a fused qkv row,
// RMSNorm every q and k head (weight, optional bias), rotate the first rope_dim dims, copy v.  Reference: the Triton kernel
// python/sgl_kernel_npu/sgl_kernel_npu/norm/split_qkv_rmsnorm_rope.py:9-372 (wrapper :374-439).  Contract restated from it:
// everything in fp32, ONE rounding at the store;  n = x * (1/sqrt(mean(x^2) + eps)) [* w [+ b]];  neox style:
// out[:h] = n[:h]*cos[:h] - n[h:r]*sin[:h], out[h:r] = n[h:r]*cos[h:r] + n[:h]*sin[h:r]  (h = rope_dim/2, r = rope_dim);
// interleaved style: pairs (2i, 2i+1) with sin[i], cos[i];  dims >= rope_dim pass through normalised.
//
// HBM-bound by its bytes (4 per element), but an SM has only ~22 issue slots and ~0.17 shared-memory wavefronts per element at
// 6.5 TB/s, so the main kernel is laid out to spend few of either.  A CTA takes 4096-element chunks of one token's fused row
// through a ring of stages filled by 1-D bulk copies; head_dim / V lanes make a head (V = 8, 16 for head_dim 512), a lane owns
// two slices of V/2 elements chosen so that a neox RoPE pair lives in ONE lane (no exchange), shared-memory reads are
// conflict-free 8/16-byte accesses, the fp32 arithmetic is issued as two-wide f32x2 instructions, results go straight to global.
// Shapes the vector layout does not fit (head_dim < 32, rope_dim not a multiple of V, unaligned pointers) take the scalar kernel
// at the end of the file.
#pragma once
#include "mlapo_numerics.cuh"
#include "sm100_ptx.cuh"

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
  // split_qkvgate_gemma_rmsnorm_rope (same reference module, :441-747): every q head is followed by a gate head of the same
  // width ([q_0 | gate_0 | q_1 | gate_1 ... | k | v]) that is copied to gate_out, and the norm multiplies by (1 + weight)
  void* gate_out;        // act [batch, q_hidden] or null
  int gated, gemma;
  // split_qkv_rmsnorm_rope_pos_cache_half_npu (reference norm/split_qkv_rmsnorm_rope_pos_cache_half_npu.py:25-230): cos / sin come
  // from row clamp(positions[t], 0, max_seq - 1) of a cache [max_seq, rope_dim] laid out [cos(rope_dim/2) | sin(rope_dim/2)] (one
  // value per neox pair), and the normalised value may be rounded to the tensor dtype before the rotation (cast_norm)
  const void* cache;      // act or f32 [max_seq, cache_stride], or null: sin / cos above are used
  const void* positions;  // int32 / int64 [batch]; null with a cache: token t uses row t
  long long cache_stride; // elements
  int pos_i64, cache_f32, max_seq, cast_norm;
};


constexpr int kSplitWarps = 8;
constexpr int kSplitMaxHeadDim = 512;

// two-wide fp32 arithmetic (FMUL2 / FFMA2 on sm_100): half the issue slots of the scalar forms
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
  unsigned long long d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;"
      : "=l"(d)
      : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)), "l"(*reinterpret_cast<unsigned long long*>(&c)));
  return *reinterpret_cast<float2*>(&d);
}

// V 16-bit elements as raw words; one load / store instruction up to 16 bytes, two for 32
template <int V>
struct Raw16 {
  uint32_t w[(V + 1) / 2];
  __device__ __forceinline__ void load(const void* p) {
    if constexpr (V == 2) w[0] = *reinterpret_cast<const uint32_t*>(p);
    else if constexpr (V == 4) { const uint2 t = *reinterpret_cast<const uint2*>(p); w[0] = t.x; w[1] = t.y; }
    else {
#pragma unroll
      for (int i = 0; i < V / 8; ++i) {
        const uint4 t = reinterpret_cast<const uint4*>(p)[i];
        w[4 * i] = t.x; w[4 * i + 1] = t.y; w[4 * i + 2] = t.z; w[4 * i + 3] = t.w;
      }
    }
  }
  __device__ __forceinline__ void store(void* p) const {
    if constexpr (V == 4) *reinterpret_cast<uint2*>(p) = make_uint2(w[0], w[1]);
#pragma unroll
    for (int i = 0; i < V / 8; ++i) reinterpret_cast<uint4*>(p)[i] = make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
  }
  __device__ __forceinline__ void zero() {
#pragma unroll
    for (int i = 0; i < (V + 1) / 2; ++i) w[i] = 0u;
  }
};

template <int HD>
struct SplitGeom {
  static constexpr int V = HD >= 512 ? 16 : 8;          // elements per lane
  static constexpr int L = HD / V;                      // lanes per head: 4 .. 32
  static constexpr int SPAN = 32 * V;                   // elements of the fused row per warp
  static constexpr int U = 16 / V;                      // spans per warp and chunk
  static constexpr int CHUNK = kSplitWarps * U * SPAN;  // elements of one token's row a CTA takes per stage (4096 = 8 KB)
  static constexpr int STAGES = 4;
  static constexpr int STAGE_BYTES = CHUNK * 2 + 4 * HD;  // chunk, then this token's sin and cos (<= HD elements each)
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 16 * HD;
};

// Work unit = (token, chunk of its fused row).  A ninth warp keeps the STAGES-deep ring full with 1-D bulk copies (row chunk, sin,
// cos) onto one mbarrier per stage; the 8 working warps pull their 16 bytes per lane into registers, release the stage at once
// (one arrive per warp) and store results straight to global, so no CTA-wide barrier sits in the loop.
// A CTA stays on ONE chunk index of the row and walks over tokens, so where a lane's heads lie and go is loop-invariant.  The
// grid is (chunks_per_tok - 1) * g_full + g_last CTAs: g_full per full chunk index, g_last for the last, possibly shorter, chunk
// (sized by the host in proportion to its length so that all CTAs finish together).
constexpr int kSplitThreads = (kSplitWarps + 1) * 32;

template <class T, int HD, bool NEOX>
__global__ void __launch_bounds__(kSplitThreads, 3) split_qkv_vec_kernel(SplitQkvArgs a, unsigned chunks_per_tok, unsigned g_full, unsigned g_last) {
  using namespace ptx;
  using G = SplitGeom<HD>;
  constexpr int V = G::V, L = G::L, SPAN = G::SPAN, U = G::U, P = V / 2, S = G::STAGES;
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ __align__(8) uint64_t s_full[S], s_empty[S];
  float* s_wb = reinterpret_cast<float*>(smem + S * G::STAGE_BYTES);  // [4][HD]: q weight, k weight, q bias, k bias
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qh = a.q_hidden, kvh = a.kv_hidden;
  const int qspan = a.gated ? 2 * qh : qh;  // columns of the fused row before k
  const int row = qspan + 2 * kvh;
  const int rd = a.rope_dim, half = rd / 2;

  const unsigned full_ctas = (chunks_per_tok - 1) * g_full;
  const bool last = blockIdx.x >= full_ctas;
  const unsigned ch = last ? chunks_per_tok - 1 : blockIdx.x / g_full;
  const unsigned tstep = last ? g_last : g_full;
  const unsigned batch = (unsigned)a.batch;
  unsigned t = last ? blockIdx.x - full_ctas : blockIdx.x - ch * g_full;
  if (threadIdx.x == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(smem_u32(&s_full[s]), 1); mbar_init(smem_u32(&s_empty[s]), kSplitWarps); }
    fence_mbar_init();
  }
  if (a.has_norm) {
    const bool has_bias = a.q_bias != nullptr;
    const float w1 = a.gemma ? 1.f : 0.f;
    for (int i = threadIdx.x; i < HD; i += kSplitThreads) {
      s_wb[i] = Act<T>::to_f(reinterpret_cast<const T*>(a.q_weight)[i]) + w1;
      s_wb[HD + i] = Act<T>::to_f(reinterpret_cast<const T*>(a.k_weight)[i]) + w1;
      s_wb[2 * HD + i] = has_bias ? Act<T>::to_f(reinterpret_cast<const T*>(a.q_bias)[i]) : 0.f;
      s_wb[3 * HD + i] = has_bias ? Act<T>::to_f(reinterpret_cast<const T*>(a.k_bias)[i]) : 0.f;
    }
  }
  __syncthreads();

  if (warp == kSplitWarps) {  // producer
    if (lane == 0) {
      for (unsigned it = 0; t < batch; ++it, t += tstep) {
        const int stage = (int)(it % S);
        if (it >= (unsigned)S) mbar_wait_light(smem_u32(&s_empty[stage]), ((it / S) - 1) & 1);
        const uint32_t bar = smem_u32(&s_full[stage]);
        const uint32_t dst = smem_u32(smem + stage * G::STAGE_BYTES);
        const int off = (int)ch * G::CHUNK;
        const uint32_t bytes = 2u * (uint32_t)(row - off < G::CHUNK ? row - off : G::CHUNK);
        const bool rope_here = off < qspan + kvh;  // a chunk of v heads only needs no sin / cos
        const uint32_t cs_bytes = !rope_here ? 0u : a.cache ? (a.cache_f32 ? 4u : 2u) * (uint32_t)rd : 4u * (uint32_t)rd;
        mbar_arrive_expect_tx(bar, bytes + cs_bytes);
        bulk_load_1d(dst, reinterpret_cast<const T*>(a.qkv) + (long long)t * row + off, bytes, bar);
        if (rope_here && a.cache) {  // one cache row [cos | sin]
          long long p = !a.positions ? (long long)t : a.pos_i64 ? reinterpret_cast<const long long*>(a.positions)[t] : (long long)reinterpret_cast<const int*>(a.positions)[t];
          p = p < 0 ? 0 : p >= a.max_seq ? a.max_seq - 1 : p;
          bulk_load_1d(dst + G::CHUNK * 2, reinterpret_cast<const uint8_t*>(a.cache) + p * a.cache_stride * (a.cache_f32 ? 4 : 2), cs_bytes, bar);
        } else if (rope_here) {
          bulk_load_1d(dst + G::CHUNK * 2, reinterpret_cast<const T*>(a.sin) + (long long)t * rd, 2u * (uint32_t)rd, bar);
          bulk_load_1d(dst + G::CHUNK * 2 + 2 * HD, reinterpret_cast<const T*>(a.cos) + (long long)t * rd, 2u * (uint32_t)rd, bar);
        }
      }
    }
    return;
  }

  // The lane's V elements of a head are two slices of W: for neox rope lanes slice A sits in the first half of the rotated dims
  // and slice B holds its partners in the second half, so the rotation needs no exchange between lanes; every other lane owns
  // V contiguous elements (behind rope_dim for neox; pairs stay inside a lane for the interleaved style).
  constexpr int W = V / 2, PW = W / 2;
  const int j = lane % L;
  int oA, oB;
  bool in_rope;
  if constexpr (NEOX) {
    in_rope = j * W < half;
    oA = in_rope ? j * W : rd + V * (j - half / W);
    oB = in_rope ? half + j * W : oA + W;
  } else {
    in_rope = j * V < rd;
    oA = j * V;
    oB = oA + W;
  }
  const int head_pos = (lane / L) * HD;  // of the lane's head inside the warp's span
  // Where the lane's head of span u lies (kind: 0 q, 1 k, 2 v or gate: copied, 3 beyond the row) and where it goes: out0 is the output position
  // for token 0, tstr the output's token stride.
  int kind[U], tstr[U];
  T* out0[U];
  const int off = (int)ch * G::CHUNK;
  {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int c = off + (u * kSplitWarps + warp) * SPAN + head_pos;
      tstr[u] = kvh;
      if (c >= row) { kind[u] = 3; out0[u] = nullptr; }
      else if (c >= qspan + kvh) { kind[u] = 2; out0[u] = reinterpret_cast<T*>(a.v_out) + (c - qspan - kvh); }
      else if (c >= qspan) { kind[u] = 1; out0[u] = reinterpret_cast<T*>(a.k_out) + (c - qspan); }
      else {
        tstr[u] = qh;
        const int hh = c / HD;  // gated rows: even heads are q, odd heads their gates
        const bool gate = a.gated && (hh & 1);
        kind[u] = gate ? 2 : 0;
        out0[u] = (gate ? reinterpret_cast<T*>(a.gate_out) : reinterpret_cast<T*>(a.q_out)) + (a.gated ? (hh >> 1) * HD : c);
      }
    }
  }
  const uint32_t full0 = smem_u32(&s_full[0]), empty0 = smem_u32(&s_empty[0]);
  const int ldA = (warp * SPAN + head_pos + oA) * 2, ldB = (warp * SPAN + head_pos + oB) * 2;  // in a stage, round 0

  for (unsigned it = 0; t < batch; ++it, t += tstep) {
    const int stage = (int)(it % S);
    mbar_wait_light(full0 + 8 * stage, (it / S) & 1);
    const uint8_t* st = smem + stage * G::STAGE_BYTES;
    Raw16<W> rawA[U], rawB[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (kind[u] < 3) { rawA[u].load(st + ldA + u * kSplitWarps * SPAN * 2); rawB[u].load(st + ldB + u * kSplitWarps * SPAN * 2); }
      else { rawA[u].zero(); rawB[u].zero(); }
    }
    // this token's sin / cos for the lane's slices; the sign of the rotation is folded into sin
    float2 snA[PW], csA[PW], snB[PW], csB[PW];
    if (in_rope && off < qspan + kvh) {
      const uint8_t* ssin = st + G::CHUNK * 2;
      const uint8_t* scos = ssin + 2 * HD;
      if (NEOX && a.cache) {  // cache row in the stage: cos then sin, one value per pair -> both slices use the same
        if (a.cache_f32) {
          const float* cf = reinterpret_cast<const float*>(ssin);
#pragma unroll
          for (int p = 0; p < PW; ++p) {
            const float2 c = *reinterpret_cast<const float2*>(cf + oA + 2 * p), sv = *reinterpret_cast<const float2*>(cf + half + oA + 2 * p);
            csA[p] = csB[p] = c;
            snA[p] = make_float2(-sv.x, -sv.y);
            snB[p] = sv;
          }
        } else {
          Raw16<W> rc, rs;
          rc.load(ssin + oA * 2);
          rs.load(ssin + (half + oA) * 2);
#pragma unroll
          for (int p = 0; p < PW; ++p) {
            csA[p] = csB[p] = make_float2(Act<T>::lo(rc.w[p]), Act<T>::hi(rc.w[p]));
            snB[p] = make_float2(Act<T>::lo(rs.w[p]), Act<T>::hi(rs.w[p]));
            snA[p] = make_float2(-snB[p].x, -snB[p].y);
          }
        }
      } else if constexpr (NEOX) {
        Raw16<W> r;
        r.load(ssin + oA * 2);
#pragma unroll
        for (int p = 0; p < PW; ++p) snA[p] = make_float2(-Act<T>::lo(r.w[p]), -Act<T>::hi(r.w[p]));
        r.load(ssin + oB * 2);
#pragma unroll
        for (int p = 0; p < PW; ++p) snB[p] = make_float2(Act<T>::lo(r.w[p]), Act<T>::hi(r.w[p]));
        r.load(scos + oA * 2);
#pragma unroll
        for (int p = 0; p < PW; ++p) csA[p] = make_float2(Act<T>::lo(r.w[p]), Act<T>::hi(r.w[p]));
        r.load(scos + oB * 2);
#pragma unroll
        for (int p = 0; p < PW; ++p) csB[p] = make_float2(Act<T>::lo(r.w[p]), Act<T>::hi(r.w[p]));
      } else {
        Raw16<W> rs, rc;  // pair i of the head uses sin[i], cos[i]: W values for the lane's W pairs
        rs.load(ssin + oA);
        rc.load(scos + oA);
#pragma unroll
        for (int p = 0; p < W; ++p) {
          const float s = (p & 1) ? Act<T>::hi(rs.w[p >> 1]) : Act<T>::lo(rs.w[p >> 1]);
          const float c = (p & 1) ? Act<T>::hi(rc.w[p >> 1]) : Act<T>::lo(rc.w[p >> 1]);
          (p < PW ? snA[p] : snB[p - PW]) = make_float2(-s, s);
          (p < PW ? csA[p] : csB[p - PW]) = make_float2(c, c);
        }
      }
    }
    __syncwarp();  // every lane holds its part of the stage: give it back
    if (lane == 0) mbar_arrive(empty0 + 8 * stage);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float2 xA[PW], xB[PW];
      float2 acc = make_float2(0.f, 0.f);
#pragma unroll
      for (int p = 0; p < PW; ++p) {
        xA[p] = make_float2(Act<T>::lo(rawA[u].w[p]), Act<T>::hi(rawA[u].w[p]));
        xB[p] = make_float2(Act<T>::lo(rawB[u].w[p]), Act<T>::hi(rawB[u].w[p]));
        acc = fma2(xA[p], xA[p], acc);
        acc = fma2(xB[p], xB[p], acc);
      }
      if (a.has_norm) {
        float ss = acc.x + acc.y;
#pragma unroll
        for (int o = L / 2; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        if (kind[u] < 2) {
          // 1/sqrt by the hardware approximation and one Newton step (error below one fp32 ulp): five instructions where the
          // IEEE sqrt and division take ~35, a third of this loop's issue slots
          const float m = __fadd_rn(ss * (1.f / HD), a.eps);
          const float y0 = rsqrtf(m);
          const float rstd = fmaf(y0, fmaf(-(0.5f * m) * y0, y0, 0.5f), y0);
          const float2 r2 = make_float2(rstd, rstd);
          const float* wk = s_wb + kind[u] * HD;
          const float* bk = s_wb + (2 + kind[u]) * HD;
#pragma unroll
          for (int p = 0; p < PW; p += 2) {
            float4 w4 = *reinterpret_cast<const float4*>(wk + oA + 2 * p), b4 = *reinterpret_cast<const float4*>(bk + oA + 2 * p);
            xA[p] = fma2(mul2(xA[p], r2), make_float2(w4.x, w4.y), make_float2(b4.x, b4.y));
            xA[p + 1] = fma2(mul2(xA[p + 1], r2), make_float2(w4.z, w4.w), make_float2(b4.z, b4.w));
            w4 = *reinterpret_cast<const float4*>(wk + oB + 2 * p);
            b4 = *reinterpret_cast<const float4*>(bk + oB + 2 * p);
            xB[p] = fma2(mul2(xB[p], r2), make_float2(w4.x, w4.y), make_float2(b4.x, b4.y));
            xB[p + 1] = fma2(mul2(xB[p + 1], r2), make_float2(w4.z, w4.w), make_float2(b4.z, b4.w));
          }
          if (a.cast_norm) {  // the rotation then starts from the normalised value as the tensor dtype holds it
#pragma unroll
            for (int p = 0; p < PW; ++p) {
              const uint32_t wa = Act<T>::pack(xA[p].x, xA[p].y), wb = Act<T>::pack(xB[p].x, xB[p].y);
              xA[p] = make_float2(Act<T>::lo(wa), Act<T>::hi(wa));
              xB[p] = make_float2(Act<T>::lo(wb), Act<T>::hi(wb));
            }
          }
        }
      }
      if (kind[u] == 3) continue;
      T* dst = out0[u] + (long long)t * tstr[u];
      if (kind[u] == 2) {  // v, gate: the bits pass through
        rawA[u].store(dst + oA);
        rawB[u].store(dst + oB);
        continue;
      }
      if (in_rope) {
#pragma unroll
        for (int p = 0; p < PW; ++p) {
          if constexpr (NEOX) {  // A: n*cos - partner*sin;  B: n*cos + partner*sin
            const float2 na = fma2(xB[p], snA[p], mul2(xA[p], csA[p]));
            xB[p] = fma2(xA[p], snB[p], mul2(xB[p], csB[p]));
            xA[p] = na;
          } else {  // (x0, x1) -> (x0 c - x1 s, x1 c + x0 s)
            xA[p] = fma2(make_float2(xA[p].y, xA[p].x), snA[p], mul2(xA[p], csA[p]));
            xB[p] = fma2(make_float2(xB[p].y, xB[p].x), snB[p], mul2(xB[p], csB[p]));
          }
        }
      }
      Raw16<W> outA, outB;
#pragma unroll
      for (int p = 0; p < PW; ++p) {
        outA.w[p] = Act<T>::pack(xA[p].x, xA[p].y);
        outB.w[p] = Act<T>::pack(xB[p].x, xB[p].y);
      }
      outA.store(dst + oA);
      outB.store(dst + oB);
    }
  }
}

// Scalar kernel for the shapes the vector layout does not fit: one warp per (token, head), element by element.
template <class T>
__global__ void __launch_bounds__(kSplitWarps * 32) split_qkv_scalar_kernel(SplitQkvArgs a) {
  __shared__ float s_x[kSplitWarps][kSplitMaxHeadDim];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = a.q_hidden / a.head_dim, nk = a.kv_hidden / a.head_dim;
  const int nqs = a.gated ? 2 * nq : nq;  // heads of the fused row before k (q and gate heads alternate when gated)
  const int heads = nqs + 2 * nk;
  const long long total = a.batch * heads;
  const int hd = a.head_dim, rd = a.rope_dim, half = a.rope_dim / 2;
  const float w1 = a.gemma ? 1.f : 0.f;
  float* x = s_x[warp];
  for (long long item = (long long)blockIdx.x * kSplitWarps + warp; item < total; item += (long long)gridDim.x * kSplitWarps) {
    const long long t = item / heads;
    const int h = (int)(item % heads);
    const T* src = reinterpret_cast<const T*>(a.qkv) + t * heads * hd + (long long)h * hd;
    T* dst;
    const T *w = nullptr, *b = nullptr;
    bool copy = false;
    if (h < nqs) {
      const int hq = a.gated ? h >> 1 : h;
      copy = a.gated && (h & 1);
      dst = reinterpret_cast<T*>(copy ? a.gate_out : a.q_out) + t * a.q_hidden + (long long)hq * hd;
      w = (const T*)a.q_weight; b = (const T*)a.q_bias;
    } else if (h < nqs + nk) {
      dst = reinterpret_cast<T*>(a.k_out) + t * a.kv_hidden + (long long)(h - nqs) * hd; w = (const T*)a.k_weight; b = (const T*)a.k_bias;
    } else {
      copy = true;
      dst = reinterpret_cast<T*>(a.v_out) + t * a.kv_hidden + (long long)(h - nqs - nk) * hd;
    }
    if (copy) {  // v, gate
      for (int e = lane; e < hd; e += 32) dst[e] = src[e];
      continue;
    }
    float ss = 0.f;
    for (int e = lane; e < hd; e += 32) {
      const float v = Act<T>::to_f(src[e]);
      x[e] = v;
      ss = fmaf(v, v, ss);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if (a.has_norm) {
      const float rstd = __fdiv_rn(1.f, __fsqrt_rn(__fadd_rn(__fdiv_rn(ss, (float)hd), a.eps)));
      for (int e = lane; e < hd; e += 32) {
        const float n = __fmul_rn(x[e], rstd);
        const float we = Act<T>::to_f(w[e]) + w1;
        const float y = b ? fmaf(n, we, Act<T>::to_f(b[e])) : __fmul_rn(n, we);
        x[e] = a.cast_norm ? Act<T>::round(y) : y;
      }
    }
    __syncwarp();
    const T* sn = reinterpret_cast<const T*>(a.sin) + t * rd;
    const T* cs = reinterpret_cast<const T*>(a.cos) + t * rd;
    long long crow = 0;  // position-cache mode: element offset of the token's cache row
    if (a.cache) {
      long long p = !a.positions ? t : a.pos_i64 ? reinterpret_cast<const long long*>(a.positions)[t] : (long long)reinterpret_cast<const int*>(a.positions)[t];
      crow = (p < 0 ? 0 : p >= a.max_seq ? a.max_seq - 1 : p) * a.cache_stride;
    }
    for (int e = lane; e < hd; e += 32) {
      float o = x[e];
      if (e < rd) {
        float rot, s, c;
        if (a.cache) {  // neox, one cache value per pair
          rot = e < half ? -x[e + half] : x[e - half];
          const long long ci = crow + (e < half ? e : e - half);
          c = a.cache_f32 ? reinterpret_cast<const float*>(a.cache)[ci] : Act<T>::to_f(reinterpret_cast<const T*>(a.cache)[ci]);
          s = a.cache_f32 ? reinterpret_cast<const float*>(a.cache)[ci + half] : Act<T>::to_f(reinterpret_cast<const T*>(a.cache)[ci + half]);
        } else if (a.neox) {
          rot = e < half ? -x[e + half] : x[e - half];
          s = Act<T>::to_f(sn[e]);
          c = Act<T>::to_f(cs[e]);
        } else {
          rot = (e & 1) ? x[e - 1] : -x[e + 1];
          s = Act<T>::to_f(sn[e >> 1]);
          c = Act<T>::to_f(cs[e >> 1]);
        }
        o = fmaf(rot, s, __fmul_rn(o, c));
      }
      dst[e] = Act<T>::from_f(o);
    }
    __syncwarp();
  }
}

// split_qkv_rmsnorm_mrope (reference norm/split_qkv_rmsnorm_mrope.py:57-333, selection as the test's golden,
// tests/python/sgl_kernel_npu/test_split_qkv_rmsnorm_mrope.py:8-50): cos_sin is [3 (t, h, w), tokens, rope_dim] with rows
// [cos(rope_dim/2) | sin(rope_dim/2)]; pair i of a token takes its cos / sin from the t, h or w row by its section.  This kernel
// writes the selected row per token, [tokens, rope_dim] in the same layout, which the kernels above then read as a cache with
// row t for token t (2 * rope_dim bytes per token beside the 4 bytes per element of the main pass).
__device__ __forceinline__ int mrope_section_of(int i, int sec_t, int sec_h, int sec_w, int interleaved) {
  if (interleaved) return (i % 3 == 1 && i <= 3 * sec_h) ? 1 : (i % 3 == 2 && i <= 3 * sec_w) ? 2 : 0;
  return i < sec_t ? 0 : i < sec_t + sec_h ? 1 : 2;
}
template <class E>
__global__ void mrope_select_kernel(const E* __restrict__ cos_sin, E* __restrict__ out, long long tokens, int rope_dim, int sec_t, int sec_h,
                                    int sec_w, int interleaved) {
  const int half = rope_dim / 2;
  const long long total = tokens * rope_dim;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const long long t = idx / rope_dim;
    const int e = (int)(idx - t * rope_dim);
    const int sec = mrope_section_of(e < half ? e : e - half, sec_t, sec_h, sec_w, interleaved);
    out[idx] = cos_sin[((long long)sec * tokens + t) * rope_dim + e];
  }
}

}  // namespace mlapo
