// split_qkv_tp_rmsnorm_rope (the tensor-parallel member of the family).  Reference: the two Triton kernels and
// the all-reduce between them in python/sgl_kernel_npu/sgl_kernel_npu/norm/split_qkv_tp_rmsnorm_rope.py:8-170 (wrapper :172-288).
// Contract restated from it: the RMSNorm runs over the WHOLE q (and k) row of a token, whose columns are sharded over the
// tensor-parallel ranks:  var = (sum over ranks of  mean_local(x^2)) / tp_world;  y = act_rne(x * (1/sqrt(var + eps)) * w[col])
// with a weight per column; the first rotary_dim dims of every head of y are rotated neox style with cos[t, i], sin[t, i],
// i < rotary_dim/2 (one value per pair), rounded once more; v is copied.
//
// The reference makes three passes over q and k in global memory (split + local sums, normalise in place, rotate in place).
// Here a token's fused row is brought into shared memory ONCE by a bulk copy (ring of stages, producer warp), the sums are taken
// from there, and the normalised, rotated row goes straight to q / k / v: 4 bytes of HBM traffic per element.  Modes:
//   kTpFused  tp_world == 1: sums and apply in one pass;
//   kTpSums   writes the local means to qk_var [batch, 2] (the host all-reduces them: NCCL, 8 bytes per token);
//   kTpApply  the one-pass apply with the reduced qk_var;
//   kTpPeer   ONE kernel for tp_world > 1: each CTA publishes a token's local means straight into every rank's exchange buffer over
//             NVLink (plain peer stores of 8-byte {value, epoch} records: data and flag arrive together, no fence needed), works on
//             the next tokens' sums meanwhile, then applies with the peers' records, which a gather warp has meanwhile collected
//             from this GPU's OWN memory into shared memory.
//             The row never leaves shared memory in between, so the collective costs no extra HBM pass and no kernel boundary.
//             Every rank adds the records in rank order: all ranks get the same bits.  Needs all CTAs co-resident (the host sizes
//             the grid so) and the same batch on every rank; records of call e live in half (e & 1) of the buffer, and a rank can
//             only be one call ahead of a peer (its next call waits for the peer's records of this one), so halves never collide.
#pragma once
#include "mlapo_split_qkv.cuh"

namespace mlapo {

enum { kTpFused = 0, kTpSums = 1, kTpApply = 2, kTpPeer = 3 };
constexpr int kTpMaxRanks = 8;

struct SplitTpArgs {
  const void* qkv;       // act [batch, q_hidden + 2 kv_hidden]
  const void* cos;       // act [batch, rotary_dim] (first half used)
  const void* sin;
  void* q_out;           // act [batch, q_hidden]
  void* k_out;           // act [batch, kv_hidden]
  void* v_out;
  const void* q_weight;  // act [q_hidden]
  const void* k_weight;  // act [kv_hidden]
  float* qk_var;         // f32 [batch, 2]: written by kTpSums, read by kTpApply
  long long batch;
  int q_hidden, kv_hidden, head_dim, rotary_dim;
  float eps, inv_tp_world;
  int mode, stages;
  // kTpPeer: exchange buffer of every rank as mapped here, [2 halves][tp_world sources][capacity tokens][2 records of 8 bytes]
  unsigned long long peer[kTpMaxRanks];
  int tp_rank, tp_world, capacity;
  unsigned* epoch_state;  // device memory {epoch of the next call (never 0: the buffer starts zeroed), CTAs done}: read at kernel
                          // start, advanced by the last CTA to finish, so that replays of a captured graph get fresh epochs too
};

// warps of a CTA: apply warps, sum warps, one producer warp, one gather warp (peer mode)
constexpr int kTpApplyWarps = 5, kTpSumWarps = 2;
constexpr int kTpProducerWarp = kTpApplyWarps + kTpSumWarps, kTpGatherWarp = kTpProducerWarp + 1;
constexpr int kTpThreads = (kTpGatherWarp + 1) * 32;

// dynamic shared memory: stages x [row | cos row | sin row], then the norm weights [q_hidden | kv_hidden] in the tensor dtype
__host__ __device__ inline int tp_stage_bytes(int row, int rd) { return 2 * row + 4 * rd; }
__host__ __device__ inline int tp_smem_bytes(int row, int rd, int qh, int kvh, int stages) {
  return stages * tp_stage_bytes(row, rd) + 2 * (qh + kvh) + 16;
}

// A record is ONE 64-bit word {value (low), epoch (high)} written and read with single 8-byte accesses at system scope: an
// aligned 64-bit access is single-copy atomic, so a reader sees the value and the epoch of the same store (a two-word vector
// access would be two scalar accesses that may tear).
__device__ __forceinline__ void st_record(unsigned long long addr, float v, unsigned epoch) {
  const unsigned long long rec = (unsigned long long)__float_as_uint(v) | ((unsigned long long)epoch << 32);
  asm volatile("st.relaxed.sys.global.b64 [%0], %1;" ::"l"(addr), "l"(rec) : "memory");
}
// Waits for a record of this call's epoch.  The wait is bounded so that a peer that never publishes (a rank that failed before its
// launch) ends as a launch failure instead of a GPU that never answers again; the bound is generous (about half a minute of polling)
// because ranks legitimately run a call or more apart: a host stall, a first graph instantiation, another kernel queued ahead.
__device__ __forceinline__ float ld_record(unsigned long long addr, unsigned epoch) {
  unsigned long long rec;
  for (unsigned n = 0;; ++n) {
    asm volatile("ld.relaxed.sys.global.b64 %0, [%1];" : "=l"(rec) : "l"(addr) : "memory");
    if ((unsigned)(rec >> 32) == epoch) break;
    if (n > 64) __nanosleep(n > 4096 ? 1000 : 200);
    if (n > (1u << 25)) __trap();
  }
  return __uint_as_float((unsigned)rec);
}

// Roles run on different tokens at the same time, coupled only by per-stage mbarriers (no CTA-wide barrier in the loop):
//   producer warp  bulk-copies token rows (+ cos / sin rows) into the ring                          -> full[s]
//   sum warps      means of x^2 over the q and k columns of a row                                   -> summed[s]  (or publish / qk_var)
//   gather warp    (peer mode) collects every rank's records of the token from this GPU's buffer    -> gathered[s]
//   apply warps    normalise with the (global) means, round, rotate, store q / k, copy v            -> empty[s] (with the sum warps)
template <class T>
__global__ void __launch_bounds__(kTpThreads, 3) split_qkv_tp_kernel(SplitTpArgs a) {
  using namespace ptx;
  constexpr int V = 8, W = 4, PW = 2;
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ __align__(8) uint64_t s_full[8], s_empty[8], s_summed[8], s_gathered[8];
  __shared__ float s_part[kTpSumWarps][2];
  __shared__ float s_sum[8][2];                   // per stage: this rank's means of the stage's token
  __shared__ float s_gather[8][kTpMaxRanks][2];   // per stage: every rank's means (peer mode)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qh = a.q_hidden, kvh = a.kv_hidden, row = qh + 2 * kvh, hd = a.head_dim;
  const int rd = a.rotary_dim, half = rd / 2;
  const int S = a.stages, stage_bytes = tp_stage_bytes(row, rd);
  const T* s_w = reinterpret_cast<const T*>(smem + S * stage_bytes);  // [q_hidden | kv_hidden] norm weights
  const unsigned batch = (unsigned)a.batch;
  const bool do_sums = a.mode != kTpApply, do_apply = a.mode != kTpSums;
  // every thread reads the call's epoch before any CTA can have finished (the last CTA to finish advances it)
  const unsigned epoch = a.mode == kTpPeer ? *reinterpret_cast<volatile unsigned*>(a.epoch_state) : 0u;
  __shared__ unsigned s_exited;
  if (threadIdx.x == 0) s_exited = 0;

  if (threadIdx.x == 0) {
    const int readers = (do_sums ? kTpSumWarps : 0) + (do_apply ? kTpApplyWarps : 0);
    for (int s = 0; s < S; ++s) {
      mbar_init(smem_u32(&s_full[s]), 1);
      mbar_init(smem_u32(&s_empty[s]), readers);
      mbar_init(smem_u32(&s_summed[s]), 1);
      mbar_init(smem_u32(&s_gathered[s]), 1);
    }
    fence_mbar_init();
  }
  if (do_apply) {
    uint32_t* dst = reinterpret_cast<uint32_t*>(smem + S * stage_bytes);
    for (int i = threadIdx.x; i < qh / 2; i += kTpThreads) dst[i] = reinterpret_cast<const uint32_t*>(a.q_weight)[i];
    for (int i = threadIdx.x; i < kvh / 2; i += kTpThreads) dst[qh / 2 + i] = reinterpret_cast<const uint32_t*>(a.k_weight)[i];
  }
  __syncthreads();
  const uint32_t full0 = smem_u32(&s_full[0]), empty0 = smem_u32(&s_empty[0]);

  auto role = [&]() {
  if (warp == kTpProducerWarp) {  // one token per stage
    if (lane == 0) {
      unsigned it = 0;
      for (unsigned t = blockIdx.x; t < batch; t += gridDim.x, ++it) {
        const int stage = (int)(it % S);
        if (it >= (unsigned)S) mbar_wait_light(empty0 + 8 * stage, ((it / S) - 1) & 1);
        const uint32_t bar = full0 + 8 * stage;
        const uint32_t dst = smem_u32(smem + stage * stage_bytes);
        mbar_arrive_expect_tx(bar, 2u * row + (do_apply ? 4u * rd : 0u));
        bulk_load_1d(dst, reinterpret_cast<const T*>(a.qkv) + (long long)t * row, 2u * row, bar);
        if (do_apply) {
          bulk_load_1d(dst + 2 * row, reinterpret_cast<const T*>(a.cos) + (long long)t * rd, 2u * rd, bar);
          bulk_load_1d(dst + 2 * row + 2 * rd, reinterpret_cast<const T*>(a.sin) + (long long)t * rd, 2u * rd, bar);
        }
      }
    }
    return;
  }

  if (warp == kTpGatherWarp) {  // peer mode: lane r waits for rank r's records of the CTA's tokens, in order, in this GPU's own
    if (a.mode == kTpPeer) {    // buffer (a poll costs an L2 round trip: kept off the working warps)
      const unsigned long long rec0 = (epoch & 1u) * ((unsigned long long)a.tp_world * a.capacity * 16ull);
      unsigned it = 0;
      for (unsigned t = blockIdx.x; t < batch; t += gridDim.x, ++it) {
        const int stage = (int)(it % S);
        if (it >= (unsigned)S) mbar_wait_light(empty0 + 8 * stage, ((it / S) - 1) & 1);  // s_gather[stage] has been read
        if (lane < a.tp_world) {
          const unsigned long long src = a.peer[a.tp_rank] + rec0 + ((unsigned long long)lane * a.capacity + t) * 16ull;
          s_gather[stage][lane][0] = ld_record(src, epoch);
          s_gather[stage][lane][1] = ld_record(src + 8, epoch);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&s_gathered[stage]));
      }
    }
    return;
  }

  if (warp >= kTpApplyWarps) {  // sum warps
    if (!do_sums) return;
    const int sw = warp - kTpApplyWarps, tid = sw * 32 + lane;
    const int nvq = qh / V, nvk = kvh / V;
    const unsigned long long rec0 = (epoch & 1u) * ((unsigned long long)a.tp_world * a.capacity * 16ull);
    unsigned it = 0;
    for (unsigned t = blockIdx.x; t < batch; t += gridDim.x, ++it) {
      const int stage = (int)(it % S);
      mbar_wait_light(full0 + 8 * stage, (it / S) & 1);
      const uint8_t* st = smem + stage * stage_bytes;
      // 16-byte vectors, fp32, two-wide accumulation; q and k columns apart
      float2 aq = make_float2(0.f, 0.f), ak = make_float2(0.f, 0.f);
      for (int v = tid; v < nvq + nvk; v += kTpSumWarps * 32) {
        const uint4 r = *reinterpret_cast<const uint4*>(st + v * 16);
        const uint32_t w4[4] = {r.x, r.y, r.z, r.w};
        float2 acc = make_float2(0.f, 0.f);
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          const float2 x = make_float2(Act<T>::lo(w4[p]), Act<T>::hi(w4[p]));
          acc = fma2(x, x, acc);
        }
        if (v < nvq) { aq.x += acc.x; aq.y += acc.y; } else { ak.x += acc.x; ak.y += acc.y; }
      }
      float sq = aq.x + aq.y, sk = ak.x + ak.y;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        sq += __shfl_xor_sync(0xffffffffu, sq, o);
        sk += __shfl_xor_sync(0xffffffffu, sk, o);
      }
      if (lane == 0) { s_part[sw][0] = sq; s_part[sw][1] = sk; }
      named_bar_sync(1, kTpSumWarps * 32);
      sq = sk = 0.f;
#pragma unroll
      for (int w = 0; w < kTpSumWarps; ++w) { sq += s_part[w][0]; sk += s_part[w][1]; }
      const float var_q = sq / (float)qh, var_k = sk / (float)kvh;
      if (a.mode == kTpPeer) {
        if (tid < a.tp_world) {  // one thread per destination rank (this rank's own buffer included)
          const unsigned long long dst = a.peer[tid] + rec0 + ((unsigned long long)a.tp_rank * a.capacity + t) * 16ull;
          st_record(dst, var_q, epoch);
          st_record(dst + 8, var_k, epoch);
        }
      } else if (tid == 0) {
        if (a.mode == kTpSums) { a.qk_var[2 * (long long)t] = var_q; a.qk_var[2 * (long long)t + 1] = var_k; }
        else { s_sum[stage][0] = var_q; s_sum[stage][1] = var_k; mbar_arrive(smem_u32(&s_summed[stage])); }
      }
      named_bar_sync(1, kTpSumWarps * 32);  // s_part may be rewritten; both warps are done with the stage
      if (lane == 0) mbar_arrive(empty0 + 8 * stage);
    }
    return;
  }

  // apply warps.  A head is hd / 8 lanes; a lane owns two 4-element slices chosen so that a neox pair sits in one lane.
  if (!do_apply) return;
  const int L = hd / V, heads_per_warp = 32 / L;
  const int j = lane % L;
  const bool in_rope = j * W < half;
  const int oA = in_rope ? j * W : rd + V * (j - half / W);
  const int oB = in_rope ? half + j * W : oA + W;
  const int nq = qh / hd, nk = kvh / hd;
  unsigned it = 0;
  for (unsigned t = blockIdx.x; t < batch; t += gridDim.x, ++it) {
    const int stage = (int)(it % S);
    const uint8_t* st = smem + stage * stage_bytes;
    mbar_wait_light(full0 + 8 * stage, (it / S) & 1);
    float var_q = 0.f, var_k = 0.f;
    if (a.mode == kTpFused) {
      mbar_wait_light(smem_u32(&s_summed[stage]), (it / S) & 1);
      var_q = s_sum[stage][0];
      var_k = s_sum[stage][1];
    } else if (a.mode == kTpPeer) {
      mbar_wait_light(smem_u32(&s_gathered[stage]), (it / S) & 1);
      for (int r = 0; r < a.tp_world; ++r) { var_q += s_gather[stage][r][0]; var_k += s_gather[stage][r][1]; }  // rank order: same bits everywhere
    } else {
      var_q = a.qk_var[2 * (long long)t];
      var_k = a.qk_var[2 * (long long)t + 1];
    }
    // 1/sqrt by the hardware approximation and one Newton step (below one fp32 ulp), as in the vector kernel
    const float mq = __fadd_rn(var_q * a.inv_tp_world, a.eps), mk = __fadd_rn(var_k * a.inv_tp_world, a.eps);
    const float yq = rsqrtf(mq), yk = rsqrtf(mk);
    const float scale_q = fmaf(yq, fmaf(-(0.5f * mq) * yq, yq, 0.5f), yq);
    const float scale_k = fmaf(yk, fmaf(-(0.5f * mk) * yk, yk, 0.5f), yk);
    {  // v: copy
      const uint4* src = reinterpret_cast<const uint4*>(st + 2 * (qh + kvh));
      uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<T*>(a.v_out) + (long long)t * kvh);
      for (int v = threadIdx.x; v < kvh / V; v += kTpApplyWarps * 32) dst[v] = src[v];
    }
    // cos / sin of the lane's pairs (one value per pair), sign folded into the first slice's sin
    float2 cs[PW], snA[PW], snB[PW];
    if (in_rope) {
      const uint2 rc = *reinterpret_cast<const uint2*>(st + 2 * row + oA * 2), rs = *reinterpret_cast<const uint2*>(st + 2 * row + 2 * rd + oA * 2);
      cs[0] = make_float2(Act<T>::lo(rc.x), Act<T>::hi(rc.x));
      cs[1] = make_float2(Act<T>::lo(rc.y), Act<T>::hi(rc.y));
      snB[0] = make_float2(Act<T>::lo(rs.x), Act<T>::hi(rs.x));
      snB[1] = make_float2(Act<T>::lo(rs.y), Act<T>::hi(rs.y));
      snA[0] = make_float2(-snB[0].x, -snB[0].y);
      snA[1] = make_float2(-snB[1].x, -snB[1].y);
    }
    for (int h0 = warp * heads_per_warp; h0 < nq + nk; h0 += kTpApplyWarps * heads_per_warp) {
      const int head = h0 + lane / L;
      if (head >= nq + nk) continue;
      const bool is_q = head < nq;
      const int col = head * hd;  // in the fused row and in [q weights | k weights]
      const float sc = is_q ? scale_q : scale_k;
      const float2 s2 = make_float2(sc, sc);
      const uint2 ra = *reinterpret_cast<const uint2*>(st + (col + oA) * 2), rb = *reinterpret_cast<const uint2*>(st + (col + oB) * 2);
      const uint2 wa = *reinterpret_cast<const uint2*>(s_w + col + oA), wb = *reinterpret_cast<const uint2*>(s_w + col + oB);
      const uint32_t xa[2] = {ra.x, ra.y}, xb[2] = {rb.x, rb.y}, wwa[2] = {wa.x, wa.y}, wwb[2] = {wb.x, wb.y};
      uint32_t oa[2], ob[2];
#pragma unroll
      for (int p = 0; p < PW; ++p) {
        // y = act_rne(x * scale * w): the reference stores the normalised row before it rotates it
        float2 ya = mul2(mul2(make_float2(Act<T>::lo(xa[p]), Act<T>::hi(xa[p])), s2), make_float2(Act<T>::lo(wwa[p]), Act<T>::hi(wwa[p])));
        float2 yb = mul2(mul2(make_float2(Act<T>::lo(xb[p]), Act<T>::hi(xb[p])), s2), make_float2(Act<T>::lo(wwb[p]), Act<T>::hi(wwb[p])));
        oa[p] = Act<T>::pack(ya.x, ya.y);
        ob[p] = Act<T>::pack(yb.x, yb.y);
        if (in_rope) {
          ya = make_float2(Act<T>::lo(oa[p]), Act<T>::hi(oa[p]));
          yb = make_float2(Act<T>::lo(ob[p]), Act<T>::hi(ob[p]));
          const float2 na = fma2(yb, snA[p], mul2(ya, cs[p]));
          const float2 nb = fma2(ya, snB[p], mul2(yb, cs[p]));
          oa[p] = Act<T>::pack(na.x, na.y);
          ob[p] = Act<T>::pack(nb.x, nb.y);
        }
      }
      T* dst = is_q ? reinterpret_cast<T*>(a.q_out) + (long long)t * qh + col : reinterpret_cast<T*>(a.k_out) + (long long)t * kvh + (col - qh);
      *reinterpret_cast<uint2*>(dst + oA) = make_uint2(oa[0], oa[1]);
      *reinterpret_cast<uint2*>(dst + oB) = make_uint2(ob[0], ob[1]);
    }
    __syncwarp();  // the warp is done with the stage
    if (lane == 0) mbar_arrive(empty0 + 8 * stage);
  }
  };
  role();
  // peer mode: the last warp of the last CTA to finish advances the epoch for the next call (0xFFFFFFFE -> 1 keeps the parity alternating)
  if (a.mode == kTpPeer && lane == 0 && atomicAdd(&s_exited, 1u) == kTpThreads / 32 - 1) {
    __threadfence();
    if (atomicAdd(a.epoch_state + 1, 1u) == gridDim.x - 1) {
      a.epoch_state[1] = 0;
      __threadfence();
      *reinterpret_cast<volatile unsigned*>(a.epoch_state) = epoch >= 0xFFFFFFFEu ? 1u : epoch + 1u;
    }
  }
}

}  // namespace mlapo
