// mla_preprocess for B200 (sm_100a): hand-written CUDA kernels + the C-ABI launcher.
// This translation unit holds the host side (validation, tensor maps, launches) and the extern "C" entry points; the
// kernels live in mlapo_quant_norm.cuh (k0 k1 k3), mlapo_mm1.cuh (k2), mlapo_q.cuh (k4), mlapo_q_pair.cuh (k7),
// on top of mlapo_numerics.cuh and the inline-PTX layer sm100_ptx.cuh.
//
// Path (reference phases P1..P6, csrc/mla_preprocess/op_kernel/mla_preprocess_mix_bf16.hpp:2773-2916):
//   k1 quant_hidden        hidden[N,D] act -> a0 int8                       (Quant, bf16.hpp:283-534)
//   k2 mm1_splitk          a0 . Wdqkv^T -> split-K int32 partial sums       (PpMatmulW8a8Aic, bf16.hpp:1515-1975)
//   k3 norm_rope_cache     sum partials, +bias, *descale -> act y0;
//                          q_a RMSNorm(+beta)+quant -> a1 int8              (RmsNormQuant, bf16.hpp:536-840)
//                          latent RMSNorm(512), RoPE(64), paged-cache scatter (RmsNormAndRopeConvergence1, :2499-2686)
//   k4 q_fused (per head)  a1 . Wuq[h]^T, dequant, RoPE -> q_out1; nope slice stays in smem and is absorbed:
//                          q_nope . Wuk[h] -> q_out0 (+ int8 re-quant)      (PpMatmulW8a8Aiv, RopeFp16, PpMatmulEinSum, EinSumQuant)
//   k7 q_pair              k4's arithmetic on CTA pairs (tcgen05 cta_group::2) streaming units, when units outnumber the SMs (prefill)
// k1..k4 are chained with programmatic dependent launch: each requests its weight tiles, then griddepcontrol.wait.  (A single
// persistent cooperative kernel holding all four phases existed in rounds 1-2; it was measured slower at every decode size, even
// with a 1.7 us grid barrier, and was removed: DESIGN.md §5.)
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

#include <atomic>
#include <map>
#include <mutex>
#include <tuple>
#include <type_traits>

#include "../../include/mla_preprocess_b200.h"
#include "sm100_ptx.cuh"
#include "mlapo_numerics.cuh"
#include "mlapo_quant_norm.cuh"
#include "mlapo_mm1.cuh"
#include "mlapo_mm1_pair.cuh"
#include "mlapo_q.cuh"
#include "mlapo_q_pair.cuh"
#include "mlapo_format.cuh"
#include "mlapo_bmt.cuh"
#include "mlapo_slots.cuh"
#include "mlapo_split_qkv.cuh"
#include "mlapo_split_qkv_tp.cuh"

namespace mlapo {


// ============================================================================================ host side
thread_local char g_err[512] = "";
// Debug flags (test / profiling only, see the header): one process-wide word; every entry point takes ONE snapshot of it into the
// calling thread's copy, so a call sees a single consistent value from its workspace layout to its last launch.
static std::atomic<int> g_debug_flags_word{0};
thread_local int g_debug_flags = 0;  // bit0: sync + check after each kernel; bit1: write the y0 tap; bit2: time each kernel
static inline void snapshot_debug_flags() { g_debug_flags = g_debug_flags_word.load(std::memory_order_relaxed); }
static unsigned* g_wait_log_host = nullptr;  // host-mapped log of stuck barrier waits (debug flag bit11)
constexpr int kMaxKernels = 8;
constexpr int kMaxDevices = 64;
// Per-kernel timing events (debug bit2) belong to the device they were created on; the names and the count of the last launch
// belong to the thread that made it.
static cudaEvent_t g_ev[kMaxDevices][kMaxKernels + 1];
static std::atomic<bool> g_ev_ready[kMaxDevices];
static std::mutex g_ev_mutex;
thread_local int g_ev_count = 0;  // events recorded by the last launch of this thread
thread_local int g_ev_dev = 0;    // ... on this device
thread_local const char* g_kernel_names[kMaxKernels] = {"quant_hidden", "mm1_splitk", "norm_rope_cache", "q_fused", "", "", "", ""};
static int current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return (dev >= 0 && dev < kMaxDevices) ? dev : 0;
}
static void record_event(int i, cudaStream_t st) {
  if (!(g_debug_flags & 4)) return;
  const int dev = current_device();
  if (!g_ev_ready[dev].load(std::memory_order_acquire)) {
    std::lock_guard<std::mutex> lock(g_ev_mutex);
    if (!g_ev_ready[dev].load(std::memory_order_relaxed)) {
      for (int k = 0; k <= kMaxKernels; ++k) cudaEventCreate(&g_ev[dev][k]);
      g_ev_ready[dev].store(true, std::memory_order_release);
    }
  }
  cudaEventRecord(g_ev[dev][i], st);
  g_ev_count = i + 1;
  g_ev_dev = dev;
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

// SM count of the CURRENT device (cached per device ordinal; the write is idempotent, so a benign race at first use is harmless)
static int num_sms() {
  static std::atomic<int> sms[kMaxDevices];
  const int dev = current_device();
  int v = sms[dev].load(std::memory_order_relaxed);
  if (!v) {
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    if (v <= 0) v = 148;
    sms[dev].store(v, std::memory_order_relaxed);
  }
  return v;
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

// Calls of more than 128 tokens run MM1 on CTA pairs (mlapo_mm1_pair.cuh): a work item then occupies two SMs.  Debug bit16
// keeps the single-CTA form (A/B).
static bool mm1_on_pairs(int n_tokens) { return n_tokens > kBlockM && !(g_debug_flags & 65536); }
// ... with 128-column weight tiles for one block of 256 tokens, 256-column tiles for more (debug bit19 keeps 128)
static int mm1_pair_bn(int n_tokens) { return (n_tokens > 2 * kBlockM && !(g_debug_flags & 524288)) ? kMm1BlockNWide : kMm1BlockN; }

static int choose_splits0(int n_tokens, int hidden_dim) {
  const int m_super = n_tokens > 0 ? (n_tokens + 2 * kBlockM - 1) / (2 * kBlockM) : 1;  // work items cover up to 256 tokens
  const int bn = mm1_on_pairs(n_tokens) ? mm1_pair_bn(n_tokens) : kMm1BlockN;
  const int n_tiles = (MLAPO_MM1_OUT + bn - 1) / bn;                                     // 17 (or 9)
  const int kb_total = hidden_dim / kBlockKBytes;
  int s = (mm1_on_pairs(n_tokens) ? num_sms() / 2 : num_sms()) / (n_tiles * m_super);  // one wave of CTAs (or CTA pairs)
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

// Resident blocks per SM of a kernel without dynamic shared memory, cached per device at the call site (`slot` = a static
// std::atomic<int>[kMaxDevices]).  Grid-stride kernels are launched with exactly one wave: a second, partial wave of blocks that
// each take the same share of the rows would run at a fraction of the occupancy for as long as the first.
template <class K>
static int resident_blocks(K kern, int threads, std::atomic<int>* slot, int fallback) {
  int v = slot->load(std::memory_order_relaxed);
  if (!v) {
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, kern, threads, 0) != cudaSuccess || v <= 0) { cudaGetLastError(); v = fallback; }
    slot->store(v, std::memory_order_relaxed);
  }
  return v;
}

// Shared-memory opt-in of one kernel instantiation, once per device.  `done` is a per-call-site latch indexed by device ordinal;
// two threads racing on the first call both set the (idempotent) attribute and both store `true`.
using DoneFlag = std::atomic<bool>;
template <class K>
static int set_smem_attr(K kern, int bytes, DoneFlag* done, const char* name) {
  if (done->load(std::memory_order_acquire)) return MLAPO_OK;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess)
    return fail(MLAPO_ERR_CUDA, "cudaFuncSetAttribute(%s, %d): %s", name, bytes, cudaGetErrorString(cudaGetLastError()));
  done->store(true, std::memory_order_release);
  return MLAPO_OK;
}

// Token rows per q-path CTA: two token tiles per CTA stream each head's weights once per 256 tokens; when that leaves most
// SMs idle (few heads, e.g. a TP slice) one tile per CTA doubles the CTA count at the price of re-reading the head's
// weights from L2.
static int q_cta_tokens(int n, int h) { return (n > kBlockM && (long long)h * ((n + 255) / 256) >= 96) ? 256 : 128; }

template <class T>
static int launch_impl(const MlapoParams* p, uint8_t* ws, const MlapoWorkspaceLayout& l, cudaStream_t st) {
  const int n = p->num_tokens, d = p->hidden_dim, h = p->num_heads;
  const bool per_token = p->quant_mode == MLAPO_QUANT_PER_TOKEN_SYMM;
  const bool int8_out = p->cache_mode == MLAPO_CACHE_INT8_NZ;
  const bool pdl = (g_debug_flags & 5) == 0;  // per-kernel sync/timing wants serialised kernels
  const int input_complete = (p->flags & MLAPO_FLAG_INPUT_COMPLETE) ? 1 : 0;
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
  const int dev = current_device();

  // ---- tensor maps and argument blocks shared by both launch shapes
  const bool wide = l.splits0 == 1 && n >= 1024;  // prefill: no split-K, 256-wide weight tiles
  const int bn = wide ? kMm1BlockNWide : kMm1BlockN;
  CUtensorMap tm_a0, tm_wdqkv, tm_part, tm_a1, tm_wuq, tm_wuk, tm_o0, tm_o1;
  if ((rc = make_tmap(&tm_a0, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a0, n, d, d, kBlockM, 128))) return rc;
  if ((rc = make_tmap(&tm_wdqkv, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, p->wdqkv_km, MLAPO_MM1_OUT, d, d, bn, 128))) return rc;
  if ((rc = make_tmap(&tm_part, CU_TENSOR_MAP_DATA_TYPE_INT32, 4, partial0, n, MLAPO_MM1_OUT, MLAPO_MM1_OUT * 4, kBlockM, 32, l.splits0,
                      (uint64_t)n * MLAPO_MM1_OUT * 4))) return rc;
  // prefill shape: MM1 dequantises in its epilogue and writes act y0 [N, 2112] where the int32 sums would have gone (or, with the
  // debug tap on, into the tap region, the raw sums then going to their usual place as well)
  T* const y0_buf = (g_debug_flags & 2) ? reinterpret_cast<T*>(ws + l.y0_tap) : reinterpret_cast<T*>(partial0);
  CUtensorMap tm_y0;
  if (wide && (rc = make_tmap(&tm_y0, act_dt, 2, y0_buf, n, MLAPO_MM1_OUT, MLAPO_MM1_OUT * 2, kBlockM, 64))) return rc;
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
  na.y0_tap = ((g_debug_flags & 2) && !wide) ? ws + l.y0_tap : nullptr;
  na.y0_in = wide ? y0_buf : nullptr;
  na.kv_cache = p->kv_cache; na.kv_cache_rope = p->kv_cache_rope;
  na.cache_mode = p->cache_mode;
  na.cache_slots = (long long)p->num_cache_blocks * p->cache_block_size;
  na.tl = tl3;
  na.input_complete = input_complete;

  QArgs qa{};
  qa.descale1 = p->descale1; qa.descale_is_i64 = p->descale_is_int64;
  qa.bias1 = per_token ? nullptr : reinterpret_cast<const int32_t*>(p->bias1);
  qa.tok_scale1 = per_token ? tok_scale + n : nullptr;
  qa.cos = p->cos; qa.sin = p->sin; qa.q_nope_scale = p->q_nope_scale;
  qa.num_heads = h; qa.m_total = n; qa.cta_tokens = cta_tokens;
  qa.debug_drain_delay = (g_debug_flags & 1048576) ? 6000 : 0;  // debug bit20: late drain groups in the pair kernel (regression test)
  const int q_tok_tiles = (n + cta_tokens - 1) / cta_tokens;

  g_kernel_names[0] = "quant_hidden";
  g_kernel_names[1] = "mm1_splitk";

  // ---- k1: quantise hiddenState (bit12 of the debug flags: also ask for the call's weights in L2 — measured slower, see profiles/README.md)
  WeightPrefetch wp{};
  if (g_debug_flags & 4096) {
    wp.ptr[0] = reinterpret_cast<const char*>(p->wdqkv_km); wp.lines[0] = (long long)MLAPO_MM1_OUT * d / 128;
    wp.ptr[1] = reinterpret_cast<const char*>(p->wuq_km);   wp.lines[1] = (long long)h * MLAPO_Q_HEAD * MLAPO_Q_LORA / 128;
    wp.ptr[2] = reinterpret_cast<const char*>(p->wuk);      wp.lines[2] = (long long)h * MLAPO_Q_NOPE * MLAPO_KV_LORA * 2 / 128;
  }
  record_event(0, st);
  if (per_token) {
    e__ = launch_k(quant_hidden_per_token_kernel<T>, dim3(n), dim3(256), 0, st, pdl, reinterpret_cast<const T*>(p->hidden_state), a0,
                   tok_scale, d, wp);
  } else {
    const long long nvec = (long long)n * d / 8;
    // Blocks of 128 threads (4 warps x 40 registers, no shared memory): small enough that one of them fits on an SM next to a
    // q_fused CTA of the previous call (10 warps x 168 registers and all but 1.25 KB of the SM's shared memory), so that under
    // MLAPO_FLAG_INPUT_COMPLETE the quantiser runs while that kernel is still busy instead of queueing for its SMs.
    constexpr int qt = 128;
    {  // same shared-memory carve-out as the q kernel, or the SM would have to drain before it can take a quantise block
      static DoneFlag done[kMaxDevices];
      if (!done[dev].load(std::memory_order_acquire)) {
        cudaFuncSetAttribute(quant_hidden_kernel<T, 1>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        cudaFuncSetAttribute(quant_hidden_kernel<T, 2>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        cudaFuncSetAttribute(quant_hidden_kernel<T, 4>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        done[dev].store(true, std::memory_order_release);
      }
    }
    const long long wave = (long long)num_sms() * 8 * 256;
    if (nvec <= wave / 2) {  // one 16-byte vector per thread
      e__ = launch_k(quant_hidden_kernel<T, 1>, dim3((unsigned)((nvec + qt - 1) / qt)), dim3(qt), 0, st, pdl,
                     reinterpret_cast<const T*>(p->hidden_state), a0, reinterpret_cast<const T*>(p->quant_scale0),
                     reinterpret_cast<const int8_t*>(p->quant_offset0), nvec, wp, tl0, input_complete);
    } else if (nvec <= wave) {  // two vectors per thread, both loads in flight
      e__ = launch_k(quant_hidden_kernel<T, 2>, dim3((unsigned)((nvec + 2 * qt - 1) / (2 * qt))), dim3(qt), 0, st, pdl,
                     reinterpret_cast<const T*>(p->hidden_state), a0, reinterpret_cast<const T*>(p->quant_scale0),
                     reinterpret_cast<const int8_t*>(p->quant_offset0), nvec, wp, tl0, input_complete);
    } else if (nvec <= 2 * wave) {  // four vectors per thread, still in the small blocks (up to ~670 tokens at D = 7168)
      e__ = launch_k(quant_hidden_kernel<T, 4>, dim3((unsigned)((nvec + 4 * qt - 1) / (4 * qt))), dim3(qt), 0, st, pdl,
                     reinterpret_cast<const T*>(p->hidden_state), a0, reinterpret_cast<const T*>(p->quant_scale0),
                     reinterpret_cast<const int8_t*>(p->quant_offset0), nvec, wp, tl0, input_complete);
    } else {                 // prefill sizes: grid-stride over one resident wave, 4 independent loads in flight per thread
      static std::atomic<int> occ[kMaxDevices];   // U = 8 selects the form on pairs of vectors (32 bytes in, 16 out per thread and step)
      e__ = launch_k(quant_hidden_kernel<T, 8>, dim3(num_sms() * resident_blocks(quant_hidden_kernel<T, 8>, 256, &occ[dev], 4)), dim3(256), 0, st, pdl,
                     reinterpret_cast<const T*>(p->hidden_state), a0, reinterpret_cast<const T*>(p->quant_scale0),
                     reinterpret_cast<const int8_t*>(p->quant_offset0), nvec, wp, tl0, input_complete);
    }
  }
  MLAPO_CHECK_LAUNCH("quant_hidden");
  record_event(1, st);

  // ---- k2: MM1, split-K partial sums
  {
    const int n_tiles = (MLAPO_MM1_OUT + bn - 1) / bn;
    if (wide) {
      static DoneFlag done[kMaxDevices];
      constexpr int smem = mm1_smem_bytes<2, kMm1BlockNWide, true>();
      if ((rc = set_smem_attr(mm1_dequant_kernel<2, kMm1BlockNWide, T>, smem, &done[dev], "mm1_dequant<2,256>"))) return rc;
      Mm1DequantArgs dq{};
      dq.bias0 = na.bias0; dq.descale0 = p->descale0; dq.tok_scale0 = na.tok_scale0; dq.descale_is_i64 = p->descale_is_int64;
      dq.tap_acc = (g_debug_flags & 2) ? 1 : 0;
      e__ = launch_k(mm1_dequant_kernel<2, kMm1BlockNWide, T>, dim3(n_tiles, 1, (n + 255) / 256), dim3(kGemmThreads), smem, st, pdl,
                     tm_a0, tm_wdqkv, tm_part, tm_y0, dq, n, d / kBlockKBytes, tl1);
      g_kernel_names[1] = "mm1_dequant";
    } else if (mm1_on_pairs(n)) {
      const int pbn = mm1_pair_bn(n), p_tiles = (MLAPO_MM1_OUT + pbn - 1) / pbn;
      CUtensorMap tm_wh;  // this form fetches the weight tile in two halves, one per CTA of the pair
      if ((rc = make_tmap(&tm_wh, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, p->wdqkv_km, MLAPO_MM1_OUT, d, d, pbn / 2, 128))) return rc;
      const dim3 pgrid(2 * p_tiles, l.splits0, (n + 255) / 256);
      if (pbn == kMm1BlockN) {
        static DoneFlag done[kMaxDevices];
        if ((rc = set_smem_attr(mm1_pair_kernel<kMm1BlockN>, Mm1Pair<kMm1BlockN>::SMEM_BYTES, &done[dev], "mm1_pair<128>"))) return rc;
        e__ = launch_k(mm1_pair_kernel<kMm1BlockN>, pgrid, dim3(kGemmThreads), Mm1Pair<kMm1BlockN>::SMEM_BYTES, st, pdl, tm_a0, tm_wh, tm_part, n,
                       d / kBlockKBytes, tl1);
      } else {
        static DoneFlag done[kMaxDevices];
        if ((rc = set_smem_attr(mm1_pair_kernel<kMm1BlockNWide>, Mm1Pair<kMm1BlockNWide>::SMEM_BYTES, &done[dev], "mm1_pair<256>"))) return rc;
        e__ = launch_k(mm1_pair_kernel<kMm1BlockNWide>, pgrid, dim3(kGemmThreads), Mm1Pair<kMm1BlockNWide>::SMEM_BYTES, st, pdl, tm_a0, tm_wh,
                       tm_part, n, d / kBlockKBytes, tl1);
      }
      g_kernel_names[1] = "mm1_pair";
    } else if (n > kBlockM) {
      static DoneFlag done[kMaxDevices];
      if ((rc = set_smem_attr(mm1_splitk_kernel<2, kMm1BlockN>, mm1_smem_bytes<2, kMm1BlockN>(), &done[dev], "mm1<2,128>"))) return rc;
      e__ = launch_k(mm1_splitk_kernel<2, kMm1BlockN>, dim3(n_tiles, l.splits0, (n + 255) / 256), dim3(kGemmThreads),
                     mm1_smem_bytes<2, kMm1BlockN>(), st, pdl, tm_a0, tm_wdqkv, tm_part, n, d / kBlockKBytes, tl1);
    } else {
      static DoneFlag done[kMaxDevices];
      if ((rc = set_smem_attr(mm1_splitk_kernel<1, kMm1BlockN>, mm1_smem_bytes<1, kMm1BlockN>(), &done[dev], "mm1<1,128>"))) return rc;
      e__ = launch_k(mm1_splitk_kernel<1, kMm1BlockN>, dim3(n_tiles, l.splits0, 1), dim3(kGemmThreads), mm1_smem_bytes<1, kMm1BlockN>(),
                     st, pdl, tm_a0, tm_wdqkv, tm_part, n, d / kBlockKBytes, tl1);
    }
    MLAPO_CHECK_LAUNCH("mm1_splitk");
  }
  record_event(2, st);

  // ---- k3: norms, quant, rope, cache scatter
  if (wide && !(g_debug_flags & 131072)) {  // prefill: one warp per row of the dequantised y0 (debug bit17 keeps the CTA-per-row form)
    const int rows_per_block = kNormWarpThreads / 32;
    static std::atomic<int> occ[kMaxDevices];
    const int want = (n + rows_per_block - 1) / rows_per_block;
    const int cap = num_sms() * resident_blocks(norm_rows_warp_kernel<T>, kNormWarpThreads, &occ[dev], 2);
    e__ = launch_k(norm_rows_warp_kernel<T>, dim3(want < cap ? want : cap), dim3(kNormWarpThreads), 0, st, pdl, na);
  } else if (wide) e__ = launch_k(norm_rope_cache_kernel<T, 0>, dim3(n), dim3(kNormThreads), 0, st, pdl, na);
  else if (l.splits0 == 1) e__ = launch_k(norm_rope_cache_kernel<T, 1>, dim3(n), dim3(kNormThreads), 0, st, pdl, na);
  else e__ = launch_k(norm_rope_cache_kernel<T, 8>, dim3(n), dim3(kNormThreads), 0, st, pdl, na);
  MLAPO_CHECK_LAUNCH("norm_rope_cache");
  record_event(3, st);

  // ---- k4 / k7: fused MM2 + dequant + RoPE + W_UK absorb.  One CTA per (head, token super-tile) while those fit one wave;
  //      beyond that, persistent CTA pairs stream the units so that consecutive units overlap.
  const int q_units = h * ((n + 255) / 256);
  // k7 wins where a pair has many units to stream (C3: 494 us for the q path against 558 us for the single-CTA streaming kernel
  // it replaced); with one or two units per pair (decode) its longer dependency chain loses to k4 (C2: 33 vs 27 us), so it runs
  // only when units outnumber the SMs.  Debug bit13 forces it for n > 128 (parity tests).
  const bool many_units = n > 256 && q_units > num_sms();
  if (many_units || ((g_debug_flags & 8192) && n > kBlockM)) {
    // ---- k7: CTA pairs.  Each CTA of a pair owns one 128-token tile and half of the head's weight rows.
    int rcx;
    CUtensorMap tm_a1p, tm_wuqp;
    if ((rcx = make_tmap(&tm_a1p, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, a1, n, MLAPO_Q_LORA, MLAPO_Q_LORA, kBlockM, 128))) return rcx;
    if ((rcx = make_tmap(&tm_wuqp, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, p->wuq_km, (uint64_t)h * MLAPO_Q_HEAD, MLAPO_Q_LORA, MLAPO_Q_LORA,
                         kPHalfB, 128))) return rcx;
    const int max_pairs = num_sms() / 2;
    dim3 grid(2 * (q_units < max_pairs ? q_units : max_pairs));
    if (int8_out) {
      static DoneFlag done[kMaxDevices];
      if ((rc = set_smem_attr(q_pair_kernel<T, true>, kPSmemBytes, &done[dev], "q_pair<int8>"))) return rc;
      e__ = launch_k(q_pair_kernel<T, true>, grid, dim3(kPThreads), kPSmemBytes, st, pdl, tm_a1p, tm_wuqp, tm_wuk, tm_o0, qa, p->q_out1, q_units, tl4);
    } else {
      static DoneFlag done[kMaxDevices];
      if ((rc = set_smem_attr(q_pair_kernel<T, false>, kPSmemBytes, &done[dev], "q_pair"))) return rc;
      e__ = launch_k(q_pair_kernel<T, false>, grid, dim3(kPThreads), kPSmemBytes, st, pdl, tm_a1p, tm_wuqp, tm_wuk, tm_o0, qa, p->q_out1, q_units, tl4);
    }
    MLAPO_CHECK_LAUNCH("q_pair");
    g_kernel_names[3] = "q_pair";
  } else {
    g_kernel_names[3] = "q_fused";
    // Few (head, token tile) units: 2 or 4 CTAs share a unit (each recomputes the nope slice of MM2 and takes its share of the
    // RoPE pairs and of the absorb chunks) as long as the grid still fits one wave.  Debug bit14 keeps one CTA per unit (A/B).
    const int units = h * q_tok_tiles;
    const int split = (g_debug_flags & 16384) ? 1 : (units * 4 <= num_sms() ? 4 : units * 2 <= num_sms() ? 2 : 1);
    CUtensorMap tm_wuq_n = tm_wuq, tm_wuq_r = tm_wuq;
    if (split > 1) {
      if ((rc = make_tmap(&tm_wuq_n, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, p->wuq_km, (uint64_t)h * MLAPO_Q_HEAD, MLAPO_Q_LORA, MLAPO_Q_LORA,
                          MLAPO_Q_NOPE, 128))) return rc;
      if ((rc = make_tmap(&tm_wuq_r, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, p->wuq_km, (uint64_t)h * MLAPO_Q_HEAD, MLAPO_Q_LORA, MLAPO_Q_LORA,
                          32 / split, 128))) return rc;
    }
    qa.q_out1 = p->q_out1;
    dim3 grid(h, q_tok_tiles, split);
#define MLAPO_LAUNCH_QF(I8, SP)                                                                                                   \
  do {                                                                                                                            \
    static DoneFlag done[kMaxDevices];                                                                                            \
    if ((rc = set_smem_attr(q_fused_kernel<T, I8, SP>, kQSmemBytes, &done[dev], "q_fused"))) return rc;                           \
    e__ = launch_k(q_fused_kernel<T, I8, SP>, grid, dim3(kQThreads), kQSmemBytes, st, pdl, tm_a1, tm_wuq, tm_wuq_n, tm_wuq_r,     \
                   tm_wuk, tm_o0, tm_o1, qa, tl4);                                                                                \
  } while (0)
    if (int8_out) {
      if (split == 4) MLAPO_LAUNCH_QF(true, 4); else if (split == 2) MLAPO_LAUNCH_QF(true, 2); else MLAPO_LAUNCH_QF(true, 1);
    } else {
      if (split == 4) MLAPO_LAUNCH_QF(false, 4); else if (split == 2) MLAPO_LAUNCH_QF(false, 2); else MLAPO_LAUNCH_QF(false, 1);
    }
#undef MLAPO_LAUNCH_QF
    MLAPO_CHECK_LAUNCH("q_fused");
  }
  record_event(4, st);
  return MLAPO_OK;
}

// launch of one instantiation of the staged split_qkv kernel; the shared-memory opt-in is made once per instantiation and device
template <class T, int HD, bool NEOX>
static cudaError_t split_vec_launch(const SplitQkvArgs& a, cudaStream_t st) {
  using G = SplitGeom<HD>;
  static DoneFlag opted[kMaxDevices];
  const int dev = current_device();
  cudaError_t e = cudaSuccess;
  if (!opted[dev].load(std::memory_order_acquire)) {
    e = cudaFuncSetAttribute(split_qkv_vec_kernel<T, HD, NEOX>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    opted[dev].store(true, std::memory_order_release);
  }
  // CTAs per chunk index of the row, in proportion to the chunk's cost (its length plus a fixed part per token), 3 CTAs per SM
  const long long row = (a.gated ? 2LL : 1LL) * a.q_hidden + 2LL * a.kv_hidden;
  const unsigned cpt = (unsigned)((row + G::CHUNK - 1) / G::CHUNK);
  const long long last_len = row - (long long)(cpt - 1) * G::CHUNK;
  const long long w_full = G::CHUNK + 512, w_last = last_len + 512;
  const unsigned ctas = (unsigned)num_sms() * 3u;
  unsigned g_last = (unsigned)((ctas * w_last + ((cpt - 1) * w_full + w_last) / 2) / ((cpt - 1) * w_full + w_last));
  if (g_last < 1) g_last = 1;
  unsigned g_full = cpt > 1 ? (ctas - g_last) / (cpt - 1) : 0;
  if (cpt > 1 && g_full < 1) g_full = 1;
  const unsigned batch = (unsigned)a.batch;
  if (g_last > batch) g_last = batch;
  if (g_full > batch) g_full = batch;
  split_qkv_vec_kernel<T, HD, NEOX><<<(cpt - 1) * g_full + g_last, kSplitThreads, G::SMEM_BYTES, st>>>(a, cpt, g_full, g_last);
  return cudaGetLastError();
}

template <class T>
static int bmt_launch_impl(const void* a, const void* b, void* c, int64_t m, int64_t batch, int64_t k, int64_t n, int b_is_nz,
                           cudaStream_t st) {
  const CUtensorMapDataType dt = std::is_same<T, __nv_bfloat16>::value ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  cudaError_t e;
  const bool tc = !b_is_nz && k % 64 == 0 && n % 64 == 0 && m < (1ll << 31) && batch <= 65535 &&
                  ((uintptr_t)a % 16 == 0) && ((uintptr_t)b % 16 == 0) && ((uintptr_t)c % 16 == 0);
  if (tc) {
    int rc;
    const int dev = current_device();
    CUtensorMap tm_a, tm_b, tm_c;  // coordinates (inner, rows, batch); the op's "transposes" are just these stride orders
    if ((rc = make_tmap(&tm_a, dt, 2, a, (uint64_t)m, (uint64_t)k, (uint64_t)batch * k * 2, kBlockM, 64, (uint64_t)batch, (uint64_t)k * 2))) return rc;
    if ((rc = make_tmap(&tm_b, dt, 2, b, (uint64_t)k, (uint64_t)n, (uint64_t)n * 2, 64, 64, (uint64_t)batch, (uint64_t)k * n * 2))) return rc;
    if ((rc = make_tmap(&tm_c, dt, 2, c, (uint64_t)m, (uint64_t)n, (uint64_t)batch * n * 2, kBlockM, 64, (uint64_t)batch, (uint64_t)n * 2))) return rc;
    const unsigned m_tiles = (unsigned)((m + 2 * kBlockM - 1) / (2 * kBlockM));
    if (n % 128 == 0) {
      static DoneFlag done[kMaxDevices];
      if ((rc = set_smem_attr(bmt_tc_kernel<T, 128>, bmt_smem_bytes<128>(), &done[dev], "bmt<128>"))) return rc;
      e = launch_k(bmt_tc_kernel<T, 128>, dim3((unsigned)(n / 128), m_tiles, (unsigned)batch), dim3(kBmtThreads), bmt_smem_bytes<128>(), st, false,
                   tm_a, tm_b, tm_c, (int)m, (int)k);
    } else {
      static DoneFlag done[kMaxDevices];
      if ((rc = set_smem_attr(bmt_tc_kernel<T, 64>, bmt_smem_bytes<64>(), &done[dev], "bmt<64>"))) return rc;
      e = launch_k(bmt_tc_kernel<T, 64>, dim3((unsigned)(n / 64), m_tiles, (unsigned)batch), dim3(kBmtThreads), bmt_smem_bytes<64>(), st, false,
                   tm_a, tm_b, tm_c, (int)m, (int)k);
    }
  } else {
    long long blocks = (m * batch * n + 255) / 256;
    if (blocks > (long long)num_sms() * 64) blocks = (long long)num_sms() * 64;
    bmt_generic_kernel<T><<<(unsigned)blocks, 256, 0, st>>>(reinterpret_cast<const T*>(a), reinterpret_cast<const T*>(b), reinterpret_cast<T*>(c),
                                                           m, batch, k, n, b_is_nz);
    e = cudaGetLastError();
  }
  return e == cudaSuccess ? MLAPO_OK : fail(MLAPO_ERR_CUDA, "batch_matmul_transpose launch: %s", cudaGetErrorString(e));
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
  const int old = g_debug_flags_word.exchange(flags);
  g_debug_flags = flags;
#ifdef MLAPO_INSTRUMENTED
  if ((flags ^ old) & 2097152) {  // bit21: random idle periods before barrier waits (current device; synchronises it)
    const unsigned jitter = (flags & 2097152) ? 3000u : 0u;
    cudaMemcpyToSymbol(ptx::g_wait_jitter, &jitter, sizeof jitter);
  }
#endif
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
int mlapo_debug_is_instrumented(void) {
#ifdef MLAPO_INSTRUMENTED
  return 1;
#else
  return 0;
#endif
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
  const int dev = g_ev_dev;
  if (!g_ev_ready[dev].load(std::memory_order_acquire) || g_ev_count < 2) return 0;
  if (cudaEventSynchronize(g_ev[dev][g_ev_count - 1]) != cudaSuccess) return fail(MLAPO_ERR_CUDA, "event sync failed");
  int n = g_ev_count - 1;
  if (n > cap) n = cap;
  for (int i = 0; i < n; ++i) cudaEventElapsedTime(&out_ms[i], g_ev[dev][i], g_ev[dev][i + 1]);
  return n;
}
const char* mlapo_kernel_name(int i) { return (i >= 0 && i < kMaxKernels) ? g_kernel_names[i] : ""; }

int mlapo_get_workspace_layout(const MlapoParams* p, MlapoWorkspaceLayout* layout) {
  snapshot_debug_flags();
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
  snapshot_debug_flags();
  if (check_params(p)) return 0;
  return 4;
}

int batch_matmul_transpose_launch(const void* a, const void* b, void* c, int64_t m, int64_t batch, int64_t k, int64_t n,
                                  int32_t act_dtype, int32_t b_format, void* stream) {
  if (!a || !b || !c) return fail(MLAPO_ERR_INVALID_ARG, "batch_matmul_transpose: null pointer");
  if (m <= 0 || batch <= 0 || k <= 0 || n <= 0) return fail(MLAPO_ERR_INVALID_ARG, "batch_matmul_transpose: m=%lld batch=%lld k=%lld n=%lld",
                                                          (long long)m, (long long)batch, (long long)k, (long long)n);
  if (b_format != MLAPO_FORMAT_ND && b_format != MLAPO_FORMAT_NZ) return fail(MLAPO_ERR_INVALID_ARG, "batch_matmul_transpose: format %d", b_format);
  if (b_format == MLAPO_FORMAT_NZ && (k % 16 || n % 16))
    return fail(MLAPO_ERR_INVALID_ARG, "batch_matmul_transpose: NZ weights need k and n to be multiples of 16 (k=%lld n=%lld)", (long long)k, (long long)n);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (act_dtype == MLAPO_ACT_BF16) return bmt_launch_impl<__nv_bfloat16>(a, b, c, m, batch, k, n, b_format == MLAPO_FORMAT_NZ, st);
  if (act_dtype == MLAPO_ACT_FP16) return bmt_launch_impl<__half>(a, b, c, m, batch, k, n, b_format == MLAPO_FORMAT_NZ, st);
  return fail(MLAPO_ERR_INVALID_ARG, "batch_matmul_transpose: tensors must be fp16 or bf16");
}

int mlapo_alloc_extend(const void* pre_lens, const void* seq_lens, const void* last_loc, const void* free_pages, int64_t page_size,
                       void* out_indices, void* values, int32_t batch, int64_t out_capacity, int32_t lens_are_int64,
                       int32_t out_is_int64, void* stream) {
  if (!pre_lens || !seq_lens || !last_loc || !free_pages || !out_indices) return fail(MLAPO_ERR_INVALID_ARG, "alloc_extend: null pointer");
  if (batch <= 0 || page_size <= 0 || out_capacity < 0) return fail(MLAPO_ERR_INVALID_ARG, "alloc_extend: batch=%d page_size=%lld", batch, (long long)page_size);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define MLAPO_AE(I, O)                                                                                                        \
  alloc_extend_kernel<I, O><<<batch, 256, 0, st>>>(reinterpret_cast<const I*>(pre_lens), reinterpret_cast<const I*>(seq_lens), \
                                                   reinterpret_cast<const I*>(last_loc), reinterpret_cast<const I*>(free_pages), \
                                                   page_size, reinterpret_cast<O*>(out_indices), reinterpret_cast<I*>(values), batch, out_capacity)
  if (lens_are_int64) { if (out_is_int64) MLAPO_AE(long long, long long); else MLAPO_AE(long long, int); }
  else { if (out_is_int64) MLAPO_AE(int, long long); else MLAPO_AE(int, int); }
#undef MLAPO_AE
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MLAPO_OK : fail(MLAPO_ERR_CUDA, "alloc_extend launch: %s", cudaGetErrorString(e));
}

int mlapo_cache_loc_assign(const void* req_indices, int32_t req_is_int64, void* token_pool, const void* start_offset, const void* end_offset,
                           void* cache_loc, int32_t batch, int64_t row_size, int32_t update, void* stream) {
  if (!req_indices || !token_pool || !start_offset || !end_offset || !cache_loc) return fail(MLAPO_ERR_INVALID_ARG, "cache_loc_assign: null pointer");
  if (batch <= 0 || row_size <= 0) return fail(MLAPO_ERR_INVALID_ARG, "cache_loc_assign: batch=%d row_size=%lld", batch, (long long)row_size);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (req_is_int64)
    cache_loc_kernel<long long><<<batch, 128, 0, st>>>(reinterpret_cast<const long long*>(req_indices), reinterpret_cast<int32_t*>(token_pool),
                                                       reinterpret_cast<const long long*>(start_offset), reinterpret_cast<const long long*>(end_offset),
                                                       reinterpret_cast<int32_t*>(cache_loc), row_size, update);
  else
    cache_loc_kernel<int><<<batch, 128, 0, st>>>(reinterpret_cast<const int*>(req_indices), reinterpret_cast<int32_t*>(token_pool),
                                                 reinterpret_cast<const long long*>(start_offset), reinterpret_cast<const long long*>(end_offset),
                                                 reinterpret_cast<int32_t*>(cache_loc), row_size, update);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MLAPO_OK : fail(MLAPO_ERR_CUDA, "cache_loc_assign launch: %s", cudaGetErrorString(e));
}

int mlapo_assign_cache(void* out, const void* src, const void* dst_start, const void* dst_end, const void* src_start, const void* src_end,
                       int32_t batch, int64_t row_elems, int32_t elem_bytes, int32_t idx_is_int64, void* stream) {
  if (!out || !src || !dst_start || !dst_end || !src_start || !src_end) return fail(MLAPO_ERR_INVALID_ARG, "assign_cache: null pointer");
  if (batch <= 0 || row_elems <= 0 || elem_bytes <= 0 || elem_bytes > 8) return fail(MLAPO_ERR_INVALID_ARG, "assign_cache: batch=%d row=%lld elem_bytes=%d", batch, (long long)row_elems, elem_bytes);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (idx_is_int64)
    assign_cache_kernel<long long><<<batch, 256, 0, st>>>(reinterpret_cast<uint8_t*>(out), reinterpret_cast<const uint8_t*>(src),
        reinterpret_cast<const long long*>(dst_start), reinterpret_cast<const long long*>(dst_end), reinterpret_cast<const long long*>(src_start),
        reinterpret_cast<const long long*>(src_end), row_elems, elem_bytes);
  else
    assign_cache_kernel<int><<<batch, 256, 0, st>>>(reinterpret_cast<uint8_t*>(out), reinterpret_cast<const uint8_t*>(src),
        reinterpret_cast<const int*>(dst_start), reinterpret_cast<const int*>(dst_end), reinterpret_cast<const int*>(src_start),
        reinterpret_cast<const int*>(src_end), row_elems, elem_bytes);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MLAPO_OK : fail(MLAPO_ERR_CUDA, "assign_cache launch: %s", cudaGetErrorString(e));
}

static unsigned grid_for(long long work);

struct SplitPosCache {
  const void* positions;
  const void* cache;
  long long cache_stride;
  int pos_i64, cache_f32, max_seq, cast_norm;
};

// shared by the split_qkv launch entry points (gate_out != null: gated rows, gemma norm; pc != null: cos / sin from a position cache)
static int split_qkv_impl(const void* qkv, const void* sin, const void* cos, void* q_out, void* k_out, void* v_out, void* gate_out,
                          const void* q_weight, const void* k_weight, const void* q_bias, const void* k_bias, int64_t batch,
                          int32_t q_hidden, int32_t kv_hidden, int32_t head_dim, int32_t rope_dim, float eps, int32_t has_norm,
                          int32_t neox_style, int32_t gemma, int32_t act_dtype, void* stream, const SplitPosCache* pc = nullptr) {
  if (!qkv || !q_out || !k_out || !v_out || (!pc && (!sin || !cos))) return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_rmsnorm_rope: null pointer");
  if (pc && (!pc->cache || pc->max_seq < 1 || pc->cache_stride < rope_dim))
    return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_rmsnorm_rope_pos_cache: positions / cos_sin_cache [max_seq >= 1, >= rope_dim]");
  if (has_norm && (!q_weight || !k_weight)) return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_rmsnorm_rope: eps given but no norm weights");
  if ((q_bias == nullptr) != (k_bias == nullptr)) return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_rmsnorm_rope: q_bias and k_bias go together");
  if (batch <= 0 || head_dim < 2 || head_dim > kSplitMaxHeadDim || (head_dim & (head_dim - 1)) || q_hidden % head_dim || kv_hidden % head_dim ||
      kv_hidden <= 0 || q_hidden <= 0 || rope_dim <= 0 || rope_dim > head_dim || (rope_dim & 1))
    return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_rmsnorm_rope: batch=%lld q_hidden=%d kv_hidden=%d head_dim=%d (power of two <= %d) rope_dim=%d",
                (long long)batch, q_hidden, kv_hidden, head_dim, kSplitMaxHeadDim, rope_dim);
  const int gated = gate_out != nullptr;
  SplitQkvArgs a{qkv, sin, cos, q_out, k_out, v_out, q_weight, k_weight, q_bias, k_bias, batch, q_hidden, kv_hidden, head_dim, rope_dim, eps,
                 has_norm, neox_style, gate_out, gated, gemma, nullptr, nullptr, 0, 0, 0, 0, 0};
  if (pc) {
    a.positions = pc->positions; a.cache = pc->cache; a.cache_stride = pc->cache_stride;
    a.pos_i64 = pc->pos_i64; a.cache_f32 = pc->cache_f32; a.max_seq = pc->max_seq; a.cast_norm = pc->cast_norm && has_norm;
  }
  if (act_dtype != MLAPO_ACT_BF16 && act_dtype != MLAPO_ACT_FP16) return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_rmsnorm_rope: tensors must be fp16 or bf16");
  const bool bf = act_dtype == MLAPO_ACT_BF16;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int vec = head_dim >= 512 ? 16 : 8;
  uintptr_t align = 0;
  for (const void* p : {qkv, sin, cos, (const void*)q_out, (const void*)k_out, (const void*)v_out, (const void*)gate_out}) align |= reinterpret_cast<uintptr_t>(p);
  if (pc) align |= reinterpret_cast<uintptr_t>(pc->cache) | (uintptr_t)(pc->cache_stride * (pc->cache_f32 ? 4 : 2));  // bulk copies of cache rows
  if (head_dim >= 32 && rope_dim % vec == 0 && (align & 15) == 0) {
    // vector kernel; token indices are 32-bit in it, so a (hypothetical) larger batch goes in pieces
    const long long row = (gated ? 2LL : 1LL) * q_hidden + 2LL * kv_hidden;
    const long long tok_per_launch = 1LL << 31;
    const size_t es = 2;
    for (long long t0 = 0; t0 < batch; t0 += tok_per_launch) {
      const long long nb = batch - t0 < tok_per_launch ? batch - t0 : tok_per_launch;
      SplitQkvArgs c = a;
      c.qkv = static_cast<const char*>(qkv) + t0 * row * es;
      if (pc && pc->positions) c.positions = static_cast<const char*>(pc->positions) + t0 * (pc->pos_i64 ? 8 : 4);
      else if (pc) c.cache = static_cast<const char*>(pc->cache) + t0 * pc->cache_stride * (pc->cache_f32 ? 4 : 2);  // row t for token t
      else {
        c.sin = static_cast<const char*>(sin) + t0 * rope_dim * es;
        c.cos = static_cast<const char*>(cos) + t0 * rope_dim * es;
      }
      c.q_out = static_cast<char*>(q_out) + t0 * q_hidden * es;
      c.k_out = static_cast<char*>(k_out) + t0 * kv_hidden * es;
      c.v_out = static_cast<char*>(v_out) + t0 * kv_hidden * es;
      if (gated) c.gate_out = static_cast<char*>(gate_out) + t0 * q_hidden * es;
      c.batch = nb;
      cudaError_t le = cudaSuccess;
#define MLAPO_SPLIT_CASE(HD)                                                                                                       \
  case HD:                                                                                                                         \
    le = bf ? (neox_style ? split_vec_launch<__nv_bfloat16, HD, true>(c, st) : split_vec_launch<__nv_bfloat16, HD, false>(c, st))   \
            : (neox_style ? split_vec_launch<__half, HD, true>(c, st) : split_vec_launch<__half, HD, false>(c, st));               \
    break;
      switch (head_dim) {
        MLAPO_SPLIT_CASE(32) MLAPO_SPLIT_CASE(64) MLAPO_SPLIT_CASE(128) MLAPO_SPLIT_CASE(256) MLAPO_SPLIT_CASE(512)
        default: return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_rmsnorm_rope: head_dim=%d", head_dim);
      }
#undef MLAPO_SPLIT_CASE
      if (le != cudaSuccess) return fail(MLAPO_ERR_CUDA, "split_qkv_rmsnorm_rope launch: %s", cudaGetErrorString(le));
    }
  } else {
    const long long items = batch * (((gated ? 2LL : 1LL) * q_hidden + 2LL * kv_hidden) / head_dim);
    long long blocks = (items + kSplitWarps - 1) / kSplitWarps;
    if (blocks > num_sms() * 32) blocks = num_sms() * 32;
    if (bf) split_qkv_scalar_kernel<__nv_bfloat16><<<(unsigned)blocks, kSplitWarps * 32, 0, st>>>(a);
    else split_qkv_scalar_kernel<__half><<<(unsigned)blocks, kSplitWarps * 32, 0, st>>>(a);
  }
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MLAPO_OK : fail(MLAPO_ERR_CUDA, "split_qkv_rmsnorm_rope launch: %s", cudaGetErrorString(e));
}

int split_qkv_rmsnorm_rope_launch(const void* qkv, const void* sin, const void* cos, void* q_out, void* k_out, void* v_out,
                                  const void* q_weight, const void* k_weight, const void* q_bias, const void* k_bias, int64_t batch,
                                  int32_t q_hidden, int32_t kv_hidden, int32_t head_dim, int32_t rope_dim, float eps, int32_t has_norm,
                                  int32_t neox_style, int32_t act_dtype, void* stream) {
  return split_qkv_impl(qkv, sin, cos, q_out, k_out, v_out, nullptr, q_weight, k_weight, q_bias, k_bias, batch, q_hidden, kv_hidden, head_dim,
                        rope_dim, eps, has_norm, neox_style, 0, act_dtype, stream);
}

int split_qkv_rmsnorm_rope_pos_cache_launch(const void* qkv, const void* positions, int32_t positions_are_int64, const void* cos_sin_cache,
                                            int32_t cache_is_f32, int64_t cache_stride, int32_t max_seq, void* q_out, void* k_out, void* v_out,
                                            const void* q_weight, const void* k_weight, const void* q_bias, const void* k_bias, int64_t batch,
                                            int32_t q_hidden, int32_t kv_hidden, int32_t head_dim, int32_t rope_dim, float eps, int32_t has_norm,
                                            int32_t cast_norm, int32_t act_dtype, void* stream) {
  if (!positions) return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_rmsnorm_rope_pos_cache: null positions");
  const SplitPosCache pc{positions, cos_sin_cache, cache_stride, positions_are_int64, cache_is_f32, max_seq, cast_norm};
  return split_qkv_impl(qkv, nullptr, nullptr, q_out, k_out, v_out, nullptr, q_weight, k_weight, q_bias, k_bias, batch, q_hidden, kv_hidden,
                        head_dim, rope_dim, eps, has_norm, 1, 0, act_dtype, stream, &pc);
}

int split_qkv_rmsnorm_mrope_launch(const void* qkv, const void* cos_sin, int32_t cos_sin_is_f32, void* workspace, void* q_out, void* k_out,
                                   void* v_out, void* gate_out, const void* q_weight, const void* k_weight, const void* q_bias,
                                   const void* k_bias, int64_t batch, int32_t num_q_heads, int32_t num_kv_heads, int32_t head_size,
                                   int32_t rope_dim, float eps, int32_t section_t, int32_t section_h, int32_t section_w,
                                   int32_t is_interleaved, int32_t act_dtype, void* stream) {
  if (!cos_sin || !workspace) return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_rmsnorm_mrope: null cos_sin / workspace");
  if (batch <= 0 || batch > (1LL << 30) || rope_dim <= 0 || (rope_dim & 1) || num_q_heads <= 0 || num_kv_heads <= 0 || head_size <= 0)
    return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_rmsnorm_mrope: batch=%lld heads=%d/%d head_size=%d rope_dim=%d", (long long)batch, num_q_heads,
                num_kv_heads, head_size, rope_dim);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const unsigned blocks = grid_for(batch * rope_dim);
  if (cos_sin_is_f32)
    mrope_select_kernel<float><<<blocks, 256, 0, st>>>(static_cast<const float*>(cos_sin), static_cast<float*>(workspace), batch, rope_dim,
                                                       section_t, section_h, section_w, is_interleaved);
  else
    mrope_select_kernel<uint16_t><<<blocks, 256, 0, st>>>(static_cast<const uint16_t*>(cos_sin), static_cast<uint16_t*>(workspace), batch,
                                                          rope_dim, section_t, section_h, section_w, is_interleaved);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(MLAPO_ERR_CUDA, "split_qkv_rmsnorm_mrope select launch: %s", cudaGetErrorString(e));
  const SplitPosCache pc{nullptr, workspace, rope_dim, 0, cos_sin_is_f32, (int)batch, 0};
  return split_qkv_impl(qkv, nullptr, nullptr, q_out, k_out, v_out, gate_out, q_weight, k_weight, q_bias, k_bias, batch, num_q_heads * head_size,
                        num_kv_heads * head_size, head_size, rope_dim, eps, 1, 1, 0, act_dtype, stream, &pc);
}

static int split_qkv_tp_impl(const void* qkv, const void* cos, const void* sin, void* q_out, void* k_out, void* v_out,
                             const void* q_weight, const void* k_weight, void* qk_var, int64_t batch, int32_t q_hidden,
                             int32_t kv_hidden, int32_t head_dim, int32_t rotary_dim, float eps, int32_t tp_world, int32_t mode,
                             int32_t act_dtype, void* stream, const void* const* peers, int32_t tp_rank, void* epoch_state, int32_t capacity) {
  if (!qkv) return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_tp_rmsnorm_rope: null qkv");
  if (mode != kTpFused && mode != kTpSums && mode != kTpApply && mode != kTpPeer) return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_tp_rmsnorm_rope: mode %d", mode);
  if (mode == kTpPeer && (!peers || tp_world < 2 || tp_world > kTpMaxRanks || tp_rank < 0 || tp_rank >= tp_world || !epoch_state || batch > capacity))
    return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_tp_rmsnorm_rope (peer exchange): tp_rank=%d tp_world=%d (2..%d) epoch_state=%p batch=%lld capacity=%d",
                tp_rank, tp_world, kTpMaxRanks, epoch_state, (long long)batch, capacity);
  if (mode != kTpSums && (!cos || !sin || !q_out || !k_out || !v_out || !q_weight || !k_weight))
    return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_tp_rmsnorm_rope: null pointer");
  if ((mode == kTpSums || mode == kTpApply) && !qk_var) return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_tp_rmsnorm_rope: qk_var is needed for the two-phase modes");
  if (mode == kTpFused && tp_world != 1) return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_tp_rmsnorm_rope: the one-pass mode is for tp_world == 1");
  if (batch <= 0 || batch > (1LL << 31) || tp_world < 1 || (head_dim != 32 && head_dim != 64 && head_dim != 128 && head_dim != 256) ||
      q_hidden <= 0 || kv_hidden <= 0 || q_hidden % head_dim || kv_hidden % head_dim || rotary_dim <= 0 || rotary_dim > head_dim || rotary_dim % 8)
    return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_tp_rmsnorm_rope: batch=%lld q_hidden=%d kv_hidden=%d head_dim=%d (32..256) rotary_dim=%d (multiple of 8)",
                (long long)batch, q_hidden, kv_hidden, head_dim, rotary_dim);
  if (act_dtype != MLAPO_ACT_BF16 && act_dtype != MLAPO_ACT_FP16) return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_tp_rmsnorm_rope: tensors must be fp16 or bf16");
  uintptr_t align = 0;
  for (const void* p : {qkv, cos, sin, (const void*)q_out, (const void*)k_out, (const void*)v_out, q_weight, k_weight}) align |= reinterpret_cast<uintptr_t>(p);
  if (align & 15) return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_tp_rmsnorm_rope: pointers must be 16-byte aligned");
  const int row = q_hidden + 2 * kv_hidden;
  int stages = 4;
  while (stages > 3 && tp_smem_bytes(row, rotary_dim, q_hidden, kv_hidden, stages) > 72 * 1024) --stages;   // three CTAs per SM if a 3-deep ring allows
  while (stages > 1 && tp_smem_bytes(row, rotary_dim, q_hidden, kv_hidden, stages) > 72 * 1024 &&
         tp_smem_bytes(row, rotary_dim, q_hidden, kv_hidden, stages) > 110 * 1024) --stages;                  // else two
  if (stages < 2) {
    stages = 2;
    if (tp_smem_bytes(row, rotary_dim, q_hidden, kv_hidden, stages) > 200 * 1024)
      return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_tp_rmsnorm_rope: a row of %d elements does not fit shared memory twice", row);
  }
  SplitTpArgs a{qkv, cos, sin, q_out, k_out, v_out, q_weight, k_weight, static_cast<float*>(qk_var), batch, q_hidden, kv_hidden, head_dim,
                rotary_dim, eps, 1.f / (float)tp_world, mode, stages, {}, tp_rank, tp_world, capacity, static_cast<unsigned*>(epoch_state)};
  if (mode == kTpPeer)
    for (int r = 0; r < tp_world; ++r) {
      if (!peers[r]) return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_tp_rmsnorm_rope (peer exchange): null buffer of rank %d", r);
      a.peer[r] = reinterpret_cast<unsigned long long>(peers[r]);
    }
  const int smem = tp_smem_bytes(row, rotary_dim, q_hidden, kv_hidden, stages);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool bf = act_dtype == MLAPO_ACT_BF16;
  // The grid is what the device holds at once (asked of the runtime, not assumed): the peer-exchange mode needs every CTA resident,
  // since a CTA waits for records that the same CTA index publishes on the other ranks.  The shared-memory opt-in and the occupancy
  // answer are cached per (dtype, shared-memory size, device): this is a decode-path operator and both are host round trips.
  static std::mutex tp_mutex;
  static std::map<std::tuple<int, int, int>, int> tp_per_sm;
  cudaError_t e = cudaSuccess;
  int per_sm = 0;
  const int sms = num_sms();
  {
    std::lock_guard<std::mutex> lock(tp_mutex);
    const auto key = std::make_tuple(bf ? 1 : 0, smem, current_device());
    auto it = tp_per_sm.find(key);
    if (it != tp_per_sm.end()) {
      per_sm = it->second;
    } else {
      // the attribute is a high-water mark per function and device: never lower it below an earlier, larger request
      int prev = 0;
      for (const auto& kv : tp_per_sm)
        if (std::get<0>(kv.first) == (bf ? 1 : 0) && std::get<2>(kv.first) == current_device() && std::get<1>(kv.first) > prev) prev = std::get<1>(kv.first);
      if (smem > prev)
        e = bf ? cudaFuncSetAttribute(split_qkv_tp_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)
               : cudaFuncSetAttribute(split_qkv_tp_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
      if (e == cudaSuccess)
        e = bf ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, split_qkv_tp_kernel<__nv_bfloat16>, kTpThreads, smem)
               : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, split_qkv_tp_kernel<__half>, kTpThreads, smem);
      if (e == cudaSuccess) tp_per_sm.emplace(key, per_sm);
    }
  }
  if (e == cudaSuccess && (per_sm < 1 || sms < 1)) return fail(MLAPO_ERR_CUDA, "split_qkv_tp_rmsnorm_rope: the kernel does not fit an SM (%d bytes of shared memory)", smem);
  if (e == cudaSuccess) {
    const unsigned resident = (unsigned)per_sm * (unsigned)sms;
    const unsigned blocks = batch < resident ? (unsigned)batch : resident;
    if (bf) split_qkv_tp_kernel<__nv_bfloat16><<<blocks, kTpThreads, smem, st>>>(a);
    else split_qkv_tp_kernel<__half><<<blocks, kTpThreads, smem, st>>>(a);
    e = cudaGetLastError();
  }
  return e == cudaSuccess ? MLAPO_OK : fail(MLAPO_ERR_CUDA, "split_qkv_tp_rmsnorm_rope launch: %s", cudaGetErrorString(e));
}

int split_qkv_tp_rmsnorm_rope_launch(const void* qkv, const void* cos, const void* sin, void* q_out, void* k_out, void* v_out,
                                     const void* q_weight, const void* k_weight, void* qk_var, int64_t batch, int32_t q_hidden,
                                     int32_t kv_hidden, int32_t head_dim, int32_t rotary_dim, float eps, int32_t tp_world, int32_t mode,
                                     int32_t act_dtype, void* stream) {
  if (mode == kTpPeer) return fail(MLAPO_ERR_INVALID_ARG, "split_qkv_tp_rmsnorm_rope: the peer-exchange mode has its own entry point");
  return split_qkv_tp_impl(qkv, cos, sin, q_out, k_out, v_out, q_weight, k_weight, qk_var, batch, q_hidden, kv_hidden, head_dim, rotary_dim, eps,
                           tp_world, mode, act_dtype, stream, nullptr, 0, nullptr, 0);
}

int split_qkv_tp_rmsnorm_rope_peer_launch(const void* qkv, const void* cos, const void* sin, void* q_out, void* k_out, void* v_out,
                                          const void* q_weight, const void* k_weight, int64_t batch, int32_t q_hidden, int32_t kv_hidden,
                                          int32_t head_dim, int32_t rotary_dim, float eps, const void* const* peer_buffers, int32_t tp_rank,
                                          int32_t tp_world, void* epoch_state, int32_t capacity, int32_t act_dtype, void* stream) {
  return split_qkv_tp_impl(qkv, cos, sin, q_out, k_out, v_out, q_weight, k_weight, nullptr, batch, q_hidden, kv_hidden, head_dim, rotary_dim, eps,
                           tp_world, kTpPeer, act_dtype, stream, peer_buffers, tp_rank, epoch_state, capacity);
}

int split_qkvgate_gemma_rmsnorm_rope_launch(const void* qkvgate, const void* sin, const void* cos, void* q_out, void* k_out, void* v_out,
                                            void* gate_out, const void* q_weight, const void* k_weight, int64_t batch, int32_t q_hidden,
                                            int32_t kv_hidden, int32_t head_dim, int32_t rope_dim, float eps, int32_t act_dtype, void* stream) {
  if (!gate_out) return fail(MLAPO_ERR_INVALID_ARG, "split_qkvgate_gemma_rmsnorm_rope: null gate_out");
  return split_qkv_impl(qkvgate, sin, cos, q_out, k_out, v_out, gate_out, q_weight, k_weight, nullptr, nullptr, batch, q_hidden, kv_hidden,
                        head_dim, rope_dim, eps, 1, 1, 1, act_dtype, stream);
}

static unsigned grid_for(long long work) {
  long long b = (work + 255) / 256;
  return (unsigned)(b < 1 ? 1 : b > 65535 ? 65535 : b);
}

int mlapo_transdata_nd_to_nz(const void* nd, void* nz, int64_t rows, int64_t cols, int32_t c0, int32_t elem_bytes, void* stream) {
  if (!nd || !nz) return fail(MLAPO_ERR_INVALID_ARG, "transdata: null pointer");
  if (rows <= 0 || cols <= 0 || c0 <= 0 || (elem_bytes != 1 && elem_bytes != 2 && elem_bytes != 4))
    return fail(MLAPO_ERR_INVALID_ARG, "transdata: rows=%lld cols=%lld c0=%d elem_bytes=%d", (long long)rows, (long long)cols, c0, elem_bytes);
  const long long rp = (rows + 15) / 16 * 16, cp = (cols + c0 - 1) / c0 * c0;
  transdata_nd_to_nz_kernel<<<grid_for(rp * cp), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const uint8_t*>(nd), reinterpret_cast<uint8_t*>(nz), rows, cols, rp, cp, c0, elem_bytes);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MLAPO_OK : fail(MLAPO_ERR_CUDA, "transdata launch: %s", cudaGetErrorString(e));
}

int mlapo_pack_descale_int64(const void* descale_f32, void* descale_i64, int64_t n, void* stream) {
  if (!descale_f32 || !descale_i64 || n <= 0) return fail(MLAPO_ERR_INVALID_ARG, "pack_descale: bad argument");
  pack_descale_kernel<<<grid_for(n), 256, 0, static_cast<cudaStream_t>(stream)>>>(reinterpret_cast<const uint32_t*>(descale_f32),
                                                                                 reinterpret_cast<unsigned long long*>(descale_i64), n);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MLAPO_OK : fail(MLAPO_ERR_CUDA, "pack_descale launch: %s", cudaGetErrorString(e));
}

int mlapo_gather_nz_cache_rows(const void* cache, const int32_t* slots, void* rows_out, int64_t n, int32_t width, int32_t c0,
                               int32_t elem_bytes, int64_t num_blocks, void* stream) {
  if (!cache || !slots || !rows_out) return fail(MLAPO_ERR_INVALID_ARG, "gather_nz_cache_rows: null pointer");
  if (n <= 0 || width <= 0 || c0 <= 0 || width % c0 || (elem_bytes != 1 && elem_bytes != 2) || num_blocks <= 0)
    return fail(MLAPO_ERR_INVALID_ARG, "gather_nz_cache_rows: n=%lld width=%d c0=%d elem_bytes=%d blocks=%lld", (long long)n, width, c0,
                elem_bytes, (long long)num_blocks);
  gather_nz_rows_kernel<<<grid_for(n * width), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const uint8_t*>(cache), slots, reinterpret_cast<uint8_t*>(rows_out), n, width, c0, elem_bytes, num_blocks);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? MLAPO_OK : fail(MLAPO_ERR_CUDA, "gather_nz_cache_rows launch: %s", cudaGetErrorString(e));
}

int mlapo_repack_nz_weight(const void* w_nz, void* w_km, int32_t k, int32_t n_out, void* stream) {
  snapshot_debug_flags();
  if (!w_nz || !w_km) return fail(MLAPO_ERR_INVALID_ARG, "repack: null pointer");
  if (k <= 0 || k % 32 != 0 || n_out <= 0) return fail(MLAPO_ERR_INVALID_ARG, "repack: k must be a multiple of 32");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long chunks = (long long)n_out * (k / 32) * 2;
  int blocks = (int)((chunks + 255) / 256);
  if (blocks > num_sms() * 16) blocks = num_sms() * 16;
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
  snapshot_debug_flags();
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
  snapshot_debug_flags();
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
