/*
 * mla_preprocess_b200.h — C ABI of the B200 (sm_100a) implementation of sgl-kernel-npu's
 * `mla_preprocess` operator (DeepSeek-V3 MLA attention prologue).
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch types.  The torch operator
 * `torch.ops.npu.mla_preprocess` (csrc/mlapo_torch_op.cpp) and any other host (ctypes, cgo, JNI ...)
 * bind exactly these symbols.  All pointers are DEVICE pointers unless stated; all calls are
 * asynchronous on `stream` (a cudaStream_t passed as void*), never synchronise the host and never
 * allocate device memory: the caller owns every buffer, including the workspace.
 *
 * Reference interface replaced (paths relative to the sgl-kernel-npu tree):
 *   - operator entry      sglang::npu_kernel::mla_preprocess   csrc/mla_preprocess/op_host/mla_preprocess.cpp:623-704
 *   - operator schema     csrc/pytorch_extensions.cpp:105-115, include/sgl_kenel_npu_ops.h:102-119
 *   - device entry        extern "C" mla_preprocess(31 GM_ADDR) csrc/mla_preprocess/op_kernel/mla_preprocess_kernel.cpp:20-25
 *   - workspace sizing    MlaPreprocessTiling::SetWorkspace     csrc/mla_preprocess/op_host/mla_preprocess.cpp:508-540
 */
#ifndef MLA_PREPROCESS_B200_H_
#define MLA_PREPROCESS_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Return codes (0 = success, negative = failure; mlapo_last_error() has the text). */
#define MLAPO_OK 0
#define MLAPO_ERR_INVALID_ARG (-1)   /* bad pointer / size / mode value                                  */
#define MLAPO_ERR_UNSUPPORTED (-2)   /* combination the path does not implement (e.g. fp16 + NZ cache)    */
#define MLAPO_ERR_WORKSPACE (-3)     /* workspace missing, misaligned or too small                        */
#define MLAPO_ERR_CUDA (-4)          /* a CUDA runtime call or launch failed                              */
#define MLAPO_ERR_DRIVER (-5)        /* cuTensorMapEncodeTiled unavailable or failed                      */

/* Activation type of hiddenState / outputs (reference: inDtype, op_host/mla_preprocess.cpp:666). */
#define MLAPO_ACT_BF16 0
#define MLAPO_ACT_FP16 1

/* cache_mode values = reference cache_mode_map (op_host/mla_preprocess.cpp:605-606). */
#define MLAPO_CACHE_KROPE_CTKV 1     /* "krope_ctkv":   ND caches [B,bs,1,512] + [B,bs,1,64]               */
#define MLAPO_CACHE_INT8_NZ 2        /* "int8_nzcache": NZ caches, int8 k_nope (C0=32) + int8 q_out0       */
#define MLAPO_CACHE_NZ 3             /* "nzcache":      NZ caches (C0=16)                                  */

/* quant_mode values = reference quant_mode_map (op_host/mla_preprocess.cpp:608-611). */
#define MLAPO_QUANT_PER_TENSOR_ASYMM 0
#define MLAPO_QUANT_PER_TOKEN_SYMM 1

/*
 * MlapoParams.flags.
 * MLAPO_FLAG_INPUT_COMPLETE: the caller states that everything the call reads (hidden_state above all) and everything it
 * writes (outputs, caches, workspace) is free of pending work by the time the PREVIOUS kernel on `stream` lets its
 * dependents start.  That holds when the previous kernel on the stream is one of this library's own mla_preprocess
 * kernels, or any kernel that does not use programmatic dependent launch (griddepcontrol.launch_dependents /
 * cudaTriggerProgrammaticLaunchCompletion) — torch's own kernels, cuBLAS, NCCL today.  It does NOT hold behind a
 * producer that triggers its dependents early, before its last store (CUTLASS / TensorRT-LLM / flashinfer kernels with
 * PDL enabled).  With the flag set the first kernel of the call (the quantiser) reads hidden_state and writes the
 * workspace without waiting for the previous kernel to complete, so that it runs under that kernel's tail (about 2.5 us
 * per call at 256 tokens x 128 heads when calls follow one another); without it (the default) every CTA of that kernel
 * waits for the previous kernel's completion first, which is always safe.
 * One more condition when the previous kernel IS an earlier mla_preprocess call on the same workspace memory: that call's q
 * kernel lets its dependents start while it still reads its a1 rows from the workspace, and this call's quantiser then writes
 * the first num_tokens x hidden_dim bytes of it.  With the same num_tokens / hidden_dim / quant_mode (same workspace layout)
 * those bytes are the previous call's a0, long consumed; a different layout may put them under its a1.  So: set the flag
 * between calls of one shape on one workspace (or on disjoint workspaces), and clear it for the first call after a change of
 * shape.  The torch operator applies exactly this rule to the workspace it allocates (mlapo_torch_op.cpp).
 */
#define MLAPO_FLAG_INPUT_COMPLETE 1

/* Fixed dimensions of the path (op_kernel/mla_preprocess.h:55-64). */
#define MLAPO_MM1_OUT 2112           /* [k_nope 512 | k_pe 64 | q_a 1536] */
#define MLAPO_KV_LORA 512
#define MLAPO_ROPE 64
#define MLAPO_Q_LORA 1536
#define MLAPO_Q_NOPE 128
#define MLAPO_Q_HEAD 192

/*
 * Arguments of one call.  Field names follow the operator schema (csrc/pytorch_extensions.cpp:105-115).
 * gamma0/beta0 of the schema are not here: the reference binds them and never reads them
 * (mla_preprocess_mix_bf16.hpp:2762, 424-486 — step 1 is quantisation only).
 */
typedef struct MlapoParams {
  uint32_t struct_bytes;        /* = sizeof(MlapoParams), ABI check                                       */
  int32_t num_tokens;           /* N  = hiddenState.size(0)        (op_host/mla_preprocess.cpp:658)       */
  int32_t num_heads;            /* H  = wuk.size(0)                (:659)                                 */
  int32_t hidden_dim;           /* D  = hiddenState.size(-1), multiple of 128 (:660)                      */
  int32_t act_dtype;            /* MLAPO_ACT_*                                                            */
  int32_t cache_mode;           /* MLAPO_CACHE_*                                                          */
  int32_t quant_mode;           /* MLAPO_QUANT_*                                                          */
  int32_t num_cache_blocks;     /* B of the cache tensors                                                 */
  int32_t cache_block_size;     /* bs; must be 128 in the NZ modes (mla_preprocess_mix_bf16.hpp:2549-2550)*/
  int32_t descale_is_int64;     /* 1: descale0/1 are int64 carrying fp32 bits in the low word (fp16 path) */
  int32_t flags;                /* MLAPO_FLAG_* (0 = none)                                                */

  const void* hidden_state;     /* act  [N, D]                                                            */
  const void* wdqkv_km;         /* int8 [2112, D]   K-major, from mlapo_repack_nz_weight()                */
  const void* descale0;         /* f32|int64 [2112]                                                       */
  const void* bias0;            /* int32 [2112]     (ignored in per-token mode)                           */
  const void* quant_scale0;     /* act  [1]         (per-tensor mode)                                     */
  const void* quant_offset0;    /* int8 [1]                                                               */
  const void* gamma1;           /* act  [1536]                                                            */
  const void* beta1;            /* act  [1536]                                                            */
  const void* quant_scale1;     /* act  [1]                                                               */
  const void* quant_offset1;    /* int8 [1]                                                               */
  const void* wuq_km;           /* int8 [192*H, 1536] K-major, from mlapo_repack_nz_weight()              */
  const void* descale1;         /* f32|int64 [192*H]                                                      */
  const void* bias1;            /* int32 [192*H]                                                          */
  const void* gamma2;           /* act  [512]                                                             */
  const void* cos;              /* act  [N, 64]                                                           */
  const void* sin;              /* act  [N, 64]                                                           */
  const void* wuk;              /* act  [H, 128, 512] ND                                                  */
  const void* slot_mapping;     /* int32 [N]; -1 = skip the kv side of that token                         */
  const void* ctkv_scale;       /* act  [1]   required in MLAPO_CACHE_INT8_NZ, else may be NULL           */
  const void* q_nope_scale;     /* act  [H]   required in MLAPO_CACHE_INT8_NZ, else may be NULL           */

  void* q_out0;                 /* act|int8 [N, H, 512]                                                   */
  void* q_out1;                 /* act  [N, H, 64]                                                        */
  void* kv_cache;               /* = kv_cache_out0 (same storage, updated in place)                       */
  void* kv_cache_rope;          /* = kv_cache_out1                                                        */
} MlapoParams;

/* Byte offsets of the intermediates inside the workspace (for tests / debugging; 0-based, 256 B aligned). */
typedef struct MlapoWorkspaceLayout {
  size_t total_bytes;
  size_t a0;                    /* int8  [N, D]            quantised hiddenState (reference s1)           */
  size_t partial0;              /* int32 [splits0, N, 2112] split-K partial sums of MM1 (reference s2)    */
  size_t a1;                    /* int8  [N, 1536]         quantised q_a (reference s1 reuse)             */
  size_t tok_scale;             /* f32   [2, N]            per-token scales (reference s5)                */
  size_t y0_tap;                /* act   [N, 2112]         dequantised MM1 rows (reference s3); only with debug bit1 */
  size_t timeline;              /* int64 [2, 4096, 16]     per-CTA clock marks of mm1_splitk / q_fused; only with debug bit3 */
  int32_t splits0;              /* split-K factor chosen for MM1                                          */
  int32_t reserved;
} MlapoWorkspaceLayout;

/* Library identification; replaces sgl_kernel_npu_version() (csrc/pytorch_extensions.cpp:26-27, csrc/utils/version.h:23-45). */
const char* mlapo_version(void);

/* Text of the last failure on the calling thread ("" if none). No reference counterpart (TORCH_CHECK messages there). */
const char* mlapo_last_error(void);

/*
 * Workspace the call needs (bytes) / where each intermediate lives in it.
 * Replaces MlaPreprocessTiling::SetWorkspace + the at::empty workspace (op_host/mla_preprocess.cpp:508-540, 677-680).
 */
int mlapo_get_workspace_size(const MlapoParams* p, size_t* bytes);
int mlapo_get_workspace_layout(const MlapoParams* p, MlapoWorkspaceLayout* layout);

/*
 * Load-time repack of an Ascend FRACTAL_NZ int8 weight [1, K/32, n_out, 32] into the K-major
 * [n_out, K] layout the TMA descriptors of this library tile with 128-byte swizzle.
 * Replaces the NZ addressing done inside the reference GEMM (GetOffsetB, mla_preprocess_mix_bf16.hpp:1581-1588);
 * the test-side packer it inverts is transdata() (tests/python/sgl_kernel_npu/test_mla_preprocess.py:38-54).
 */
int mlapo_repack_nz_weight(const void* w_nz, void* w_km, int32_t k, int32_t n_out, void* stream);

/*
 * batch_matmul_transpose: C[m, batch, n] = sum_k A[m, batch, k] * B[batch, k, n], the W_UV half of MLA weight
 * absorption.  Replaces sglang::npu_kernel::batch_matmul_transpose (csrc/batch_matmul_transpose/op_host/batch_matmul_transpose.cpp:46-114,
 * declared include/sgl_kenel_npu_ops.h:121-124, schema csrc/pytorch_extensions.cpp:117-119).  a [m, batch, k], c [m, batch, n],
 * b ND [batch, k, n] or NZ [batch, n/16, k, 16] (k, n multiples of 16); act_dtype MLAPO_ACT_*; fp32 accumulation, one rounding.
 * No limit on m (the reference stops at 1024).  ND weights with k % 64 == 0 and n % 64 == 0 run on the tensor cores.
 */
#define MLAPO_FORMAT_ND 0
#define MLAPO_FORMAT_NZ 1
int batch_matmul_transpose_launch(const void* a, const void* b, void* c, int64_t m, int64_t batch, int64_t k, int64_t n,
                                  int32_t act_dtype, int32_t b_format, void* stream);

/*
 * split_qkv_rmsnorm_rope: split a fused qkv row [q_hidden | kv_hidden | kv_hidden], RMSNorm every q and k head
 * (weight, optional bias; skipped when has_norm = 0), rotate the first rope_dim dims (neox_style 1: halves, 0: interleaved pairs),
 * copy v.  Replaces the Triton kernel behind sgl_kernel_npu.norm.split_qkv_rmsnorm_rope.split_qkv_rmsnorm_rope
 * (python/sgl_kernel_npu/sgl_kernel_npu/norm/split_qkv_rmsnorm_rope.py:374-439).  fp32 math, one rounding at the store.
 * qkv act [batch, q_hidden + 2 kv_hidden], sin/cos act [batch, rope_dim], weights/biases act [head_dim]; head_dim a power of two.
 */
int split_qkv_rmsnorm_rope_launch(const void* qkv, const void* sin, const void* cos, void* q_out, void* k_out, void* v_out,
                                  const void* q_weight, const void* k_weight, const void* q_bias, const void* k_bias, int64_t batch,
                                  int32_t q_hidden, int32_t kv_hidden, int32_t head_dim, int32_t rope_dim, float eps, int32_t has_norm,
                                  int32_t neox_style, int32_t act_dtype, void* stream);

/*
 * split_qkv_rmsnorm_rope_pos_cache_half_npu (python/sgl_kernel_npu/sgl_kernel_npu/norm/split_qkv_rmsnorm_rope_pos_cache_half_npu.py:232-407):
 * as split_qkv_rmsnorm_rope, neox style, but cos / sin are row clamp(positions[t], 0, max_seq - 1) of cos_sin_cache
 * [max_seq, cache_stride >= rope_dim] laid out [cos(rope_dim/2) | sin(rope_dim/2)], tensor dtype or float32; cast_norm = 1 rounds the
 * normalised value to the tensor dtype before the rotation (the reference's cast_norm_to_bf16, its default).
 */
int split_qkv_rmsnorm_rope_pos_cache_launch(const void* qkv, const void* positions, int32_t positions_are_int64, const void* cos_sin_cache,
                                            int32_t cache_is_f32, int64_t cache_stride, int32_t max_seq, void* q_out, void* k_out, void* v_out,
                                            const void* q_weight, const void* k_weight, const void* q_bias, const void* k_bias, int64_t batch,
                                            int32_t q_hidden, int32_t kv_hidden, int32_t head_dim, int32_t rope_dim, float eps, int32_t has_norm,
                                            int32_t cast_norm, int32_t act_dtype, void* stream);

/*
 * split_qkv_rmsnorm_mrope (python/sgl_kernel_npu/sgl_kernel_npu/norm/split_qkv_rmsnorm_mrope.py:335-419, function
 * triton_split_qkv_rmsnorm_mrope): as split_qkv_rmsnorm_rope with weight (+ bias), neox rotation, optional gate heads (gate_out
 * non-null: rows [q_0 | gate_0 | ... | k | v]); cos_sin is [3 (t, h, w), batch, rope_dim] with rows [cos | sin], tensor dtype or
 * float32, and pair i takes the t, h or w row by mrope section (contiguous sections, or interleaved: i % 3 == 1 and i <= 3 h -> h,
 * i % 3 == 2 and i <= 3 w -> w, else t).  workspace: [batch, rope_dim] elements of cos_sin's type, device memory.
 */
int split_qkv_rmsnorm_mrope_launch(const void* qkv, const void* cos_sin, int32_t cos_sin_is_f32, void* workspace, void* q_out, void* k_out,
                                   void* v_out, void* gate_out, const void* q_weight, const void* k_weight, const void* q_bias,
                                   const void* k_bias, int64_t batch, int32_t num_q_heads, int32_t num_kv_heads, int32_t head_size,
                                   int32_t rope_dim, float eps, int32_t section_t, int32_t section_h, int32_t section_w,
                                   int32_t is_interleaved, int32_t act_dtype, void* stream);

/*
 * split_qkv_tp_rmsnorm_rope (python/sgl_kernel_npu/sgl_kernel_npu/norm/split_qkv_tp_rmsnorm_rope.py:172-288): the RMSNorm spans the
 * whole q (k) row, whose columns are sharded over tp_world ranks: var = sum over ranks of mean_local(x^2), divided by tp_world;
 * y = act_rne(x / sqrt(var + eps) * weight[col]); the first rotary_dim dims of every head rotated neox style with cos[t, i],
 * sin[t, i], i < rotary_dim / 2; v copied.  mode 0: one pass, tp_world == 1.  mode 1: write the local means to qk_var f32
 * [batch, 2]; the caller all-reduces (sums) qk_var over the tensor-parallel group (the reference's dist.all_reduce, :253-254);
 * mode 2: one-pass apply with the reduced qk_var.  q_weight act [q_hidden], k_weight act [kv_hidden]; head_dim 32..256,
 * rotary_dim a multiple of 8, pointers 16-byte aligned.
 */
int split_qkv_tp_rmsnorm_rope_launch(const void* qkv, const void* cos, const void* sin, void* q_out, void* k_out, void* v_out,
                                     const void* q_weight, const void* k_weight, void* qk_var, int64_t batch, int32_t q_hidden,
                                     int32_t kv_hidden, int32_t head_dim, int32_t rotary_dim, float eps, int32_t tp_world, int32_t mode,
                                     int32_t act_dtype, void* stream);

/*
 * The same operator for tp_world > 1 as ONE kernel: the ranks exchange the per-token means inside the kernel, through buffers in
 * each other's memory (NVLink peer stores of 8-byte {value, epoch} records), instead of an all-reduce between two launches.  No
 * reference counterpart (the reference calls dist.all_reduce, split_qkv_tp_rmsnorm_rope.py:253-254).  peer_buffers: host array of
 * tp_world device pointers, entry r = rank r's exchange buffer as mapped in THIS process (entry tp_rank is this GPU's own), each
 * 2 * tp_world * capacity * 16 bytes, zeroed once before the first call.  epoch_state: two uint32 in THIS GPU's memory, set to
 * {1, 0} once on every rank; the kernel reads the call's epoch there and its last CTA advances it (its parity picks the buffer
 * half), so replays of a captured CUDA graph work like eager calls.  Every rank must make the same sequence of calls with the
 * same batch <= capacity; a peer that never arrives becomes a trap (launch failure) after a bounded wait.
 */
int split_qkv_tp_rmsnorm_rope_peer_launch(const void* qkv, const void* cos, const void* sin, void* q_out, void* k_out, void* v_out,
                                          const void* q_weight, const void* k_weight, int64_t batch, int32_t q_hidden, int32_t kv_hidden,
                                          int32_t head_dim, int32_t rotary_dim, float eps, const void* const* peer_buffers, int32_t tp_rank,
                                          int32_t tp_world, void* epoch_state, int32_t capacity, int32_t act_dtype, void* stream);

/*
 * split_qkvgate_gemma_rmsnorm_rope (same reference module, split_qkv_rmsnorm_rope.py:441-747): the fused row is
 * [q_0 | gate_0 | q_1 | gate_1 ... | k | v] (every q head followed by a gate head of head_dim), q and k heads are normalised with
 * (1 + weight) and no bias, the first rope_dim dims rotated neox style, v and the gate heads copied.
 * qkvgate act [batch, 2 q_hidden + 2 kv_hidden], gate_out act [batch, q_hidden]; the rest as above.
 */
int split_qkvgate_gemma_rmsnorm_rope_launch(const void* qkvgate, const void* sin, const void* cos, void* q_out, void* k_out, void* v_out,
                                            void* gate_out, const void* q_weight, const void* k_weight, int64_t batch, int32_t q_hidden,
                                            int32_t kv_hidden, int32_t head_dim, int32_t rope_dim, float eps, int32_t act_dtype, void* stream);

/*
 * Slot producers: the ops that create and book-keep the slot_mapping this operator scatters by.  Integer work,
 * bit-exact against the reference tests' goldens; asynchronous on `stream`, device pointers, no scratch memory.
 *  - mlapo_alloc_extend: paged allocation of the new tokens of `batch` requests.  Replaces sglang::npu_kernel::alloc_extend
 *    (schema csrc/pytorch_extensions.cpp:31-33; golden tests/python/sgl_kernel_npu/test_alloc_extend_slot.py:9-85).
 *    pre_lens/seq_lens/last_loc/free_pages int32 or int64 (lens_are_int64), out_indices [sum(seq - pre)] int32 or int64,
 *    values[0] <- number of pages taken (may be null); out_capacity = elements out_indices can hold.
 *  - mlapo_cache_loc_assign: token_pool[req[i], start_i : end_i] <- cache_loc[packed run i] (update = 0, reference
 *    cache_loc_assign) or the reverse copy (update = 1, cache_loc_update); csrc/cache_location_assign/op_host/cache_loc_assign.cpp:88-114.
 *    req int32|int64, token_pool int32 [pool, row_size], start/end int64, cache_loc int32.
 *  - mlapo_assign_cache: out[i, dst_start_i : dst_end_i] <- src[src_start_i : src_end_i], any element size 1..8 bytes, index
 *    tensors int32 or int64; csrc/assign_cache_op/op_host/assign_cache.cpp:40-82.
 */
int mlapo_alloc_extend(const void* pre_lens, const void* seq_lens, const void* last_loc, const void* free_pages, int64_t page_size,
                       void* out_indices, void* values, int32_t batch, int64_t out_capacity, int32_t lens_are_int64,
                       int32_t out_is_int64, void* stream);
int mlapo_cache_loc_assign(const void* req_indices, int32_t req_is_int64, void* token_pool, const void* start_offset, const void* end_offset,
                           void* cache_loc, int32_t batch, int64_t row_size, int32_t update, void* stream);
int mlapo_assign_cache(void* out, const void* src, const void* dst_start, const void* dst_end, const void* src_start, const void* src_end,
                       int32_t batch, int64_t row_elems, int32_t elem_bytes, int32_t idx_is_int64, void* stream);

/*
 * Format tools.  The reference has them only as test helpers; they are exported so that Ascend-format
 * checkpoints and caches can be prepared and read on the device.  All asynchronous on `stream`, device pointers.
 *  - mlapo_transdata_nd_to_nz: ND [rows, cols] -> FRACTAL_NZ [1, ceil(cols/c0), ceil16(rows), c0], zero padded; int8 weights
 *    use c0 = 32, 16-bit tensors c0 = 16.  Replaces transdata() (tests/python/sgl_kernel_npu/test_mla_preprocess.py:38-54).
 *  - mlapo_pack_descale_int64: fp32 descale -> the int64 carrier of the fp16 path, (bits & 0xFFFFE000) | 1 << 46.
 *    Replaces trans_descale_param() (test_mla_preprocess.py:110-119).
 *  - mlapo_gather_nz_cache_rows: rows named by `slots` out of a paged NZ cache [blocks, width/c0, 128, c0] -> dense [n, width]
 *    (slot -1 or out of range -> zeros).  Replaces extract_from_nzcache() (test_mla_preprocess.py:122-135).
 */
int mlapo_transdata_nd_to_nz(const void* nd, void* nz, int64_t rows, int64_t cols, int32_t c0, int32_t elem_bytes, void* stream);
int mlapo_pack_descale_int64(const void* descale_f32, void* descale_i64, int64_t n, void* stream);
int mlapo_gather_nz_cache_rows(const void* cache, const int32_t* slots, void* rows_out, int64_t n, int32_t width, int32_t c0,
                               int32_t elem_bytes, int64_t num_blocks, void* stream);

/*
 * The operator.  Enqueues the whole prologue on `stream`:
 *   quant -> W_DQKV int8 GEMM -> dequant -> { q_a RMSNorm+quant -> W_UQ int8 GEMM -> dequant -> RoPE | W_UK absorb }
 *                                        -> { latent RMSNorm(512), RoPE(64) -> paged-cache scatter by slot_mapping }.
 * Replaces sglang::npu_kernel::mla_preprocess (op_host/mla_preprocess.cpp:623-704) and the device entry
 * mla_preprocess (op_kernel/mla_preprocess_kernel.cpp:20-25).  `workspace` must be 256-byte aligned and at
 * least mlapo_get_workspace_size() bytes.  num_tokens == 0 is a no-op.  There is no token cap
 * (the reference's is 1024: op_host/mla_preprocess.cpp:63, 689-695).
 */
int mla_preprocess_launch(const MlapoParams* p, void* workspace, size_t workspace_bytes, void* stream);

/*
 * The same operator for hosts that cannot hand over a workspace.  (No counterpart in the reference: its host always
 * allocates a fresh workspace tensor per call, op_host/mla_preprocess.cpp:677-680.)
 *
 * mla_preprocess_launch() leaves the workspace to you, which normally means a cudaMalloc or a framework allocator call.
 * Some integrations have neither at the point where they run this operator: a serving engine that claimed the whole
 * device for its own KV-cache pool at start-up and has no bytes left to ask the driver for; a request loop that must not
 * risk the device-wide stall a driver allocation can cause; a plugin interface that only passes inputs and outputs and
 * offers no scratch argument or allocator handle.  For them the library carries its own scratch: one fixed array of
 * MLAPO_STATIC_SCRATCH_BYTES in device global memory, part of the loaded module, present on every GPU that loads this
 * library from load until the process exits.  This entry point uses that array as the workspace and otherwise behaves
 * exactly like mla_preprocess_launch(): same parameters, same kernels, same results, asynchronous on `stream`.  It never
 * calls cudaMalloc/cudaFree or any allocating routine.
 *
 * Limits: num_tokens <= MLAPO_STATIC_MAX_TOKENS and hidden_dim <= MLAPO_STATIC_MAX_HIDDEN (any head count); anything
 * that would need more than the array returns MLAPO_ERR_WORKSPACE with a message naming the sizes, and nothing is launched.
 * First use: the first call on each device looks the array's device address up once (the device that owns the buffers in
 * `p` must be current, as for every call); make that first call during warm-up, before any stream capture.
 * Sharing: there is one array per device and every call uses it, so two calls that can overlap on the GPU (different
 * streams, or different host threads driving the same device) race on it — keep calls to this entry point on one stream
 * or order them yourself.  The array keeps whatever the previous call left in it; that is harmless because every byte
 * the operator reads from its workspace is written earlier in the same call, so nothing has to be cleared between calls.
 */
#define MLAPO_STATIC_MAX_TOKENS 256
#define MLAPO_STATIC_MAX_HIDDEN 8192
#define MLAPO_STATIC_SCRATCH_BYTES (20u << 20)
int mla_preprocess_launch_static(const MlapoParams* p, void* stream);

/* Number of kernels one mla_preprocess_launch() enqueues for these parameters (for launch accounting). */
int mlapo_launches_per_call(const MlapoParams* p);

/*
 * Debug / profiling hooks (no reference counterpart — the reference has DEBUG_MODE printf builds only,
 * CMakeLists.txt:48-53).  TEST AND PROFILING ONLY: the flags are one process-wide word; set them from one thread while
 * no call is being enqueued by another.  Bits 1 and 3 enlarge the workspace, so mlapo_get_workspace_size() and the launch
 * must see the same flags (a launch takes one snapshot of the word and refuses a workspace that is too small for it).
 * flags: bit0 = synchronise and check after every kernel (errors name the kernel; also serialises the kernels),
 * bit1 = also write the dequantised MM1 rows y0 (act [N,2112]) after the last workspace region (and, at prefill sizes, the raw
 * int32 sums where the split-K partial sums normally go), bit2 = record CUDA events around every kernel of the next launches,
 * bit3 = per-CTA clock marks (timeline region), bit11 = log stuck barrier waits to a host-mapped buffer before trapping,
 * bit12 = the quantise kernel also asks for the call's weights in L2 (measured slower; kept for A/B), bit13 = use the CTA-pair
 * q kernel for every call of more than 128 tokens, bit14 = one CTA per (head, token tile) unit even when there are few units,
 * bit16 = MM1 on single CTAs (8 K-splits) where CTA pairs (4 K-splits) would run, bit17 = the CTA-per-row norm kernel at prefill sizes,
 * bit19 = 128-column weight tiles for MM1 on CTA pairs where 256-column tiles would be used (more than 256 tokens).
 * bit20 = the CTA-pair q kernel's drain groups of every second CTA idle 6000 cycles before chunks 2 and 3 (regression test for a
 *         barrier-phase race between consecutive units),
 * bit21 = race hunting: every fourth mbarrier wait of every kernel is preceded by a random idle period of up to 3000 cycles (set
 *         and cleared on the current device).
 * Bits 3, 20 and 21 act only in the INSTRUMENTED build of the same sources (-DMLAPO_INSTRUMENTED ->
 * libmla_preprocess_b200_instr.so; mlapo_debug_is_instrumented() tells): the shipped library compiles the clock marks, the delay
 * hook and the jitter out — even untaken, those branches cost 3-4 % of a decode call.  tools/timeline.py, the chain spans of
 * bench.py and tests/test_jitter_gpu.py load the instrumented library (MLAPO_LIB); everything else, the timed paths above all,
 * runs the shipped one.
 * Returns the previous flags.
 */
int mlapo_set_debug_flags(int flags);
/* 1 if this is the instrumented build (debug bits 3, 20, 21 act), else 0. */
int mlapo_debug_is_instrumented(void);
/* Debug (flag bit11 = log stuck barrier waits instead of trapping): {count, {cta, thread, barrier, parity} x count}. */
int mlapo_debug_read_wait_log(unsigned* out, int cap_words);
/* With bit2 set: per-kernel durations (ms) of the last launch, in launch order; returns how many were written. */
int mlapo_get_kernel_times_ms(float* out_ms, int cap);
const char* mlapo_kernel_name(int index);

#ifdef __cplusplus
}
#endif
#endif /* MLA_PREPROCESS_B200_H_ */
