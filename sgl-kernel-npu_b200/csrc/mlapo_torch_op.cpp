// torch.ops.npu.mla_preprocess for B200 — the reference's operator surface, re-hosted on the CUDA dispatch key.
//
// Mirrors (same schema string, same C++ signature, same mode strings / defaults / error texts):
//   schema      csrc/pytorch_extensions.cpp:105-115   (TORCH_LIBRARY_FRAGMENT(npu, m), m.def("mla_preprocess(...)"))
//   impl        csrc/pytorch_extensions.cpp:205, 268  (PrivateUse1 there, CUDA here)
//   signature   include/sgl_kenel_npu_ops.h:102-119
//   host logic  csrc/mla_preprocess/op_host/mla_preprocess.cpp:605-704
//   version ops csrc/pytorch_extensions.cpp:26-27, csrc/utils/version.h:23-45
// All compute goes through the C ABI in include/mla_preprocess_b200.h; there is no CPU or eager fallback.
#include <ATen/ATen.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/library.h>

#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <string>
#include <unordered_map>

#include "mla_preprocess_b200.h"

namespace sglang {
namespace npu_kernel {

namespace {

// op_host/mla_preprocess.cpp:605-611
const std::unordered_map<std::string, int> cache_mode_map = {{"krope_ctkv", 1}, {"int8_nzcache", 2}, {"nzcache", 3}};
const std::unordered_map<std::string, int> quant_mode_map = {{"per_tensor_quant_asymm", 0}, {"per_token_quant_symm", 1}};

int get_op_mode(const std::unordered_map<std::string, int>& mode_map, const c10::optional<c10::string_view>& mode_opt,
                const char* default_mode, const char* mode_name) {
  const std::string mode_str = mode_opt.has_value() ? std::string(mode_opt->data(), mode_opt->size()) : std::string(default_mode);
  auto it = mode_map.find(mode_str);
  TORCH_CHECK(it != mode_map.end(), "Unsupported ", mode_name, " value: '", mode_str, "'");  // same text as :619
  return it->second;
}

void check_cabi(int rc, const char* what) { TORCH_CHECK(rc == MLAPO_OK, what, " failed (", rc, "): ", mlapo_last_error()); }

// The op signature has no "prepare" hook, so FRACTAL_NZ weights are repacked on first sight and cached.  The key is the
// source tensor's StorageImpl (kept addressable by a weak reference, so a recycled device address or a recycled
// StorageImpl cannot alias a dead entry), its offset, shape and version counter (in-place updates invalidate; tensors
// created under torch.inference_mode() carry no counter and count as version 0, so an in-place update of such a weight
// needs a new tensor).  Weights are static in serving, so this runs once per layer at warm-up; entries die with their
// source tensor.  The repack is enqueued on the stream of the first call; an event recorded behind it orders every other
// stream that later receives the cached tensor.  Nothing is added to the cache while a stream is being captured (the
// repacked tensor would live in the graph's private pool and the repack would be replayed with the graph): warm up with
// one eager call, or repack explicitly with mla_preprocess_repack_weight and pass the K-major tensor.
struct RepackKey {
  const void* storage;
  int64_t offset, k, n;
  uint32_t version;
  bool operator==(const RepackKey& o) const {
    return storage == o.storage && offset == o.offset && k == o.k && n == o.n && version == o.version;
  }
};
struct RepackKeyHash {
  size_t operator()(const RepackKey& k) const {
    return std::hash<const void*>()(k.storage) ^ (std::hash<int64_t>()(k.k * 1315423911u + k.n + k.offset) << 1) ^ k.version;
  }
};
struct RepackEntry {
  c10::weak_intrusive_ptr<c10::StorageImpl> source;
  at::Tensor km;
  cudaStream_t made_on;
  cudaEvent_t ready;  // recorded on made_on behind the repack kernel (never destroyed: a handful per process)
};
std::mutex g_repack_mutex;
std::unordered_map<RepackKey, RepackEntry, RepackKeyHash> g_repack_cache;

at::Tensor repack_nz(const at::Tensor& w_nz, cudaStream_t stream) {
  TORCH_CHECK(w_nz.is_cuda() && w_nz.scalar_type() == at::kChar && w_nz.is_contiguous(), "weight must be a contiguous int8 CUDA tensor");
  TORCH_CHECK(w_nz.dim() == 4 && w_nz.size(0) == 1 && w_nz.size(3) == 32, "expected FRACTAL_NZ weight [1, K/32, n_out, 32], got ",
              w_nz.sizes());
  const int64_t k = w_nz.size(1) * 32, n = w_nz.size(2);
  at::Tensor km = at::empty({n, k}, w_nz.options());
  check_cabi(mlapo_repack_nz_weight(w_nz.data_ptr(), km.data_ptr(), (int32_t)k, (int32_t)n, stream), "mlapo_repack_nz_weight");
  return km;
}

// Accepts the reference's NZ layout [1, K/32, n_out, 32] (repacked + cached) or an already K-major [n_out, K] tensor.
at::Tensor kmajor_weight(const at::Tensor& w, int64_t n_out, int64_t k, const char* name, cudaStream_t stream) {
  TORCH_CHECK(w.scalar_type() == at::kChar, name, " must be int8 (W8A8 path)");
  if (w.dim() == 2) {
    TORCH_CHECK(w.size(0) == n_out && w.size(1) == k && w.is_contiguous(), name, " (K-major) must be contiguous [", n_out, ", ", k, "]");
    return w;
  }
  TORCH_CHECK(w.dim() == 4 && w.size(0) == 1 && w.size(1) * 32 == k && w.size(2) == n_out && w.size(3) == 32, name,
              " must be FRACTAL_NZ [1, ", k / 32, ", ", n_out, ", 32], got ", w.sizes());
  const uint32_t version = w.is_inference() ? 0u : (uint32_t)w._version();
  RepackKey key{w.storage().unsafeGetStorageImpl(), w.storage_offset(), k, n_out, version};
  std::lock_guard<std::mutex> lock(g_repack_mutex);
  auto it = g_repack_cache.find(key);
  if (it != g_repack_cache.end()) {
    if (!it->second.source.expired()) {
      if (it->second.made_on != stream)
        TORCH_CHECK(cudaStreamWaitEvent(stream, it->second.ready, 0) == cudaSuccess, "mla_preprocess: cannot order the stream behind the weight repack");
      return it->second.km;
    }
    g_repack_cache.erase(it);
  }
  cudaStreamCaptureStatus capturing = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(stream, &capturing);
  TORCH_CHECK(capturing == cudaStreamCaptureStatusNone, "mla_preprocess: ", name,
              " is a FRACTAL_NZ weight seen for the first time inside a CUDA graph capture; run one eager warm-up call first, or "
              "repack it with torch.ops.npu.mla_preprocess_repack_weight and pass the K-major tensor");
  for (auto e = g_repack_cache.begin(); e != g_repack_cache.end();)  // drop entries whose source weight is gone
    e = e->second.source.expired() ? g_repack_cache.erase(e) : std::next(e);
  at::Tensor km = repack_nz(w, stream);
  cudaEvent_t ready = nullptr;
  TORCH_CHECK(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming) == cudaSuccess && cudaEventRecord(ready, stream) == cudaSuccess,
              "mla_preprocess: cannot record the weight-repack event");
  g_repack_cache.emplace(key, RepackEntry{w.storage().getWeakStorageImpl(), km, stream, ready});
  return km;
}

}  // namespace

// include/sgl_kenel_npu_ops.h:102-119 — identical signature
std::tuple<at::Tensor&, at::Tensor&, at::Tensor&, at::Tensor&> mla_preprocess(
    const at::Tensor& hiddenState, const at::Tensor& gamma0, const at::Tensor& beta0, const at::Tensor& wdqkv,
    const at::Tensor& descale0, const at::Tensor& gamma1, const at::Tensor& beta1, const at::Tensor& wuq,
    const at::Tensor& descale1, const at::Tensor& gamma2, const at::Tensor& cos, const at::Tensor& sin, const at::Tensor& wuk,
    const at::Tensor& kv_cache, const at::Tensor& kv_cache_rope, const at::Tensor& slotmapping, const at::Tensor& quant_scale0,
    const at::Tensor& quant_offset0, const at::Tensor& bias0, const at::Tensor& quant_scale1, const at::Tensor& quant_offset1,
    const at::Tensor& bias1, const c10::optional<at::Tensor>& ctkv_scale, const c10::optional<at::Tensor>& q_nope_scale,
    c10::optional<c10::string_view> cache_mode, c10::optional<c10::string_view> quant_mode, at::Tensor& q_out0,
    at::Tensor& kv_cache_out0, at::Tensor& q_out1, at::Tensor& kv_cache_out1) {
  // op_host/mla_preprocess.cpp:634-635 (same defaults)
  const int cacheMode = get_op_mode(cache_mode_map, cache_mode, "krope_ctkv", "cache_mode");
  const int quantMode = get_op_mode(quant_mode_map, quant_mode, "per_token_quant_symm", "quant_mode");
  (void)gamma0;  // bound and never read by the reference either (mla_preprocess_mix_bf16.hpp:2762, 424-486)
  (void)beta0;

  TORCH_CHECK(hiddenState.is_cuda(), "mla_preprocess: hiddenState must be a CUDA tensor (no CPU fallback)");
  const auto act = hiddenState.scalar_type();
  TORCH_CHECK(act == at::kBFloat16 || act == at::kHalf, "hiddenState must be bfloat16 or float16");
  TORCH_CHECK(hiddenState.dim() == 2 && hiddenState.is_contiguous(), "hiddenState must be contiguous [tokenNum, hiddenSize]");
  const int64_t N = hiddenState.size(0), D = hiddenState.size(1);  // :658-660
  TORCH_CHECK(wuk.dim() == 3 && wuk.size(1) == MLAPO_Q_NOPE && wuk.size(2) == MLAPO_KV_LORA, "wuk must be [headNum, 128, 512]");
  const int64_t H = wuk.size(0);
  const c10::cuda::CUDAGuard guard(hiddenState.device());
  cudaStream_t stream = at::cuda::getCurrentCUDAStream();

  auto dev_ok = [&](const at::Tensor& t, const char* name, at::ScalarType dt, c10::IntArrayRef shape = {}) {
    TORCH_CHECK(t.is_cuda() && t.device() == hiddenState.device(), name, " must be on ", hiddenState.device());
    TORCH_CHECK(t.is_contiguous(), name, " must be contiguous");
    TORCH_CHECK(t.scalar_type() == dt, name, " must be ", dt, ", got ", t.scalar_type());
    if (!shape.empty()) TORCH_CHECK(t.sizes() == shape, name, " must have shape ", shape, ", got ", t.sizes());
  };
  const auto dsc_dt = descale0.scalar_type();
  TORCH_CHECK(dsc_dt == at::kFloat || dsc_dt == at::kLong, "descale0 must be float32 (bf16 path) or int64 (fp16 path)");
  dev_ok(descale0, "descale0", dsc_dt, {MLAPO_MM1_OUT});
  dev_ok(descale1, "descale1", dsc_dt, {H * MLAPO_Q_HEAD});
  dev_ok(gamma1, "gamma1", act, {MLAPO_Q_LORA});
  dev_ok(beta1, "beta1", act, {MLAPO_Q_LORA});
  dev_ok(gamma2, "gamma2", act, {MLAPO_KV_LORA});
  dev_ok(cos, "cos", act, {N, MLAPO_ROPE});
  dev_ok(sin, "sin", act, {N, MLAPO_ROPE});
  dev_ok(wuk, "wuk", act);
  dev_ok(slotmapping, "slotmapping", at::kInt, {N});
  const bool per_tensor = quantMode == MLAPO_QUANT_PER_TENSOR_ASYMM;
  if (per_tensor) {
    dev_ok(quant_scale0, "quant_scale0", act);
    dev_ok(quant_offset0, "quant_offset0", at::kChar);
    dev_ok(bias0, "bias0", at::kInt, {MLAPO_MM1_OUT});
    dev_ok(quant_scale1, "quant_scale1", act);
    dev_ok(quant_offset1, "quant_offset1", at::kChar);
    dev_ok(bias1, "bias1", at::kInt, {H * MLAPO_Q_HEAD});
  }
  const auto q0_dt = cacheMode == MLAPO_CACHE_INT8_NZ ? at::kChar : act;
  dev_ok(q_out0, "q_out0", q0_dt, {N, H, MLAPO_KV_LORA});
  dev_ok(q_out1, "q_out1", act, {N, H, MLAPO_ROPE});
  dev_ok(kv_cache_out0, "kv_cache_out0", q0_dt);
  dev_ok(kv_cache_out1, "kv_cache_out1", act);
  TORCH_CHECK(kv_cache_out0.data_ptr() == kv_cache.data_ptr() && kv_cache_out1.data_ptr() == kv_cache_rope.data_ptr(),
              "kv_cache_out0/1 must be the same storage as kv_cache/kv_cache_rope");
  TORCH_CHECK(kv_cache.dim() == 4 && kv_cache_rope.dim() == 4, "caches must be 4-D");
  int64_t blocks, block_size;
  if (cacheMode == MLAPO_CACHE_KROPE_CTKV) {
    blocks = kv_cache.size(0);
    block_size = kv_cache.size(1);
    TORCH_CHECK(kv_cache.size(2) * kv_cache.size(3) == MLAPO_KV_LORA && kv_cache_rope.numel() == blocks * block_size * MLAPO_ROPE,
                "krope_ctkv caches must be [B,bs,1,512] and [B,bs,1,64]");
  } else {
    blocks = kv_cache.size(0);
    block_size = kv_cache.size(2);
    const int64_t c0 = cacheMode == MLAPO_CACHE_INT8_NZ ? 32 : 16;
    TORCH_CHECK(kv_cache.size(1) == MLAPO_KV_LORA / c0 && block_size == 128 && kv_cache.size(3) == c0 && kv_cache_rope.size(0) == blocks &&
                    kv_cache_rope.size(1) == 4 && kv_cache_rope.size(2) == 128 && kv_cache_rope.size(3) == 16,
                "NZ caches must be [B,", MLAPO_KV_LORA / c0, ",128,", c0, "] and [B,4,128,16] (blockSize 128)");
  }
  if (cacheMode == MLAPO_CACHE_INT8_NZ) {
    // the reference substitutes uninitialised placeholders here (op_host/mla_preprocess.cpp:636-643); that is undefined
    // in int8_nzcache mode, so this implementation insists on the real tensors
    TORCH_CHECK(ctkv_scale.has_value() && q_nope_scale.has_value(), "ctkv_scale and q_nope_scale are required for cache_mode 'int8_nzcache'");
    dev_ok(*ctkv_scale, "ctkv_scale", act);
    dev_ok(*q_nope_scale, "q_nope_scale", act, {H});
  }

  at::Tensor wdqkv_km = kmajor_weight(wdqkv, MLAPO_MM1_OUT, D, "wdqkv", stream);
  at::Tensor wuq_km = kmajor_weight(wuq, H * MLAPO_Q_HEAD, MLAPO_Q_LORA, "wuq", stream);

  MlapoParams p{};
  p.struct_bytes = sizeof(MlapoParams);
  p.num_tokens = (int32_t)N;
  p.num_heads = (int32_t)H;
  p.hidden_dim = (int32_t)D;
  p.act_dtype = act == at::kBFloat16 ? MLAPO_ACT_BF16 : MLAPO_ACT_FP16;
  p.cache_mode = cacheMode;
  p.quant_mode = quantMode;
  p.num_cache_blocks = (int32_t)blocks;
  p.cache_block_size = (int32_t)block_size;
  p.descale_is_int64 = dsc_dt == at::kLong;
  // The operator cannot see what precedes it on the stream, so it only takes the no-wait start (MLAPO_FLAG_INPUT_COMPLETE, see
  // include/mla_preprocess_b200.h) when the deployment says that no kernel that triggers its dependents early feeds this op.
  // Even then the flag is set per call, below, once the workspace is known: only when the previous operator call of this thread
  // ran on the same stream with the same shape and got the same workspace memory (see the header: a different layout could put
  // this call's first stores under the a1 rows the previous call's q kernel is still reading).
  static const bool input_complete = [] { const char* e = std::getenv("SGL_KERNEL_NPU_MLAPO_INPUT_COMPLETE"); return e && e[0] == '1'; }();
  p.flags = 0;
  p.hidden_state = hiddenState.data_ptr();
  p.wdqkv_km = wdqkv_km.data_ptr();
  p.descale0 = descale0.data_ptr();
  p.gamma1 = gamma1.data_ptr();
  p.beta1 = beta1.data_ptr();
  p.wuq_km = wuq_km.data_ptr();
  p.descale1 = descale1.data_ptr();
  p.gamma2 = gamma2.data_ptr();
  p.cos = cos.data_ptr();
  p.sin = sin.data_ptr();
  p.wuk = wuk.data_ptr();
  p.slot_mapping = slotmapping.data_ptr();
  if (per_tensor) {
    p.quant_scale0 = quant_scale0.data_ptr();
    p.quant_offset0 = quant_offset0.data_ptr();
    p.bias0 = bias0.data_ptr();
    p.quant_scale1 = quant_scale1.data_ptr();
    p.quant_offset1 = quant_offset1.data_ptr();
    p.bias1 = bias1.data_ptr();
  }
  if (cacheMode == MLAPO_CACHE_INT8_NZ) {
    p.ctkv_scale = ctkv_scale->data_ptr();
    p.q_nope_scale = q_nope_scale->data_ptr();
  }
  p.q_out0 = q_out0.data_ptr();
  p.q_out1 = q_out1.data_ptr();
  p.kv_cache = kv_cache_out0.data_ptr();
  p.kv_cache_rope = kv_cache_out1.data_ptr();

  if (N > 0) {
    size_t ws_bytes = 0;
    check_cabi(mlapo_get_workspace_size(&p, &ws_bytes), "mlapo_get_workspace_size");
    // transient workspace from the caching allocator on the current stream (reference: at::empty, :677-680)
    at::Tensor workspace = at::empty({(int64_t)std::max<size_t>(ws_bytes, 256)}, hiddenState.options().dtype(at::kByte));
    struct LastCall { cudaStream_t stream; const void* ws; int64_t n, h, d; int quant, cache, dev; };
    static thread_local LastCall last{nullptr, nullptr, 0, 0, 0, -1, -1, -1};
    const LastCall now{stream, workspace.data_ptr(), N, (int64_t)p.num_heads, D, quantMode, cacheMode, (int)hiddenState.get_device()};
    const bool same_as_last = last.ws == now.ws && last.stream == now.stream && last.n == now.n && last.h == now.h && last.d == now.d &&
                              last.quant == now.quant && last.cache == now.cache && last.dev == now.dev;
    if (input_complete && same_as_last) p.flags = MLAPO_FLAG_INPUT_COMPLETE;
    last = now;
    check_cabi(mla_preprocess_launch(&p, workspace.data_ptr(), (size_t)workspace.numel(), stream), "mla_preprocess_launch");
  }
  return std::forward_as_tuple(q_out0, kv_cache_out0, q_out1, kv_cache_out1);  // :703
}

// Load-time helper for callers that want to repack explicitly and pass K-major weights (no cache involved).
at::Tensor mla_preprocess_repack_weight(const at::Tensor& w_nz) {
  const c10::cuda::CUDAGuard guard(w_nz.device());
  return repack_nz(w_nz, at::cuda::getCurrentCUDAStream());
}

// csrc/batch_matmul_transpose/op_host/batch_matmul_transpose.cpp:46-114 (same signature as include/sgl_kenel_npu_ops.h:121-124):
// tensor_c[m, batch, :] = tensor_a[m, batch, :] . tensor_b[batch]; the mode strings and their defaults are the reference's
// (:16-25, 65-67); quant_mode is validated and, as in the reference's fp16/bf16 kernel, does not change the arithmetic.
void batch_matmul_transpose(const at::Tensor& tensor_a, const at::Tensor& tensor_b, at::Tensor& tensor_c,
                            c10::optional<c10::string_view> format_mode, c10::optional<c10::string_view> quant_mode) {
  const std::string fmt(format_mode.value_or("ND")), qm(quant_mode.value_or("per_channel_symm"));
  TORCH_CHECK(fmt == "ND" || fmt == "NZ", "format_mode: Unsupported mode value ", fmt);
  TORCH_CHECK(qm == "per_channel_symm" || qm == "per_channel_asymm" || qm == "per_token_symm", "quant_mode: Unsupported mode value ", qm);
  TORCH_CHECK(tensor_a.scalar_type() == tensor_b.scalar_type() && tensor_b.scalar_type() == tensor_c.scalar_type(), "tensor type is not the same");
  TORCH_CHECK(tensor_a.scalar_type() == at::kBFloat16 || tensor_a.scalar_type() == at::kHalf, "tensor type only support half or bf16");
  TORCH_CHECK(tensor_a.is_cuda() && tensor_b.is_cuda() && tensor_c.is_cuda(), "batch_matmul_transpose: tensors must be CUDA tensors");
  TORCH_CHECK(tensor_a.is_contiguous() && tensor_b.is_contiguous() && tensor_c.is_contiguous(), "batch_matmul_transpose: tensors must be contiguous");
  TORCH_CHECK(tensor_a.dim() == 3, "batch size is not same between srcTensor and dstTensor");
  int64_t n;
  if (fmt == "ND") {
    TORCH_CHECK(tensor_b.dim() == 3, "tensor shape should be dim3 in ND format");
    TORCH_CHECK(tensor_a.size(2) == tensor_b.size(1), "tensor shape is wrong");
    n = tensor_b.size(2);
  } else {
    TORCH_CHECK(tensor_b.dim() == 4, "tensor shape should be dim4 in nz format");
    TORCH_CHECK(tensor_a.size(2) == tensor_b.size(2), "tensor shape is wrong");
    n = tensor_b.size(1) * tensor_b.size(3);
  }
  TORCH_CHECK(tensor_a.size(1) == tensor_b.size(0), "tensor shape is wrong");
  TORCH_CHECK(tensor_c.numel() == tensor_a.size(0) * tensor_a.size(1) * n, "tensor_c must hold [m, batch, n] elements");
  if (tensor_c.numel() == 0) return;
  const c10::cuda::CUDAGuard guard(tensor_a.device());
  check_cabi(batch_matmul_transpose_launch(tensor_a.data_ptr(), tensor_b.data_ptr(), tensor_c.data_ptr(), tensor_a.size(0), tensor_a.size(1),
                                           tensor_a.size(2), n, tensor_a.scalar_type() == at::kBFloat16 ? MLAPO_ACT_BF16 : MLAPO_ACT_FP16,
                                           fmt == "NZ" ? MLAPO_FORMAT_NZ : MLAPO_FORMAT_ND, at::cuda::getCurrentCUDAStream()),
             "batch_matmul_transpose_launch");
}

// ---- slot producers (csrc/pytorch_extensions.cpp:31-45): same schemas, CUDA tensors
static int is_i64(const at::Tensor& t, const char* name) {
  TORCH_CHECK(t.scalar_type() == at::kInt || t.scalar_type() == at::kLong, name, " must be int32 or int64");
  return t.scalar_type() == at::kLong ? 1 : 0;
}
static void need_cuda_contig(const at::Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda() && t.is_contiguous(), name, " must be a contiguous CUDA tensor");
}

void alloc_extend(const at::Tensor& pre_lens, const at::Tensor& seq_lens, const at::Tensor& last_loc, const at::Tensor& free_pages,
                  int64_t page_size, at::Tensor& out_indices, at::Tensor& values) {
  for (const at::Tensor* t : {&pre_lens, &seq_lens, &last_loc, &free_pages, (const at::Tensor*)&out_indices, (const at::Tensor*)&values})
    need_cuda_contig(*t, "alloc_extend tensor");
  const int i64 = is_i64(pre_lens, "pre_lens");
  TORCH_CHECK(seq_lens.scalar_type() == pre_lens.scalar_type() && last_loc.scalar_type() == pre_lens.scalar_type() &&
                  free_pages.scalar_type() == pre_lens.scalar_type() && values.scalar_type() == pre_lens.scalar_type(),
              "pre_lens, seq_lens, last_loc, free_pages and values must share one integer dtype");
  if (pre_lens.numel() == 0) return;
  const c10::cuda::CUDAGuard guard(pre_lens.device());
  check_cabi(mlapo_alloc_extend(pre_lens.data_ptr(), seq_lens.data_ptr(), last_loc.data_ptr(), free_pages.data_ptr(), page_size,
                                out_indices.data_ptr(), values.data_ptr(), (int32_t)pre_lens.numel(), out_indices.numel(), i64,
                                is_i64(out_indices, "out_indices"), at::cuda::getCurrentCUDAStream()),
             "mlapo_alloc_extend");
}

static void cache_loc_common(const at::Tensor& req, const at::Tensor& token_pool, const at::Tensor& start, const at::Tensor& end,
                             const at::Tensor& cache_loc, int update) {
  // csrc/cache_location_assign/op_host/cache_loc_assign.cpp:73-86
  TORCH_CHECK((req.scalar_type() == at::kInt || req.scalar_type() == at::kLong) && token_pool.scalar_type() == at::kInt &&
                  start.scalar_type() == at::kLong && end.scalar_type() == at::kLong && cache_loc.scalar_type() == at::kInt,
              "Only support inputTensor combo1: int64, int32, int64, int64, int32; combo2: int32, int32, int64, int64, int32");
  for (const at::Tensor* t : {&req, &token_pool, &start, &end, &cache_loc}) need_cuda_contig(*t, "cache_loc tensor");
  TORCH_CHECK(token_pool.dim() == 2, "token_pool must be [pool, row]");
  if (req.numel() == 0) return;
  const c10::cuda::CUDAGuard guard(token_pool.device());
  check_cabi(mlapo_cache_loc_assign(req.data_ptr(), req.scalar_type() == at::kLong, token_pool.data_ptr(), start.data_ptr(), end.data_ptr(),
                                    cache_loc.data_ptr(), (int32_t)req.numel(), token_pool.size(1), update, at::cuda::getCurrentCUDAStream()),
             "mlapo_cache_loc_assign");
}
at::Tensor cache_loc_assign(const at::Tensor& req, const at::Tensor& token_pool, const at::Tensor& start, const at::Tensor& end,
                            const at::Tensor& cache_loc) {
  cache_loc_common(req, token_pool, start, end, cache_loc, 0);
  return token_pool;
}
at::Tensor cache_loc_update(const at::Tensor& req, const at::Tensor& token_pool, const at::Tensor& start, const at::Tensor& end,
                            const at::Tensor& cache_loc) {
  cache_loc_common(req, token_pool, start, end, cache_loc, 1);
  return cache_loc;
}

bool assign_cache_op(at::Tensor& out, const at::Tensor& src, const at::Tensor& dst_start, const at::Tensor& dst_end,
                     const at::Tensor& src_start, const at::Tensor& src_end) {
  // csrc/assign_cache_op/op_host/assign_cache.cpp:40-82: malformed arguments return false instead of throwing
  const at::Tensor* idx[4] = {&dst_start, &dst_end, &src_start, &src_end};
  if (out.dim() < 2 || src.dim() < 1 || out.scalar_type() != src.scalar_type()) return false;
  for (auto* t : idx)
    if (t->dim() < 1 || t->size(0) != out.size(0) || t->scalar_type() != dst_start.scalar_type()) return false;
  if (dst_start.scalar_type() != at::kInt && dst_start.scalar_type() != at::kLong) return false;
  need_cuda_contig(out, "out");
  need_cuda_contig(src, "src");
  for (auto* t : idx) need_cuda_contig(*t, "index tensor");
  if (out.size(0) == 0) return true;
  const c10::cuda::CUDAGuard guard(out.device());
  check_cabi(mlapo_assign_cache(out.data_ptr(), src.data_ptr(), dst_start.data_ptr(), dst_end.data_ptr(), src_start.data_ptr(), src_end.data_ptr(),
                                (int32_t)out.size(0), out.stride(0), (int32_t)out.element_size(), dst_start.scalar_type() == at::kLong,
                                at::cuda::getCurrentCUDAStream()),
             "mlapo_assign_cache");
  return true;
}

std::string sgl_kernel_npu_version() { return std::string(mlapo_version()); }
void sgl_kernel_npu_print_version() { std::printf("%s\n", mlapo_version()); }

}  // namespace npu_kernel
}  // namespace sglang

// csrc/pytorch_extensions.cpp:24-27, 105-115 — schema strings verbatim (FRAGMENT so it coexists with other npu ops)
TORCH_LIBRARY_FRAGMENT(npu, m) {
  m.def("sgl_kernel_npu_print_version() -> ()", []() { sglang::npu_kernel::sgl_kernel_npu_print_version(); });
  m.def("sgl_kernel_npu_version() -> str", []() { return sglang::npu_kernel::sgl_kernel_npu_version(); });
  m.def(
      "mla_preprocess(Tensor hiddenState, Tensor gamma0, Tensor beta0, Tensor wdqkv, "
      "Tensor descale0, Tensor gamma1, Tensor beta1, Tensor wuq, "
      "Tensor descale1, Tensor gamma2, Tensor cos, Tensor sin, Tensor wuk,"
      "Tensor kv_cache, Tensor kv_cache_rope, Tensor slotmapping, "
      "Tensor quant_scale0, Tensor quant_offset0, Tensor bias0, "
      "Tensor quant_scale1, Tensor quant_offset1, Tensor bias1, *, "
      "Tensor? ctkv_scale=None, Tensor? q_nope_scale=None, "
      "str? cache_mode=None, str? quant_mode=None, "
      "Tensor(a!) q_out0, Tensor(b!) kv_cache_out0, Tensor(c!) q_out1, Tensor(d!) kv_cache_out1) "
      "-> (Tensor(a!), Tensor(b!), Tensor(c!), Tensor(d!))");
  m.def("mla_preprocess_repack_weight(Tensor w_nz) -> Tensor");
  m.def(  // csrc/pytorch_extensions.cpp:117-119, verbatim
      "batch_matmul_transpose(Tensor tensor_a, Tensor tensor_b, Tensor(a!) tensor_c, "
      "str? format_mode=None, str? quant_mode=None) -> ()");
  // csrc/pytorch_extensions.cpp:31-45, verbatim
  m.def(
      "alloc_extend(Tensor pre_lens, Tensor seq_lens, Tensor last_loc, Tensor free_pages, int page_size, "
      "Tensor(a!) out_indices, Tensor(b!) values) -> ()");
  m.def(
      "cache_loc_assign(Tensor req_indices, Tensor token_pool, Tensor start_offset, Tensor end_offset, Tensor "
      "out_cache_loc) -> Tensor");
  m.def(
      "cache_loc_update(Tensor req_indices, Tensor token_pool, Tensor start_offset, Tensor end_offset, Tensor "
      "out_cache_loc) -> Tensor");
  m.def(
      "assign_cache_op(Tensor! out, Tensor src, Tensor dst_start_idx, Tensor dst_end_idx, Tensor src_start_idx, "
      "Tensor src_end_idx) -> bool");
}

// csrc/pytorch_extensions.cpp:205, 268 register under PrivateUse1 (the NPU); here the device is CUDA.
TORCH_LIBRARY_IMPL(npu, CUDA, m) {
  m.impl("mla_preprocess", TORCH_FN(sglang::npu_kernel::mla_preprocess));
  m.impl("mla_preprocess_repack_weight", TORCH_FN(sglang::npu_kernel::mla_preprocess_repack_weight));
  m.impl("batch_matmul_transpose", TORCH_FN(sglang::npu_kernel::batch_matmul_transpose));
  m.impl("alloc_extend", TORCH_FN(sglang::npu_kernel::alloc_extend));
  m.impl("cache_loc_assign", TORCH_FN(sglang::npu_kernel::cache_loc_assign));
  m.impl("cache_loc_update", TORCH_FN(sglang::npu_kernel::cache_loc_update));
  m.impl("assign_cache_op", TORCH_FN(sglang::npu_kernel::assign_cache_op));
}
