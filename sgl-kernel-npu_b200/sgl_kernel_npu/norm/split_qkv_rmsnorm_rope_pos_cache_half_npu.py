"""Mirror of the reference module python/sgl_kernel_npu/sgl_kernel_npu/norm/split_qkv_rmsnorm_rope_pos_cache_half_npu.py (function
split_qkv_rmsnorm_rope_pos_cache_half_npu, :232-407) on the CUDA kernels behind split_qkv_rmsnorm_rope_pos_cache_launch.  Same
arguments, defaults, checks and return values; `wide_grid_min_tokens` (an Ascend launch-grid switch) is accepted and has no effect.
`cast_norm_to_bf16` rounds the normalised value to the tensor's dtype (bf16 in the reference's use)."""
import ctypes

import torch

from .. import cabi

cabi.lib.split_qkv_rmsnorm_rope_pos_cache_launch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_int32,
                                                             ctypes.c_int64, ctypes.c_int32] + [ctypes.c_void_p] * 7 + [ctypes.c_int64] + \
    [ctypes.c_int32] * 4 + [ctypes.c_float] + [ctypes.c_int32] * 3 + [ctypes.c_void_p]


def split_qkv_rmsnorm_rope_pos_cache_half_npu(input_tensor, positions, cos_sin_cache, q_hidden_size, kv_hidden_size, head_dim, eps=None,
                                              q_weight=None, k_weight=None, q_bias=None, k_bias=None, rope_dim=None,
                                              cast_norm_to_bf16=True, wide_grid_min_tokens=1024):
    assert input_tensor.dim() == 2
    B, total_hidden = input_tensor.shape
    if rope_dim is None:
        rope_dim = head_dim
    assert rope_dim % 2 == 0 and rope_dim <= head_dim
    assert total_hidden == q_hidden_size + 2 * kv_hidden_size
    pos = positions
    assert pos.numel() == B, f"positions must be [B], got numel={pos.numel()} B={B}"
    if pos.dtype not in (torch.int32, torch.int64):
        pos = pos.to(torch.int32)
    pos = pos.contiguous()
    cache = cos_sin_cache.contiguous()
    max_seq = cache.shape[0]
    assert max_seq >= 1, "cos_sin_cache must be non-empty"
    assert head_dim & (head_dim - 1) == 0, "this kernel assumes head_dim is power-of-2"
    assert q_hidden_size % kv_hidden_size == 0
    norms, bias = eps is not None, q_bias is not None
    for name, t, need in (("q_weight", q_weight, norms), ("k_weight", k_weight, norms), ("q_bias", q_bias, bias), ("k_bias", k_bias, bias)):
        if need and (t is None or t.numel() < head_dim):
            raise ValueError(f"{name} must have at least head_dim={head_dim} elements, got {t.numel() if t is not None else 0}.")
    if input_tensor.dtype not in (torch.bfloat16, torch.float16):
        raise RuntimeError("input_tensor must be bfloat16 or float16")
    if cache.dtype not in (input_tensor.dtype, torch.float32) or cache.dim() != 2 or cache.shape[1] < rope_dim:
        raise RuntimeError("cos_sin_cache must be [max_seq, >= rope_dim] in the input's dtype or float32")
    for t in [input_tensor, pos, cache] + [w for w in (q_weight, k_weight, q_bias, k_bias) if w is not None]:
        if not t.is_cuda or not t.is_contiguous():
            raise RuntimeError("all tensors must be contiguous CUDA tensors")
    for w in (q_weight, k_weight, q_bias, k_bias):
        if w is not None and w.dtype != input_tensor.dtype:
            raise RuntimeError("norm weights and biases must have the input's dtype")
    new = lambda w: torch.empty((B, w), device=input_tensor.device, dtype=input_tensor.dtype)
    q, k, v = new(q_hidden_size), new(kv_hidden_size), new(kv_hidden_size)
    if B == 0:
        return q, k, v
    ptr = lambda t: None if t is None else t.data_ptr()
    with torch.cuda.device(input_tensor.device):
        cabi._check(cabi.lib.split_qkv_rmsnorm_rope_pos_cache_launch(
            input_tensor.data_ptr(), pos.data_ptr(), 1 if pos.dtype == torch.int64 else 0, cache.data_ptr(),
            1 if cache.dtype == torch.float32 else 0, cache.stride(0), max_seq, q.data_ptr(), k.data_ptr(), v.data_ptr(),
            ptr(q_weight) if norms else None, ptr(k_weight) if norms else None, ptr(q_bias) if norms else None, ptr(k_bias) if norms else None,
            B, q_hidden_size, kv_hidden_size, head_dim, rope_dim, float(eps) if norms else 0.0, 1 if norms else 0,
            1 if cast_norm_to_bf16 else 0, 0 if input_tensor.dtype == torch.bfloat16 else 1, cabi._stream_ptr(input_tensor)),
            "split_qkv_rmsnorm_rope_pos_cache_launch")
    return q, k, v
