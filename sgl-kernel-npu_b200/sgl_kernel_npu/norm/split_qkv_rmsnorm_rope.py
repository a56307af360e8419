"""Mirror of the reference module python/sgl_kernel_npu/sgl_kernel_npu/norm/split_qkv_rmsnorm_rope.py (function
split_qkv_rmsnorm_rope, :374-439) on the CUDA kernel behind split_qkv_rmsnorm_rope_launch.  Same arguments, defaults and
return values.  split_qkvgate_gemma_rmsnorm_rope (:686-747) is the module's gated variant: rows [q_0 | gate_0 | q_1 | gate_1 ... | k | v],
RMSNorm with (1 + weight), returns (q, k, v, gate)."""
import ctypes

import torch

from .. import cabi

cabi.lib.split_qkv_rmsnorm_rope_launch.argtypes = [ctypes.c_void_p] * 10 + [ctypes.c_int64] + [ctypes.c_int32] * 4 + [ctypes.c_float] + \
    [ctypes.c_int32] * 3 + [ctypes.c_void_p]
cabi.lib.split_qkvgate_gemma_rmsnorm_rope_launch.argtypes = [ctypes.c_void_p] * 9 + [ctypes.c_int64] + [ctypes.c_int32] * 4 + [ctypes.c_float] + \
    [ctypes.c_int32, ctypes.c_void_p]


def _check_tensors(input, others, width):
    if input.dtype not in (torch.bfloat16, torch.float16):
        raise RuntimeError("input must be bfloat16 or float16")
    if input.dim() != 2 or input.shape[1] != width:
        raise RuntimeError(f"input must be [batch, {width}]")
    for t in [input] + others:
        if not t.is_cuda or not t.is_contiguous() or t.dtype != input.dtype:
            raise RuntimeError("all tensors must be contiguous CUDA tensors of the input's dtype")


def split_qkvgate_gemma_rmsnorm_rope(input, sin, cos, q_hidden_size, kv_hidden_size, head_dim, rope_dim, eps, q_weight, k_weight):
    if head_dim & (head_dim - 1):
        raise AssertionError("head_dim must be a power of two")          # reference: assert KV_BLOCK_SIZE == head_dim
    if q_hidden_size % kv_hidden_size:
        raise AssertionError("q_hidden_size must be a multiple of kv_hidden_size")
    _check_tensors(input, [sin, cos, q_weight, k_weight], 2 * q_hidden_size + 2 * kv_hidden_size)
    batch = input.shape[0]
    if sin.numel() != batch * rope_dim or cos.numel() != batch * rope_dim:
        raise RuntimeError("sin / cos must hold [batch, rope_dim] elements")
    new = lambda w: torch.empty(batch, w, device=input.device, dtype=input.dtype)
    q, k, v, gate = new(q_hidden_size), new(kv_hidden_size), new(kv_hidden_size), new(q_hidden_size)
    if batch == 0:
        return q, k, v, gate
    with torch.cuda.device(input.device):
        cabi._check(cabi.lib.split_qkvgate_gemma_rmsnorm_rope_launch(
            input.data_ptr(), sin.data_ptr(), cos.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), gate.data_ptr(), q_weight.data_ptr(),
            k_weight.data_ptr(), batch, q_hidden_size, kv_hidden_size, head_dim, rope_dim, float(eps),
            0 if input.dtype == torch.bfloat16 else 1, cabi._stream_ptr(input)), "split_qkvgate_gemma_rmsnorm_rope_launch")
    return q, k, v, gate


def split_qkv_rmsnorm_rope(input, sin, cos, q_hidden_size, kv_hidden_size, head_dim, eps=None, q_weight=None, k_weight=None,
                           q_bias=None, k_bias=None, is_neox_style=True):
    if head_dim & (head_dim - 1):
        raise AssertionError("head_dim must be a power of two")          # reference: assert KV_BLOCK_SIZE == head_dim
    if q_hidden_size % kv_hidden_size:
        raise AssertionError("q_hidden_size must be a multiple of kv_hidden_size")
    if input.dtype not in (torch.bfloat16, torch.float16):
        raise RuntimeError("input must be bfloat16 or float16")
    batch = input.shape[0]
    if input.dim() != 2 or input.shape[1] != q_hidden_size + 2 * kv_hidden_size:
        raise RuntimeError("input must be [batch, q_hidden_size + 2 * kv_hidden_size]")
    rope_dim = sin.shape[-1]
    tensors = [input, sin, cos] + [t for t in (q_weight, k_weight, q_bias, k_bias) if t is not None]
    for t in tensors:
        if not t.is_cuda or not t.is_contiguous() or t.dtype != input.dtype:
            raise RuntimeError("all tensors must be contiguous CUDA tensors of the input's dtype")
    if sin.numel() != batch * rope_dim or cos.numel() != batch * rope_dim:
        raise RuntimeError("sin / cos must hold [batch, rope_dim] elements")
    q = torch.empty(batch, q_hidden_size, device=input.device, dtype=input.dtype)
    k = torch.empty(batch, kv_hidden_size, device=input.device, dtype=input.dtype)
    v = torch.empty(batch, kv_hidden_size, device=input.device, dtype=input.dtype)
    if batch == 0:
        return q, k, v
    ptr = lambda t: None if t is None else t.data_ptr()
    with torch.cuda.device(input.device):
        cabi._check(cabi.lib.split_qkv_rmsnorm_rope_launch(
            input.data_ptr(), sin.data_ptr(), cos.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), ptr(q_weight), ptr(k_weight),
            ptr(q_bias), ptr(k_bias), batch, q_hidden_size, kv_hidden_size, head_dim, rope_dim, 0.0 if eps is None else float(eps),
            1 if eps is not None else 0, 1 if is_neox_style else 0, 0 if input.dtype == torch.bfloat16 else 1,
            cabi._stream_ptr(input)), "split_qkv_rmsnorm_rope_launch")
    return q, k, v
