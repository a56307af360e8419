"""Mirror of the reference module python/sgl_kernel_npu/sgl_kernel_npu/norm/split_qkv_rmsnorm_mrope.py (function
triton_split_qkv_rmsnorm_mrope, :335-419; the name is the reference's public one, nothing here is Triton) on the CUDA kernels
behind split_qkv_rmsnorm_mrope_launch.  Same arguments, defaults and return values (q, k, v, gate; gate is [tokens, 0] without
has_gate)."""
import ctypes

import torch

from .. import cabi

cabi.lib.split_qkv_rmsnorm_mrope_launch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32] + [ctypes.c_void_p] * 9 + \
    [ctypes.c_int64] + [ctypes.c_int32] * 4 + [ctypes.c_float] + [ctypes.c_int32] * 5 + [ctypes.c_void_p]


def triton_split_qkv_rmsnorm_mrope(qkv, q_weight, k_weight, cos_sin, num_q_heads, num_kv_heads, head_size, eps, mrope_section, is_interleaved,
                                   rope_dim=None, q_bias=None, k_bias=None, has_gate=False):
    q_size, kv_size = num_q_heads * head_size, num_kv_heads * head_size
    num_tokens = qkv.shape[0]
    gate_size = q_size if has_gate else 0
    if rope_dim is None:
        rope_dim = head_size
    if qkv.dtype not in (torch.bfloat16, torch.float16):
        raise RuntimeError("qkv must be bfloat16 or float16")
    if qkv.dim() != 2 or qkv.shape[1] != q_size + gate_size + 2 * kv_size:
        raise RuntimeError("qkv must be [tokens, q_size + gate_size + 2 * kv_size]")
    if tuple(cos_sin.shape) != (3, num_tokens, rope_dim) or cos_sin.dtype not in (qkv.dtype, torch.float32):
        raise RuntimeError("cos_sin must be [3, tokens, rope_dim] in qkv's dtype or float32")
    if len(mrope_section) != 3:
        raise RuntimeError("mrope_section must hold the t, h and w section sizes")
    for t in [qkv, cos_sin, q_weight, k_weight] + [b for b in (q_bias, k_bias) if b is not None]:
        if not t.is_cuda or not t.is_contiguous():
            raise RuntimeError("all tensors must be contiguous CUDA tensors")
    for t in (q_weight, k_weight, q_bias, k_bias):
        if t is not None and t.dtype != qkv.dtype:
            raise RuntimeError("norm weights and biases must have qkv's dtype")
    new = lambda w: torch.empty(num_tokens, w, device=qkv.device, dtype=qkv.dtype)
    q, k, v, gate = new(q_size), new(kv_size), new(kv_size), new(gate_size)
    if num_tokens == 0:
        return q, k, v, gate
    selected = torch.empty(num_tokens, rope_dim, device=qkv.device, dtype=cos_sin.dtype)  # the kernel's workspace
    ptr = lambda t: None if t is None else t.data_ptr()
    with torch.cuda.device(qkv.device):
        cabi._check(cabi.lib.split_qkv_rmsnorm_mrope_launch(
            qkv.data_ptr(), cos_sin.data_ptr(), 1 if cos_sin.dtype == torch.float32 else 0, selected.data_ptr(), q.data_ptr(), k.data_ptr(),
            v.data_ptr(), gate.data_ptr() if has_gate else None, q_weight.data_ptr(), k_weight.data_ptr(), ptr(q_bias), ptr(k_bias), num_tokens,
            num_q_heads, num_kv_heads, head_size, rope_dim, float(eps), int(mrope_section[0]), int(mrope_section[1]), int(mrope_section[2]),
            1 if is_interleaved else 0, 0 if qkv.dtype == torch.bfloat16 else 1, cabi._stream_ptr(qkv)), "split_qkv_rmsnorm_mrope_launch")
    return q, k, v, gate
