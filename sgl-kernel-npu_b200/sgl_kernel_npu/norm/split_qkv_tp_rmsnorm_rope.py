"""Mirror of the reference module python/sgl_kernel_npu/sgl_kernel_npu/norm/split_qkv_tp_rmsnorm_rope.py (function
split_qkv_tp_rmsnorm_rope, :172-288) on the CUDA kernel behind split_qkv_tp_rmsnorm_rope_launch.  Same arguments and return values.
tp_world == 1: one pass.  tp_world > 1: local means -> dist.all_reduce over tp_group (8 bytes per token, as the reference does
between its two kernels, :253-254) -> one-pass apply."""
import ctypes

import torch
import torch.distributed as dist

from .. import cabi

cabi.lib.split_qkv_tp_rmsnorm_rope_launch.argtypes = [ctypes.c_void_p] * 9 + [ctypes.c_int64] + [ctypes.c_int32] * 4 + [ctypes.c_float] + \
    [ctypes.c_int32] * 3 + [ctypes.c_void_p]

MODE_FUSED, MODE_SUMS, MODE_APPLY = 0, 1, 2


def _launch(mode, input, cos, sin, q, k, v, q_weight, k_weight, qk_var, q_hidden_size, kv_hidden_size, head_dim, rotary_dim, eps, tp_world):
    with torch.cuda.device(input.device):
        cabi._check(cabi.lib.split_qkv_tp_rmsnorm_rope_launch(
            input.data_ptr(), cos.data_ptr(), sin.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), q_weight.data_ptr(), k_weight.data_ptr(),
            None if qk_var is None else qk_var.data_ptr(), input.shape[0], q_hidden_size, kv_hidden_size, head_dim, rotary_dim, float(eps),
            tp_world, mode, 0 if input.dtype == torch.bfloat16 else 1, cabi._stream_ptr(input)), "split_qkv_tp_rmsnorm_rope_launch")


def split_qkv_tp_rmsnorm_rope(input, cos, sin, q_hidden_size, kv_hidden_size, head_dim, eps, q_weight, k_weight, rotary_dim, tp_world, tp_group):
    if head_dim & (head_dim - 1):
        raise AssertionError("head_dim must be a power of two")          # reference: assert KV_BLOCK_SIZE == head_dim
    if q_hidden_size % kv_hidden_size:
        raise AssertionError("q_hidden_size must be a multiple of kv_hidden_size")
    if input.dtype not in (torch.bfloat16, torch.float16):
        raise RuntimeError("input must be bfloat16 or float16")
    batch = input.shape[0]
    if input.dim() != 2 or input.shape[1] != q_hidden_size + 2 * kv_hidden_size:
        raise RuntimeError("input must be [batch, q_hidden_size + 2 * kv_hidden_size]")
    for t in (input, cos, sin, q_weight, k_weight):
        if not t.is_cuda or not t.is_contiguous() or t.dtype != input.dtype:
            raise RuntimeError("all tensors must be contiguous CUDA tensors of the input's dtype")
    if cos.numel() != batch * rotary_dim or sin.numel() != batch * rotary_dim:
        raise RuntimeError("cos / sin must hold [batch, rotary_dim] elements")
    if q_weight.numel() != q_hidden_size or k_weight.numel() != kv_hidden_size:
        raise RuntimeError("q_weight / k_weight hold one value per column of q / k")
    new = lambda w: torch.empty(batch, w, device=input.device, dtype=input.dtype)
    q, k, v = new(q_hidden_size), new(kv_hidden_size), new(kv_hidden_size)
    if batch == 0:
        return q, k, v
    args = (q_hidden_size, kv_hidden_size, head_dim, rotary_dim, eps, tp_world)
    if tp_world == 1:
        _launch(MODE_FUSED, input, cos, sin, q, k, v, q_weight, k_weight, None, *args)
        return q, k, v
    qk_var = torch.empty(batch, 2, dtype=torch.float32, device=input.device)
    _launch(MODE_SUMS, input, cos, sin, q, k, v, q_weight, k_weight, qk_var, *args)
    dist.all_reduce(qk_var, group=tp_group)
    _launch(MODE_APPLY, input, cos, sin, q, k, v, q_weight, k_weight, qk_var, *args)
    return q, k, v
