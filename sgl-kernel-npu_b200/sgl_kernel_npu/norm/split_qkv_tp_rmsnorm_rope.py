"""Mirror of the reference module python/sgl_kernel_npu/sgl_kernel_npu/norm/split_qkv_tp_rmsnorm_rope.py (function
split_qkv_tp_rmsnorm_rope, :172-288) on the CUDA kernel behind split_qkv_tp_rmsnorm_rope_launch.  Same arguments and return values.
tp_world == 1: one pass.  tp_world > 1, EXCHANGE == "allreduce" (default; also batches beyond the buffer's capacity, or no symmetric
memory on this system): local means -> dist.all_reduce over tp_group (8 bytes per token, as the reference does between its two
kernels, :253-254) -> one-pass apply.  EXCHANGE == "peer" (opt-in: set the module attribute or SGL_KERNEL_NPU_TP_EXCHANGE=peer): ONE
kernel per rank; the ranks exchange the per-token means inside it through a symmetric-memory buffer
(torch.distributed._symmetric_memory: every rank's buffer mapped into every process, plain NVLink peer stores of 8-byte records).
The one-kernel form halves the latency but asks more of the caller, which is why it is not the default: every rank of the group
must make the same sequence of calls with the same batch; the kernel's CTAs must all be resident at once (no other kernel may hold
SMs for good on any rank); one call per group may be in flight; an error return on one rank (so that it skips a launch its peers
make) leaves the group's epochs apart until the process group is rebuilt; and a rank whose peer never launches traps after its
bounded wait (about half a minute), which takes that rank's CUDA context down."""
import ctypes
import os
import warnings

import torch
import torch.distributed as dist

from .. import cabi

EXCHANGE = os.environ.get("SGL_KERNEL_NPU_TP_EXCHANGE", "allreduce")  # "allreduce" | "peer"
CAPACITY = 16384  # tokens per call the exchange buffer holds (2 * tp_world * CAPACITY * 16 bytes per rank)
_state = {}

cabi.lib.split_qkv_tp_rmsnorm_rope_peer_launch.argtypes = [ctypes.c_void_p] * 8 + [ctypes.c_int64] + [ctypes.c_int32] * 4 + [ctypes.c_float] + \
    [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]


def _exchange_state(tp_group, tp_world, device):
    """the exchange buffer of this (group, device): allocated, zeroed and rendezvoused once; None when symmetric memory is not available"""
    group = tp_group if tp_group is not None else dist.group.WORLD
    key = (id(group), device.index)
    if key not in _state:
        try:
            import torch.distributed._symmetric_memory as symm_mem
            buf = symm_mem.empty(2 * tp_world * CAPACITY * 4, dtype=torch.float32, device=device)
            buf.zero_()
            epoch = torch.tensor([1, 0], dtype=torch.int32, device=device)  # {next call's epoch, CTAs done}: advanced by the kernel
            hdl = symm_mem.rendezvous(buf, group)
            torch.cuda.synchronize(device)
            dist.barrier(group=group)  # every rank's buffer is zeroed before any rank publishes into it
            ptrs = (ctypes.c_void_p * tp_world)(*[int(p) for p in hdl.buffer_ptrs])
            _state[key] = dict(buf=buf, hdl=hdl, ptrs=ptrs, epoch=epoch, rank=dist.get_rank(group))
        except Exception as e:  # no peer access / no symmetric memory here: the all-reduce path is the reference's own
            warnings.warn(f"split_qkv_tp_rmsnorm_rope: symmetric memory unavailable ({e!r}); using dist.all_reduce")
            _state[key] = None
    return _state[key]

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
    st = _exchange_state(tp_group, tp_world, input.device) if EXCHANGE == "peer" and batch <= CAPACITY and tp_world <= 8 else None
    if st is not None:
        with torch.cuda.device(input.device):
            cabi._check(cabi.lib.split_qkv_tp_rmsnorm_rope_peer_launch(
                input.data_ptr(), cos.data_ptr(), sin.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), q_weight.data_ptr(),
                k_weight.data_ptr(), batch, q_hidden_size, kv_hidden_size, head_dim, rotary_dim, float(eps), st["ptrs"], st["rank"], tp_world,
                st["epoch"].data_ptr(), CAPACITY, 0 if input.dtype == torch.bfloat16 else 1, cabi._stream_ptr(input)), "split_qkv_tp_rmsnorm_rope_peer_launch")
        return q, k, v
    qk_var = torch.empty(batch, 2, dtype=torch.float32, device=input.device)
    _launch(MODE_SUMS, input, cos, sin, q, k, v, q_weight, k_weight, qk_var, *args)
    dist.all_reduce(qk_var, group=tp_group)
    _launch(MODE_APPLY, input, cos, sin, q, k, v, q_weight, k_weight, qk_var, *args)
    return q, k, v
