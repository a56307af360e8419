"""ctypes binding of the C ABI in include/mla_preprocess_b200.h.

Host-side mirror of the reference operator for callers that do not go through the torch
dispatcher (tests, bench, other runtimes): `mla_preprocess(...)` below has the argument
list of `torch.ops.npu.mla_preprocess` (reference csrc/pytorch_extensions.cpp:105-115),
validates like the op host should (reference op_host/mla_preprocess.cpp:623-704 only checks
the two mode strings and N) and calls `mla_preprocess_launch` with raw device pointers.

There is no CPU fallback: if the CUDA library is missing, importing this module raises.
"""
import ctypes
import os

import torch

# MLAPO_LIB overrides the library file (A/B timing of two builds in one session; the instrumented build for the timeline and
# race-hunting tools); it must still be a build of this ABI.
_LIB_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib")
_LIB_PATH = os.environ.get("MLAPO_LIB") or os.path.join(_LIB_DIR, "libmla_preprocess_b200.so")
# Same sources with -DMLAPO_INSTRUMENTED (per-CTA clock marks, wait jitter, test hooks): for profiling tools and tests only.
INSTRUMENTED_LIB = os.path.join(_LIB_DIR, "libmla_preprocess_b200_instr.so")

CACHE_MODE_MAP = {"krope_ctkv": 1, "int8_nzcache": 2, "nzcache": 3}            # op_host/mla_preprocess.cpp:605-606
QUANT_MODE_MAP = {"per_tensor_quant_asymm": 0, "per_token_quant_symm": 1}      # :608-611
DEFAULT_CACHE_MODE = "krope_ctkv"                                              # :634
DEFAULT_QUANT_MODE = "per_token_quant_symm"                                    # :635

EXPORTS = ["mlapo_version", "mlapo_last_error", "mlapo_get_workspace_size", "mlapo_get_workspace_layout",
           "mlapo_repack_nz_weight", "mla_preprocess_launch", "mla_preprocess_launch_static", "mlapo_launches_per_call"]


class MlapoParams(ctypes.Structure):
    _fields_ = [(n, ctypes.c_uint32 if n == "struct_bytes" else ctypes.c_int32) for n in (
        "struct_bytes", "num_tokens", "num_heads", "hidden_dim", "act_dtype", "cache_mode", "quant_mode",
        "num_cache_blocks", "cache_block_size", "descale_is_int64", "flags")] + [
        (n, ctypes.c_void_p) for n in (
            "hidden_state", "wdqkv_km", "descale0", "bias0", "quant_scale0", "quant_offset0", "gamma1", "beta1",
            "quant_scale1", "quant_offset1", "wuq_km", "descale1", "bias1", "gamma2", "cos", "sin", "wuk",
            "slot_mapping", "ctkv_scale", "q_nope_scale", "q_out0", "q_out1", "kv_cache", "kv_cache_rope")]


class MlapoWorkspaceLayout(ctypes.Structure):
    _fields_ = [(n, ctypes.c_size_t) for n in ("total_bytes", "a0", "partial0", "a1", "tok_scale", "y0_tap", "timeline")] + [
        ("splits0", ctypes.c_int32), ("reserved", ctypes.c_int32)]


def _load():
    if not os.path.exists(_LIB_PATH):
        raise ImportError(f"{_LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback for mla_preprocess)")
    lib = ctypes.CDLL(_LIB_PATH)
    lib.mlapo_version.restype = ctypes.c_char_p
    lib.mlapo_last_error.restype = ctypes.c_char_p
    lib.mlapo_get_workspace_size.argtypes = [ctypes.POINTER(MlapoParams), ctypes.POINTER(ctypes.c_size_t)]
    lib.mlapo_get_workspace_layout.argtypes = [ctypes.POINTER(MlapoParams), ctypes.POINTER(MlapoWorkspaceLayout)]
    lib.mlapo_repack_nz_weight.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]
    lib.mlapo_transdata_nd_to_nz.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int32,
                                             ctypes.c_int32, ctypes.c_void_p]
    lib.mlapo_pack_descale_int64.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]
    lib.mlapo_gather_nz_cache_rows.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32,
                                               ctypes.c_int32, ctypes.c_int32, ctypes.c_int64, ctypes.c_void_p]
    lib.batch_matmul_transpose_launch.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int64,
                                                  ctypes.c_int64, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]
    V, I32, I64 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64
    lib.mlapo_alloc_extend.argtypes = [V, V, V, V, I64, V, V, I32, I64, I32, I32, V]
    lib.mlapo_cache_loc_assign.argtypes = [V, I32, V, V, V, V, I32, I64, I32, V]
    lib.mlapo_assign_cache.argtypes = [V, V, V, V, V, V, I32, I64, I32, I32, V]
    lib.mla_preprocess_launch.argtypes = [ctypes.POINTER(MlapoParams), ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]
    lib.mla_preprocess_launch_static.argtypes = [ctypes.POINTER(MlapoParams), ctypes.c_void_p]
    lib.mlapo_launches_per_call.argtypes = [ctypes.POINTER(MlapoParams)]
    lib.mlapo_set_debug_flags.argtypes = [ctypes.c_int]
    # MLAPO_DEBUG_FLAGS: debug flags from the first moment (e.g. the existing tests of the other operators run under the barrier
    # jitter of the instrumented build: tests/test_jitter_gpu.py).  Test and profiling use only, like the flags themselves.
    if os.environ.get("MLAPO_DEBUG_FLAGS"):
        lib.mlapo_set_debug_flags(int(os.environ["MLAPO_DEBUG_FLAGS"], 0))
    return lib


lib = _load()


class MlapoError(RuntimeError):
    pass


def _check(rc, what):
    if rc != 0:
        raise MlapoError(f"{what} failed ({rc}): {lib.mlapo_last_error().decode()}")


def version() -> str:
    return lib.mlapo_version().decode()


def set_debug_flags(flags: int) -> int:
    """bit0: synchronise + check after each kernel; bit1: also write the y0 tap at the end of the workspace."""
    return lib.mlapo_set_debug_flags(flags)


def _stream_ptr(t: torch.Tensor):
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def repack_nz_weight(w_nz: torch.Tensor) -> torch.Tensor:
    """FRACTAL_NZ int8 [1, K/32, n_out, 32] -> K-major [n_out, K] (load-time, once per weight)."""
    if w_nz.dtype != torch.int8 or w_nz.dim() != 4 or w_nz.shape[0] != 1 or w_nz.shape[3] != 32:
        raise ValueError(f"expected int8 FRACTAL_NZ weight [1, K/32, n_out, 32], got {tuple(w_nz.shape)} {w_nz.dtype}")
    if not w_nz.is_cuda or not w_nz.is_contiguous():
        raise ValueError("weight must be a contiguous CUDA tensor")
    k, n_out = w_nz.shape[1] * 32, w_nz.shape[2]
    out = torch.empty((n_out, k), dtype=torch.int8, device=w_nz.device)
    with torch.cuda.device(w_nz.device):
        _check(lib.mlapo_repack_nz_weight(w_nz.data_ptr(), out.data_ptr(), k, n_out, _stream_ptr(w_nz)), "mlapo_repack_nz_weight")
    return out


def _need_cuda(t, what):
    if not t.is_cuda or not t.is_contiguous():
        raise ValueError(f"{what} must be a contiguous CUDA tensor")


def transdata_nd_to_nz(nd: torch.Tensor, c0: int = None) -> torch.Tensor:
    """ND [rows, cols] -> FRACTAL_NZ [1, ceil(cols/c0), ceil16(rows), c0], zero padded (reference test helper `transdata`,
    tests/python/sgl_kernel_npu/test_mla_preprocess.py:38-54).  c0 defaults to 32 for 1-byte and 16 for 2-byte elements."""
    if nd.dim() != 2:
        raise ValueError(f"expected a 2-D tensor, got {tuple(nd.shape)}")
    _need_cuda(nd, "nd")
    eb = nd.element_size()
    c0 = (32 if eb == 1 else 16) if c0 is None else c0
    rows, cols = nd.shape
    out = torch.empty((1, -(-cols // c0), -(-rows // 16) * 16, c0), dtype=nd.dtype, device=nd.device)
    with torch.cuda.device(nd.device):
        _check(lib.mlapo_transdata_nd_to_nz(nd.data_ptr(), out.data_ptr(), rows, cols, c0, eb, _stream_ptr(nd)), "mlapo_transdata_nd_to_nz")
    return out


def pack_descale_int64(descale_f32: torch.Tensor) -> torch.Tensor:
    """fp32 descale -> the int64 carrier the fp16 path expects: (bits & 0xFFFFE000) | 1 << 46 (`trans_descale_param`,
    test_mla_preprocess.py:110-119)."""
    if descale_f32.dtype != torch.float32:
        raise ValueError("descale must be float32")
    _need_cuda(descale_f32, "descale")
    out = torch.empty(descale_f32.shape, dtype=torch.int64, device=descale_f32.device)
    with torch.cuda.device(descale_f32.device):
        _check(lib.mlapo_pack_descale_int64(descale_f32.data_ptr(), out.data_ptr(), descale_f32.numel(), _stream_ptr(descale_f32)),
               "mlapo_pack_descale_int64")
    return out


def gather_nz_cache_rows(cache: torch.Tensor, slots: torch.Tensor) -> torch.Tensor:
    """Rows of a paged NZ cache [blocks, width/c0, 128, c0] named by `slots` (int32) -> dense [n, width]; slot -1 -> zeros
    (`extract_from_nzcache`, test_mla_preprocess.py:122-135)."""
    if cache.dim() != 4 or cache.shape[2] != 128:
        raise ValueError(f"expected an NZ cache [blocks, width/c0, 128, c0], got {tuple(cache.shape)}")
    if slots.dtype != torch.int32 or slots.dim() != 1:
        raise ValueError("slots must be a 1-D int32 tensor")
    _need_cuda(cache, "cache")
    _need_cuda(slots, "slots")
    blocks, groups, _, c0 = cache.shape
    width = groups * c0
    out = torch.empty((slots.numel(), width), dtype=cache.dtype, device=cache.device)
    if slots.numel():
        with torch.cuda.device(cache.device):
            _check(lib.mlapo_gather_nz_cache_rows(cache.data_ptr(), slots.data_ptr(), out.data_ptr(), slots.numel(), width, c0,
                                                  cache.element_size(), blocks, _stream_ptr(cache)), "mlapo_gather_nz_cache_rows")
    return out


BMT_FORMAT_MAP = {"ND": 0, "NZ": 1}                                                      # batch_matmul_transpose.cpp:22-25
BMT_QUANT_MODES = ("per_channel_symm", "per_channel_asymm", "per_token_symm")             # :16-20


def batch_matmul_transpose(tensor_a, tensor_b, tensor_c, format_mode=None, quant_mode=None) -> None:
    """tensor_c[m, batch, :] = tensor_a[m, batch, :] @ tensor_b[batch] through the C ABI; argument meaning, mode strings, defaults
    and checks of csrc/batch_matmul_transpose/op_host/batch_matmul_transpose.cpp:46-80."""
    fmt = "ND" if format_mode is None else format_mode
    qm = "per_channel_symm" if quant_mode is None else quant_mode
    if fmt not in BMT_FORMAT_MAP:
        raise RuntimeError(f"format_mode: Unsupported mode value {fmt}")
    if qm not in BMT_QUANT_MODES:
        raise RuntimeError(f"quant_mode: Unsupported mode value {qm}")
    if not (tensor_a.dtype == tensor_b.dtype == tensor_c.dtype):
        raise RuntimeError("tensor type is not the same")
    if tensor_a.dtype not in (torch.float16, torch.bfloat16):
        raise RuntimeError("tensor type only support half or bf16")
    for t, nm in ((tensor_a, "tensor_a"), (tensor_b, "tensor_b"), (tensor_c, "tensor_c")):
        _need_cuda(t, nm)
    if tensor_a.dim() != 3:
        raise RuntimeError("batch size is not same between srcTensor and dstTensor")
    if fmt == "ND":
        if tensor_b.dim() != 3:
            raise RuntimeError("tensor shape should be dim3 in ND format")
        if tensor_a.shape[2] != tensor_b.shape[1]:
            raise RuntimeError("tensor shape is wrong")
        n = tensor_b.shape[2]
    else:
        if tensor_b.dim() != 4:
            raise RuntimeError("tensor shape should be dim4 in nz format")
        if tensor_a.shape[2] != tensor_b.shape[2]:
            raise RuntimeError("tensor shape is wrong")
        n = tensor_b.shape[1] * tensor_b.shape[3]
    if tensor_a.shape[1] != tensor_b.shape[0]:
        raise RuntimeError("tensor shape is wrong")
    m, batch, k = tensor_a.shape
    if tensor_c.numel() != m * batch * n:
        raise RuntimeError("tensor_c must hold [m, batch, n] elements")
    if tensor_c.numel() == 0:
        return
    with torch.cuda.device(tensor_a.device):
        _check(lib.batch_matmul_transpose_launch(tensor_a.data_ptr(), tensor_b.data_ptr(), tensor_c.data_ptr(), m, batch, k, n,
                                                 (0 if tensor_a.dtype == torch.bfloat16 else 1), BMT_FORMAT_MAP[fmt], _stream_ptr(tensor_a)),
               "batch_matmul_transpose_launch")


def _ptr(t):
    return None if t is None else t.data_ptr()


def build_params(hiddenState, wdqkv_km, descale0, gamma1, beta1, wuq_km, descale1, gamma2, cos, sin, wuk,
                 kv_cache, kv_cache_rope, slotmapping, quant_scale0, quant_offset0, bias0, quant_scale1,
                 quant_offset1, bias1, ctkv_scale, q_nope_scale, cache_mode, quant_mode, q_out0, q_out1, flags=0) -> MlapoParams:
    cache_mode = DEFAULT_CACHE_MODE if cache_mode is None else cache_mode
    quant_mode = DEFAULT_QUANT_MODE if quant_mode is None else quant_mode
    if cache_mode not in CACHE_MODE_MAP:
        raise RuntimeError(f"Unsupported cache_mode value: '{cache_mode}'")
    if quant_mode not in QUANT_MODE_MAP:
        raise RuntimeError(f"Unsupported quant_mode value: '{quant_mode}'")
    act = hiddenState.dtype
    if act not in (torch.bfloat16, torch.float16):
        raise RuntimeError(f"hiddenState must be bfloat16 or float16, got {act}")
    n, d = hiddenState.shape
    h = wuk.shape[0]
    cm = CACHE_MODE_MAP[cache_mode]

    def need(name, t, dtype, shape=None, optional=False):
        if t is None:
            if optional:
                return
            raise RuntimeError(f"{name} is required")
        if not t.is_cuda or t.device != hiddenState.device:
            raise RuntimeError(f"{name} must be on {hiddenState.device}")
        if not t.is_contiguous():
            raise RuntimeError(f"{name} must be contiguous")
        if dtype is not None and t.dtype != dtype:
            raise RuntimeError(f"{name} must be {dtype}, got {t.dtype}")
        if shape is not None and tuple(t.shape) != tuple(shape):
            raise RuntimeError(f"{name} must have shape {tuple(shape)}, got {tuple(t.shape)}")

    need("hiddenState", hiddenState, act)
    need("wdqkv (K-major)", wdqkv_km, torch.int8, (2112, d))
    need("wuq (K-major)", wuq_km, torch.int8, (192 * h, 1536))
    dsc_dt = descale0.dtype
    if dsc_dt not in (torch.float32, torch.int64):
        raise RuntimeError("descale0 must be float32 (bf16 path) or int64 (fp16 path)")
    need("descale0", descale0, dsc_dt, (2112,))
    need("descale1", descale1, dsc_dt, (192 * h,))
    need("gamma1", gamma1, act, (1536,))
    need("beta1", beta1, act, (1536,))
    need("gamma2", gamma2, act, (512,))
    need("cos", cos, act, (n, 64))
    need("sin", sin, act, (n, 64))
    need("wuk", wuk, act, (h, 128, 512))
    need("slotmapping", slotmapping, torch.int32, (n,))
    per_tensor = QUANT_MODE_MAP[quant_mode] == 0
    need("quant_scale0", quant_scale0, act, optional=not per_tensor)
    need("quant_offset0", quant_offset0, torch.int8, optional=not per_tensor)
    need("bias0", bias0, torch.int32, (2112,), optional=not per_tensor)
    need("quant_scale1", quant_scale1, act, optional=not per_tensor)
    need("quant_offset1", quant_offset1, torch.int8, optional=not per_tensor)
    need("bias1", bias1, torch.int32, (192 * h,), optional=not per_tensor)
    q0_dt = torch.int8 if cm == 2 else act
    need("q_out0", q_out0, q0_dt, (n, h, 512))
    need("q_out1", q_out1, act, (n, h, 64))
    need("kv_cache", kv_cache, q0_dt)
    need("kv_cache_rope", kv_cache_rope, act)
    need("ctkv_scale", ctkv_scale, act, optional=cm != 2)
    need("q_nope_scale", q_nope_scale, act, (h,), optional=cm != 2)
    if kv_cache.dim() != 4:
        raise RuntimeError("kv_cache must be 4-D")
    if cm == 1:
        blocks, bs = kv_cache.shape[0], kv_cache.shape[1]
        if kv_cache.shape[2] * kv_cache.shape[3] != 512 or kv_cache_rope.numel() != blocks * bs * 64:
            raise RuntimeError("krope_ctkv caches must be [B,bs,1,512] and [B,bs,1,64]")
    else:
        blocks, bs = kv_cache.shape[0], kv_cache.shape[2]
        c0 = 32 if cm == 2 else 16
        if tuple(kv_cache.shape[1:]) != (512 // c0, 128, c0) or tuple(kv_cache_rope.shape) != (blocks, 4, 128, 16):
            raise RuntimeError(f"NZ caches must be [B,{512 // c0},128,{c0}] and [B,4,128,16] (block size 128)")

    p = MlapoParams()
    p.struct_bytes = ctypes.sizeof(MlapoParams)
    p.num_tokens, p.num_heads, p.hidden_dim = n, h, d
    p.act_dtype = 0 if act == torch.bfloat16 else 1
    p.cache_mode, p.quant_mode = cm, QUANT_MODE_MAP[quant_mode]
    p.num_cache_blocks, p.cache_block_size = blocks, bs
    p.descale_is_int64 = 1 if dsc_dt == torch.int64 else 0
    p.flags = int(flags)
    p.hidden_state, p.wdqkv_km, p.descale0, p.bias0 = _ptr(hiddenState), _ptr(wdqkv_km), _ptr(descale0), _ptr(bias0)
    p.quant_scale0, p.quant_offset0 = _ptr(quant_scale0), _ptr(quant_offset0)
    p.gamma1, p.beta1 = _ptr(gamma1), _ptr(beta1)
    p.quant_scale1, p.quant_offset1 = _ptr(quant_scale1), _ptr(quant_offset1)
    p.wuq_km, p.descale1, p.bias1, p.gamma2 = _ptr(wuq_km), _ptr(descale1), _ptr(bias1), _ptr(gamma2)
    p.cos, p.sin, p.wuk, p.slot_mapping = _ptr(cos), _ptr(sin), _ptr(wuk), _ptr(slotmapping)
    p.ctkv_scale, p.q_nope_scale = _ptr(ctkv_scale), _ptr(q_nope_scale)
    p.q_out0, p.q_out1, p.kv_cache, p.kv_cache_rope = _ptr(q_out0), _ptr(q_out1), _ptr(kv_cache), _ptr(kv_cache_rope)
    return p


def workspace_layout(p: MlapoParams) -> MlapoWorkspaceLayout:
    l = MlapoWorkspaceLayout()
    _check(lib.mlapo_get_workspace_layout(ctypes.byref(p), ctypes.byref(l)), "mlapo_get_workspace_layout")
    return l


def launch(p: MlapoParams, workspace: torch.Tensor, stream=None):
    s = ctypes.c_void_p(stream if stream is not None else torch.cuda.current_stream(workspace.device).cuda_stream)
    _check(lib.mla_preprocess_launch(ctypes.byref(p), workspace.data_ptr(), workspace.numel() * workspace.element_size(), s),
           "mla_preprocess_launch")


FLAG_INPUT_COMPLETE = 1   # MLAPO_FLAG_INPUT_COMPLETE: nothing the call touches is still pending behind an early-triggering (PDL) producer
STATIC_MAX_TOKENS, STATIC_MAX_HIDDEN, STATIC_SCRATCH_BYTES = 256, 8192, 20 << 20   # MLAPO_STATIC_* in the header


def launch_static(p: MlapoParams, device, stream=None):
    """`mla_preprocess_launch_static`: the operator with the library's built-in scratch instead of a caller workspace.

    For hosts that have no way to allocate one (all device memory already pledged to their own pool, no allocator in
    the call path, or a driver allocation is not acceptable inside the request loop).  Up to 256 tokens at hidden <= 8192;
    larger calls raise MlapoError and launch nothing.  All calls on a device share the one scratch array, so keep them on a
    single stream; make the first call per device during warm-up, before capturing anything."""
    s = ctypes.c_void_p(stream if stream is not None else torch.cuda.current_stream(device).cuda_stream)
    with torch.cuda.device(device):
        _check(lib.mla_preprocess_launch_static(ctypes.byref(p), s), "mla_preprocess_launch_static")


def mla_preprocess(hiddenState, gamma0, beta0, wdqkv_km, descale0, gamma1, beta1, wuq_km, descale1, gamma2,
                   cos, sin, wuk, kv_cache, kv_cache_rope, slotmapping,
                   quant_scale0, quant_offset0, bias0, quant_scale1, quant_offset1, bias1, *,
                   ctkv_scale=None, q_nope_scale=None, cache_mode=None, quant_mode=None,
                   q_out0, kv_cache_out0, q_out1, kv_cache_out1, workspace=None, flags=0):
    """Operator-shaped call through the C ABI.  `wdqkv_km` / `wuq_km` are the repacked (K-major) weights.

    gamma0/beta0 are accepted and ignored, as in the reference (bf16.hpp:2762, 424-486).
    kv_cache_out0/1 must alias kv_cache/kv_cache_rope (reference test_mla:575-577).
    Returns (q_out0, kv_cache_out0, q_out1, kv_cache_out1) like op_host/mla_preprocess.cpp:703.
    """
    if kv_cache_out0.data_ptr() != kv_cache.data_ptr() or kv_cache_out1.data_ptr() != kv_cache_rope.data_ptr():
        raise RuntimeError("kv_cache_out0/1 must be the same storage as kv_cache/kv_cache_rope")
    p = build_params(hiddenState, wdqkv_km, descale0, gamma1, beta1, wuq_km, descale1, gamma2, cos, sin, wuk,
                     kv_cache, kv_cache_rope, slotmapping, quant_scale0, quant_offset0, bias0, quant_scale1,
                     quant_offset1, bias1, ctkv_scale, q_nope_scale, cache_mode, quant_mode, q_out0, q_out1, flags=flags)
    with torch.cuda.device(hiddenState.device):
        if workspace is None:
            workspace = torch.empty(max(workspace_layout(p).total_bytes, 256), dtype=torch.uint8, device=hiddenState.device)
        launch(p, workspace)
    return q_out0, kv_cache_out0, q_out1, kv_cache_out1


# ------------------------------------------------------------------------------------------------ slot producers
def _is_i64(t, what):
    if t.dtype not in (torch.int32, torch.int64):
        raise RuntimeError(f"{what} must be int32 or int64")
    return 1 if t.dtype == torch.int64 else 0


def alloc_extend(pre_lens, seq_lens, last_loc, free_pages, page_size, out_indices, values) -> None:
    """Same arguments as torch.ops.npu.alloc_extend (csrc/pytorch_extensions.cpp:31-33): writes the slots of the new tokens
    of every request into out_indices (packed in request order) and the number of pages taken into values[0]."""
    for t, nm in ((pre_lens, "pre_lens"), (seq_lens, "seq_lens"), (last_loc, "last_loc"), (free_pages, "free_pages"),
                  (out_indices, "out_indices"), (values, "values")):
        _need_cuda(t, nm)
    i64 = _is_i64(pre_lens, "pre_lens")
    if not (seq_lens.dtype == last_loc.dtype == free_pages.dtype == values.dtype == pre_lens.dtype):
        raise RuntimeError("pre_lens, seq_lens, last_loc, free_pages and values must share one integer dtype")
    if pre_lens.numel() == 0:
        return
    with torch.cuda.device(pre_lens.device):
        _check(lib.mlapo_alloc_extend(pre_lens.data_ptr(), seq_lens.data_ptr(), last_loc.data_ptr(), free_pages.data_ptr(), int(page_size),
                                      out_indices.data_ptr(), values.data_ptr(), pre_lens.numel(), out_indices.numel(), i64,
                                      _is_i64(out_indices, "out_indices"), _stream_ptr(pre_lens)), "mlapo_alloc_extend")


def _cache_loc(req_indices, token_pool, start_offset, end_offset, out_cache_loc, update):
    ok = (req_indices.dtype in (torch.int32, torch.int64) and token_pool.dtype == torch.int32 and start_offset.dtype == torch.int64
          and end_offset.dtype == torch.int64 and out_cache_loc.dtype == torch.int32)
    if not ok:  # csrc/cache_location_assign/op_host/cache_loc_assign.cpp:73-86
        raise RuntimeError("Only support inputTensor combo1: int64, int32, int64, int64, int32; combo2: int32, int32, int64, int64, int32")
    for t, nm in ((req_indices, "req_indices"), (token_pool, "token_pool"), (start_offset, "start_offset"), (end_offset, "end_offset"),
                  (out_cache_loc, "out_cache_loc")):
        _need_cuda(t, nm)
    if req_indices.numel():
        with torch.cuda.device(token_pool.device):
            _check(lib.mlapo_cache_loc_assign(req_indices.data_ptr(), _is_i64(req_indices, "req_indices"), token_pool.data_ptr(),
                                              start_offset.data_ptr(), end_offset.data_ptr(), out_cache_loc.data_ptr(), req_indices.numel(),
                                              token_pool.shape[1], update, _stream_ptr(token_pool)), "mlapo_cache_loc_assign")


def cache_loc_assign(req_indices, token_pool, start_offset, end_offset, out_cache_loc):
    """token_pool[req[i], start_i:end_i] = out_cache_loc[packed run i]; returns token_pool (cache_loc_assign.cpp:88-100)."""
    _cache_loc(req_indices, token_pool, start_offset, end_offset, out_cache_loc, 0)
    return token_pool


def cache_loc_update(req_indices, token_pool, start_offset, end_offset, out_cache_loc):
    """out_cache_loc[packed run i] = token_pool[req[i], start_i:end_i]; returns out_cache_loc (cache_loc_assign.cpp:102-114)."""
    _cache_loc(req_indices, token_pool, start_offset, end_offset, out_cache_loc, 1)
    return out_cache_loc


def assign_cache_op(out, src, dst_start_idx, dst_end_idx, src_start_idx, src_end_idx) -> bool:
    """out[i, dst_start_i:dst_end_i] = src[src_start_i:src_end_i] (assign_cache.cpp:40-82); False on malformed arguments, as
    the reference."""
    idx = (dst_start_idx, dst_end_idx, src_start_idx, src_end_idx)
    if out.dim() < 2 or src.dim() < 1 or any(t.dim() < 1 for t in idx):
        return False
    if any(t.shape[0] != out.shape[0] for t in idx) or out.dtype != src.dtype:
        return False
    if any(t.dtype != idx[0].dtype for t in idx) or idx[0].dtype not in (torch.int32, torch.int64):
        return False
    for t, nm in ((out, "out"), (src, "src")) + tuple((t, "index tensor") for t in idx):
        _need_cuda(t, nm)
    if out.shape[0] == 0:
        return True
    with torch.cuda.device(out.device):
        _check(lib.mlapo_assign_cache(out.data_ptr(), src.data_ptr(), idx[0].data_ptr(), idx[1].data_ptr(), idx[2].data_ptr(), idx[3].data_ptr(),
                                      out.shape[0], out.stride(0), out.element_size(), _is_i64(idx[0], "index"), _stream_ptr(out)),
               "mlapo_assign_cache")
    return True
