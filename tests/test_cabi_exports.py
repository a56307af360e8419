"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/mla_preprocess_b200.h declares, and its host-only entry points (parameter validation,
workspace sizing) behave.  No kernel is launched here."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "mla_preprocess_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.findall(r"^\s*(?:const\s+char\s*\*|int)\s+(\w+)\s*\(", src, flags=re.M)


def test_header_declares_expected_entry_points():
    syms = declared_symbols()
    for s in ["mla_preprocess_launch", "mlapo_get_workspace_size", "mlapo_get_workspace_layout",
              "mlapo_repack_nz_weight", "mlapo_version", "mlapo_last_error", "mlapo_launches_per_call"]:
        assert s in syms


def test_library_exports_every_declared_symbol():
    from sgl_kernel_npu import cabi
    for s in declared_symbols():
        assert hasattr(cabi.lib, s), f"{s} declared in the header but not exported"
    assert "sm_100a" in cabi.version()


def test_struct_size_matches_and_validation_errors():
    from sgl_kernel_npu import cabi
    p = cabi.MlapoParams()
    sz = ctypes.c_size_t()
    # wrong struct size -> rejected with a message
    p.struct_bytes = 8
    assert cabi.lib.mlapo_get_workspace_size(ctypes.byref(p), ctypes.byref(sz)) == -1
    assert b"struct_bytes" in cabi.lib.mlapo_last_error()
    p.struct_bytes = ctypes.sizeof(cabi.MlapoParams)
    p.num_tokens, p.num_heads, p.hidden_dim = 256, 128, 7168
    p.act_dtype, p.cache_mode, p.quant_mode = 0, 1, 0
    p.num_cache_blocks, p.cache_block_size = 16, 128
    assert cabi.lib.mlapo_get_workspace_size(ctypes.byref(p), ctypes.byref(sz)) == 0
    lay = cabi.workspace_layout(p)
    assert lay.total_bytes == sz.value and lay.splits0 >= 1
    assert lay.a0 % 256 == 0 and lay.partial0 >= 256 * 7168 and lay.tok_scale > lay.a1 > lay.partial0
    # empty batch sizes fine on the host side (no division by the tile count)
    p.num_tokens = 0
    assert cabi.lib.mlapo_get_workspace_size(ctypes.byref(p), ctypes.byref(sz)) == 0
    p.num_tokens = 256
    # bad modes / geometry
    p.cache_mode = 7
    assert cabi.lib.mlapo_get_workspace_size(ctypes.byref(p), ctypes.byref(sz)) == -1
    assert b"Unsupported cache_mode" in cabi.lib.mlapo_last_error()
    p.cache_mode, p.hidden_dim = 1, 7000
    assert cabi.lib.mlapo_get_workspace_size(ctypes.byref(p), ctypes.byref(sz)) == -1
    p.hidden_dim, p.cache_mode, p.cache_block_size = 7168, 3, 64
    assert cabi.lib.mlapo_get_workspace_size(ctypes.byref(p), ctypes.byref(sz)) == -1
    assert b"128" in cabi.lib.mlapo_last_error()


def test_python_mirror_rejects_bad_mode_strings_like_the_reference():
    import torch
    from sgl_kernel_npu import cabi
    x = torch.zeros(1, 128, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="Unsupported cache_mode value: 'bogus'"):
        cabi.build_params(x, *([None] * 20), None, "bogus", None, None, None)
    with pytest.raises(RuntimeError, match="Unsupported quant_mode value: 'nope'"):
        cabi.build_params(x, *([None] * 20), None, "krope_ctkv", "nope", None, None)
    assert cabi.DEFAULT_QUANT_MODE == "per_token_quant_symm" and cabi.DEFAULT_CACHE_MODE == "krope_ctkv"
