"""`mla_preprocess_launch_static` — the operator call for hosts that cannot provide (or allocate) a workspace.

Who needs it: an engine that gave every byte of device memory to its own pool at start-up, a request loop where a driver
allocation (and the stall it can bring) is not allowed, a plugin interface with no scratch argument.  The library then
works out of one fixed 20 MiB array that is part of the loaded module on each GPU.

What is checked, through the C ABI: two calls in a row (different batches, the second one re-using whatever the first
left in the shared scratch) give byte-identical outputs to `mla_preprocess_launch` with a caller workspace — same
kernels, so the bar is equality, tighter than the tolerances of the oracle tests; oversize requests are refused with a
clear message and launch nothing.  The file sorts last so a stop-at-first-failure run has already run everything else;
it also passes on its own in a fresh process.
"""
import ctypes

import pytest
import torch

from oracle import mla_oracle as orc   # only its input generator is used here
import mla_harness as hz

pytestmark = pytest.mark.gpu
PT = "per_tensor_quant_asymm"
OUTS = ["q_out0", "q_out1", "kv_cache", "kv_cache_rope"]


def _params(cabi, g, case):
    return cabi.build_params(g["hiddenState"], g["wdqkv_km"], g["descale0"], g["gamma1"], g["beta1"], g["wuq_km"], g["descale1"],
                             g["gamma2"], g["cos"], g["sin"], g["wuk"], g["kv_cache"], g["kv_cache_rope"], g["slotmapping"],
                             g["quant_scale0"], g["quant_offset0"], g["bias0"], g["quant_scale1"], g["quant_offset1"], g["bias1"],
                             g["ctkv_scale"], g["q_nope_scale"], case.cache_mode, PT, g["q_out0"], g["q_out1"])


def _device_copy(cabi, case, device):
    g = hz.to_device(case, device)
    g["wdqkv_km"] = cabi.repack_nz_weight(g["wdqkv"])
    g["wuq_km"] = cabi.repack_nz_weight(g["wuq"])
    return g


def test_static_scratch_matches_caller_workspace_twice_in_a_row(cuda_device):
    from sgl_kernel_npu import cabi
    assert cabi.set_debug_flags(0) == 0
    # two different batches; the second is the largest the built-in scratch admits (256 tokens, D = 7168, 8 K-splits)
    for n, h, mode, seed in [(40, 4, "krope_ctkv", 7), (256, 8, "int8_nzcache", 8)]:
        case = orc.gen_case(N=n, H=h, D=7168, dtype=torch.bfloat16, cache_mode=mode, seed=seed, num_blocks=6, neg_slot_frac=0.1)
        ref, got = _device_copy(cabi, case, cuda_device), _device_copy(cabi, case, cuda_device)
        p_ref, p_got = _params(cabi, ref, case), _params(cabi, got, case)
        ws = torch.empty(cabi.workspace_layout(p_ref).total_bytes, dtype=torch.uint8, device=cuda_device)
        cabi.launch(p_ref, ws)                       # existing entry point, caller-owned workspace
        cabi.launch_static(p_got, cuda_device)       # new entry point, no workspace argument, nothing allocated
        torch.cuda.synchronize(cuda_device)
        for k in OUTS:
            assert torch.equal(ref[k], got[k]), f"{k} differs for N={n}"
        assert int(got["q_out1"].float().abs().sum()) > 0   # the call really produced output


def test_static_scratch_rejects_what_does_not_fit(cuda_device):
    from sgl_kernel_npu import cabi
    case = orc.gen_case(N=257, H=1, D=256, dtype=torch.bfloat16, seed=3, num_blocks=4)
    g = _device_copy(cabi, case, cuda_device)
    before = g["q_out1"].clone()
    with pytest.raises(cabi.MlapoError, match=r"257 tokens.*built-in scratch holds 20971520 bytes"):
        cabi.launch_static(_params(cabi, g, case), cuda_device)
    torch.cuda.synchronize(cuda_device)
    assert torch.equal(g["q_out1"], before)          # nothing was launched
    # the refusal is a host-side check: return code MLAPO_ERR_WORKSPACE (-3)
    rc = cabi.lib.mla_preprocess_launch_static(ctypes.byref(_params(cabi, g, case)), None)
    assert rc == -3
