"""Calls launched back to back on one stream with one workspace.  Consecutive calls overlap on the GPU (programmatic dependent
launch: a call's quantise kernel stores its a0 while the previous call's q kernel is still running, its MM1 has its weight tiles in
flight), so what one call leaves in the workspace must never reach another's results: two different inputs alternate, every
call has its own outputs and caches, and each result must equal, bit for bit, what the same input gives when it runs alone."""
import pytest
import torch

from oracle import mla_oracle as orc

pytestmark = pytest.mark.gpu


def _params(cabi, g, case, flags=0):
    return cabi.build_params(g["hiddenState"], g["wdqkv_km"], g["descale0"], g["gamma1"], g["beta1"], g["wuq_km"], g["descale1"],
                             g["gamma2"], g["cos"], g["sin"], g["wuk"], g["kv_cache"], g["kv_cache_rope"], g["slotmapping"],
                             g["quant_scale0"], g["quant_offset0"], g["bias0"], g["quant_scale1"], g["quant_offset1"], g["bias1"],
                             g["ctkv_scale"], g["q_nope_scale"], case.cache_mode, "per_tensor_quant_asymm", g["q_out0"], g["q_out1"], flags=flags)


@pytest.mark.parametrize("flags", [0, 1], ids=["default_start", "input_complete"])
@pytest.mark.parametrize("shape", [(256, 128, 7168, "krope_ctkv"), (64, 16, 7168, "int8_nzcache"), (1100, 40, 7168, "nzcache"),
                                   (8, 16, 7168, "krope_ctkv")],
                         ids=lambda s: "n%d_h%d_d%d_%s" % s)
def test_alternating_inputs_back_to_back(shape, flags, cuda_device):
    """flags = MLAPO_FLAG_INPUT_COMPLETE: the quantise kernel starts under the previous call's tail (what bench.py times); 0: every
    call waits for the previous kernel first.  Either way each result must equal the stand-alone one."""
    import mla_harness as hz
    from sgl_kernel_npu import cabi
    n, h, d, mode = shape
    cases = [orc.gen_case(N=n, H=h, D=d, dtype=torch.bfloat16, cache_mode=mode, seed=s, num_blocks=max(12, n // 64)) for s in (11, 12)]
    # what each input gives alone
    alone = []
    for c in cases:
        g, _ = hz.run_cuda(c, cuda_device, debug=False, cache_fill=0)
        alone.append({k: g[k].clone() for k in ("q_out0", "q_out1", "kv_cache", "kv_cache_rope")})
    rounds = 6
    runs = []
    for r in range(2 * rounds):
        c = cases[r % 2]
        g = hz.to_device(c, cuda_device)
        g["kv_cache"].zero_()
        g["kv_cache_rope"].zero_()
        g["wdqkv_km"] = cabi.repack_nz_weight(g["wdqkv"])
        g["wuq_km"] = cabi.repack_nz_weight(g["wuq"])
        runs.append((g, _params(cabi, g, c, flags)))
    lay = cabi.workspace_layout(runs[0][1])
    ws = torch.zeros(lay.total_bytes, dtype=torch.uint8, device=cuda_device)
    torch.cuda.synchronize(cuda_device)
    for _, p in runs:  # no synchronisation in between
        cabi.launch(p, ws)
    torch.cuda.synchronize(cuda_device)
    for r, (g, _) in enumerate(runs):
        for k, want in alone[r % 2].items():
            assert torch.equal(g[k], want), (r, k, int((g[k] != want).sum()))


@pytest.mark.parametrize("n_tokens", [64, 8], ids=["n64", "n8"])
def test_waits_for_a_producer_that_releases_its_dependents_early(n_tokens, cuda_device):
    """hiddenState (and cos / sin / slot_mapping) are written by a kernel that triggers programmatic launch completion FIRST and
    stores ~100 us LATER — what a PDL-enabled norm kernel in front of this operator may do.  With the default flags the operator's
    first kernel waits for that producer to complete before it reads anything, so the results equal those of a call on finished
    inputs; before that wait existed the quantiser read the buffer while it still held the previous contents."""
    import ctypes, os
    import mla_harness as hz
    from sgl_kernel_npu import cabi
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "microbench", "libmlapo_microbench.so")
    if not os.path.exists(path):
        pytest.skip("tools/microbench/libmlapo_microbench.so not built")
    mb = ctypes.CDLL(path)
    mb.mlapo_mb_late_copy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_longlong, ctypes.c_void_p]
    case = orc.gen_case(N=n_tokens, H=16, D=7168, dtype=torch.bfloat16, cache_mode="krope_ctkv", seed=21, num_blocks=8, neg_slot_frac=0.1)
    g_ref, _ = hz.run_cuda(case, cuda_device, debug=False, cache_fill=0)
    want = {k: g_ref[k].clone() for k in ("q_out0", "q_out1", "kv_cache", "kv_cache_rope")}
    g = hz.to_device(case, cuda_device)
    g["kv_cache"].zero_()
    g["kv_cache_rope"].zero_()
    g["wdqkv_km"], g["wuq_km"] = cabi.repack_nz_weight(g["wdqkv"]), cabi.repack_nz_weight(g["wuq"])
    late = {}
    for k in ("hiddenState", "cos", "sin", "slotmapping"):
        late[k] = g[k].clone()
        g[k].fill_(1)                     # stale contents the operator must not see
    p = _params(cabi, g, case, flags=0)
    ws = torch.zeros(cabi.workspace_layout(p).total_bytes, dtype=torch.uint8, device=cuda_device)
    torch.cuda.synchronize(cuda_device)
    st = torch.cuda.current_stream(cuda_device).cuda_stream
    keys = ["cos", "sin", "slotmapping", "hiddenState"]   # the operator's direct predecessor on the stream writes hiddenState
    for k in keys:
        nbytes = g[k].numel() * g[k].element_size()
        assert nbytes % 16 == 0
        assert mb.mlapo_mb_late_copy(g[k].data_ptr(), late[k].data_ptr(), nbytes // 16, 200000, st) == 0
    cabi.launch(p, ws)
    torch.cuda.synchronize(cuda_device)
    for k, w in want.items():
        assert torch.equal(g[k], w), (k, int((g[k] != w).sum()))
