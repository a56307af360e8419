"""GPU parity tests: CUDA path (called through the C ABI) vs the CPU oracle, stage by stage.

Bars (BASELINE.json north_star): bit-exact for int8 activations (a0, a1), int32 accumulators
(acc0), the dequantised rows y0, q_rope, both caches incl. slot placement and untouched rows;
q_out0 (bf16 tensor-core absorb) within 1e-2 relative (int8 q_out0: |diff| <= 1).
"""
import pytest
import torch

from oracle import mla_oracle as orc
import mla_harness as hz

pytestmark = pytest.mark.gpu
DT = {"bf16": torch.bfloat16, "fp16": torch.float16}
PT, TOK = "per_tensor_quant_asymm", "per_token_quant_symm"

# (N, H, D, dtype, cache_mode, quant_mode, neg_slot_frac)
CASES = [
    # the reference's own shapes (test_mla:487-498) x its dtype/cache-mode cases (:746-808)
    (1, 32, 7168, "bf16", "krope_ctkv", PT, 0.0),
    (16, 64, 7168, "bf16", "int8_nzcache", PT, 0.0),
    (31, 128, 7168, "bf16", "nzcache", PT, 0.0),
    (31, 128, 6144, "bf16", "krope_ctkv", PT, 0.0),
    (16, 32, 7168, "fp16", "krope_ctkv", PT, 0.0),
    # BASELINE C1 (TP8 slice) and beyond-reference cases: slot -1, ragged N, M-tiling, unpinned per-token mode
    (8, 16, 7168, "bf16", "krope_ctkv", PT, 0.0),
    (129, 16, 7168, "bf16", "nzcache", PT, 0.1),
    (300, 8, 7168, "bf16", "int8_nzcache", PT, 0.1),
    (64, 4, 7168, "bf16", "krope_ctkv", TOK, 0.1),
    (33, 4, 6144, "fp16", "nzcache", TOK, 0.0),
    # tile-boundary sizes and the ends of the reference's hiddenSize range [2048, 8192] (csrc/mla_preprocess/README.md:8)
    (128, 3, 2048, "bf16", "krope_ctkv", PT, 0.0),
    (257, 5, 8192, "bf16", "int8_nzcache", PT, 0.02),
    (256, 1, 7168, "fp16", "nzcache", PT, 0.0),
    # two token tiles per CTA and >= 96 (head, 256-token) units in one wave (the decode configuration of the bench)
    (130, 96, 6144, "bf16", "int8_nzcache", PT, 0.05),
    (256, 100, 7168, "fp16", "krope_ctkv", TOK, 0.0),
    (200, 128, 7168, "bf16", "nzcache", PT, 0.0),
    # prefill-sized: no split-K, 256-wide MM1 tiles, many token super-tiles (the reference stops at 1024 tokens)
    (2400, 2, 7168, "bf16", "krope_ctkv", PT, 0.05),
    # more (head, 256-token) units than SMs: the q path runs in its streaming form (persistent CTAs, overlapped units)
    (2400, 16, 7168, "bf16", "nzcache", PT, 0.05),
    (1300, 32, 6144, "bf16", "int8_nzcache", PT, 0.0),
    (1100, 40, 7168, "fp16", "krope_ctkv", TOK, 0.0),
]


def check(res, case, quant_mode, q0_close):
    exact = ["a0", "acc0", "y0", "a1", "q_out1", "kv_cache", "kv_cache_rope"]
    for k in exact:
        if quant_mode == TOK and k in ("q_out1", "kv_cache"):
            # per-token mode is parity-unpinned (no reference test): allow isolated 1-ulp differences
            assert res[k][0] <= 1e-5 * res[k][1] and res[k][2] <= 1e-3, f"{k}: {res[k]}"
        else:
            assert res[k][0] == 0, f"{k}: {res[k]}"
    if case.cache_mode == "int8_nzcache":
        assert res["q_out0"][2] <= 1 and res["q_out0"][0] <= 2e-3 * res["q_out0"][1], res["q_out0"]
    else:
        # north_star: within 1e-2 relative; the absolute floor is the reference test's own atol (test_mla:712-720),
        # which is what bounds outputs whose 128 products nearly cancel
        assert q0_close, res["q_out0_relerr"]
        assert res["q_out0"][0] <= 2e-3 * res["q_out0"][1], res["q_out0"]  # and almost always bit-identical


VARIANTS = {"chained": (False, 0), "persistent": (True, 0), "single_cta_q": (False, 128), "pair_q": (False, 8192)}


@pytest.mark.parametrize("variant", list(VARIANTS))
@pytest.mark.parametrize("cfg", CASES, ids=[f"N{c[0]}-H{c[1]}-D{c[2]}-{c[3]}-{c[4]}-{c[5][:9]}" for c in CASES])
def test_parity_vs_oracle(cfg, variant, cuda_device):
    """`chained` = the default launch shape (four PDL-chained kernels; the q path is the per-head kernel, or the CTA-pair
    kernel when there are more (head, 256-token) units than SMs); `persistent` = the opt-in single cooperative kernel for
    calls of <= 256 tokens; `single_cta_q` = the streaming q kernel the pair kernel replaced; `pair_q` = the pair kernel
    forced on every call of more than 128 tokens.  All meet the same bars."""
    n, h, d, dt, mode, qm, neg = cfg
    persistent, extra = VARIANTS[variant]
    if persistent and n > 256:
        pytest.skip("the persistent decode kernel covers <= 256 tokens")
    many_units = n > 256 and h * ((n + 255) // 256) > torch.cuda.get_device_properties(cuda_device).multi_processor_count
    if variant == "single_cta_q" and not many_units:
        pytest.skip("same kernels as `chained` for this shape")
    if variant == "pair_q" and (n <= 128 or many_units):
        pytest.skip("same kernels as `chained` for this shape")
    case = orc.gen_case(N=n, H=h, D=d, dtype=DT[dt], cache_mode=mode, seed=42, num_blocks=max(12, n // 64), neg_slot_frac=neg)
    case.t["kv_cache"].fill_(5)
    case.t["kv_cache_rope"].fill_(5)
    g, taps = hz.run_cuda(case, cuda_device, quant_mode=qm, persistent=persistent, extra_flags=extra)
    res, ocase, _ = hz.compare(case, qm, g, taps)
    q0_close = mode == "int8_nzcache" or torch.allclose(g["q_out0"].cpu().float(), ocase.t["q_out0"].float(), rtol=1e-2, atol=1e-3)
    check(res, case, qm, q0_close)


def test_golden_fixture_on_gpu(cuda_device):
    """CUDA output vs the reference golden fixture itself (not via the oracle)."""
    import glob, os
    import numpy as np
    path = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "golden_N31_H32_D7168_bf16_nzcache.npz")))[0]
    gz = np.load(path)
    case = orc.gen_case(N=31, H=32, D=7168, dtype=torch.bfloat16, cache_mode="nzcache", seed=42)
    g, _ = hz.run_cuda(case, cuda_device, debug=False)
    heads = gz["heads"].tolist()
    as_bf16 = lambda a: torch.from_numpy(a.view(np.int16).copy()).view(torch.bfloat16)
    assert torch.equal(g["q_out1"].cpu()[:, heads], as_bf16(gz["q_pe"]))
    kn, kp = hz.extract_cache_rows(case, g["kv_cache"].cpu(), g["kv_cache_rope"].cpu())
    assert torch.equal(torch.stack(kp).unsqueeze(1), as_bf16(gz["k_pe"]))
    assert torch.equal(torch.stack(kn).unsqueeze(1), as_bf16(gz["k_nope"]))
    # reference tolerance for q_nope: allclose(atol=rtol=1e-3) (test_mla:712-720)
    assert torch.allclose(g["q_out0"].cpu()[:, heads].float(), as_bf16(gz["q_nope"]).float(), rtol=1e-2, atol=1e-3)


def test_repack_is_exact_permutation(cuda_device):
    from sgl_kernel_npu import cabi
    w = torch.randint(-128, 128, (384, 1536), dtype=torch.int8)
    nz = orc.nz_pack(w).to(cuda_device)
    assert torch.equal(cabi.repack_nz_weight(nz).cpu(), w)


def test_full_range_weights_and_idempotence(cuda_device):
    """int32 accumulators stay exact with full-range int8 weights; a second call gives identical bytes."""
    case = orc.gen_case(N=40, H=2, D=7168, seed=7, num_blocks=4, full_range_weights=True)
    g, taps = hz.run_cuda(case, cuda_device)
    res, _, _ = hz.compare(case, PT, g, taps)
    assert res["acc0"][0] == 0 and res["a1"][0] == 0 and res["q_out1"][0] == 0
    g2, _ = hz.run_cuda(case, cuda_device)
    for k in ["q_out0", "q_out1", "kv_cache", "kv_cache_rope"]:
        assert torch.equal(g[k], g2[k])


def test_baseline_size_properties(cuda_device):
    """BASELINE C2 size (N=256, H=128): too big for the oracle's bmm in seconds on every head, so check
    size-independent properties: token-permutation equivariance and head-slice (TP) consistency."""
    case = orc.gen_case(N=256, H=128, D=7168, seed=11, num_blocks=8)
    g, _ = hz.run_cuda(case, cuda_device, debug=False)
    perm = torch.randperm(256)
    c2 = orc.MlaCase(256, 128, 7168, case.dtype, case.cache_mode, dict(case.t))
    for k in ["hiddenState", "cos", "sin", "slotmapping"]:
        c2.t[k] = case.t[k][perm].contiguous()
    g2, _ = hz.run_cuda(c2, cuda_device, debug=False)
    assert torch.equal(g2["q_out0"].cpu(), g["q_out0"].cpu()[perm])
    assert torch.equal(g2["q_out1"].cpu(), g["q_out1"].cpu()[perm])
    assert torch.equal(g2["kv_cache"], g["kv_cache"]) and torch.equal(g2["kv_cache_rope"], g["kv_cache_rope"])
    # TP slice: heads 32..47 computed alone must equal the same heads of the full call (C4 sharding)
    sl = slice(32, 48)
    c3 = orc.MlaCase(256, 16, 7168, case.dtype, case.cache_mode, dict(case.t))
    c3.t["wuq"] = orc.nz_pack(case.t["wuq_nd"][32 * 192:48 * 192])
    c3.t["descale1"] = case.t["descale1"][32 * 192:48 * 192].contiguous()
    c3.t["bias1"] = case.t["bias1"][32 * 192:48 * 192].contiguous()
    c3.t["wuk"] = case.t["wuk"][sl].contiguous()
    c3.t["q_nope_scale"] = case.t["q_nope_scale"][sl].contiguous()
    c3.t["q_out0"] = torch.zeros(256, 16, 512, dtype=case.dtype)
    c3.t["q_out1"] = torch.zeros(256, 16, 64, dtype=case.dtype)
    g3, _ = hz.run_cuda(c3, cuda_device, debug=False)
    assert torch.equal(g3["q_out0"].cpu(), g["q_out0"].cpu()[:, sl])
    assert torch.equal(g3["q_out1"].cpu(), g["q_out1"].cpu()[:, sl])
    # and a sampled oracle check of 4 tokens x all heads at this size
    sub = orc.MlaCase(4, 128, 7168, case.dtype, case.cache_mode, dict(case.t))
    idx = torch.tensor([0, 100, 128, 255])
    for k in ["hiddenState", "cos", "sin", "slotmapping"]:
        sub.t[k] = case.t[k][idx].contiguous()
    sub.t["q_out0"] = torch.zeros(4, 128, 512, dtype=case.dtype)
    sub.t["q_out1"] = torch.zeros(4, 128, 64, dtype=case.dtype)
    sub.t["kv_cache"] = torch.zeros_like(case.t["kv_cache"])
    sub.t["kv_cache_rope"] = torch.zeros_like(case.t["kv_cache_rope"])
    orc.run_case(sub, quant_mode=PT)
    assert torch.equal(g["q_out1"].cpu()[idx], sub.t["q_out1"])
    rel = ((g["q_out0"].cpu()[idx].float() - sub.t["q_out0"].float()).abs() / sub.t["q_out0"].float().abs().clamp_min(1e-3)).max()
    assert rel <= 1e-2


def test_empty_batch_is_noop(cuda_device):
    from sgl_kernel_npu import cabi
    case = orc.gen_case(N=1, H=2, D=256, seed=1, num_blocks=2)
    g = hz.to_device(case, cuda_device)
    km0, km1 = cabi.repack_nz_weight(g["wdqkv"]), cabi.repack_nz_weight(g["wuq"])
    e = lambda *s, dt=torch.bfloat16: torch.zeros(*s, dtype=dt, device=cuda_device)
    cabi.mla_preprocess(e(0, 256), None, None, km0, g["descale0"], g["gamma1"], g["beta1"], km1, g["descale1"], g["gamma2"],
                        e(0, 64), e(0, 64), g["wuk"], g["kv_cache"], g["kv_cache_rope"], e(0, dt=torch.int32),
                        g["quant_scale0"], g["quant_offset0"], g["bias0"], g["quant_scale1"], g["quant_offset1"], g["bias1"],
                        cache_mode="krope_ctkv", quant_mode=PT, q_out0=e(0, 2, 512), kv_cache_out0=g["kv_cache"],
                        q_out1=e(0, 2, 64), kv_cache_out1=g["kv_cache_rope"])
    torch.cuda.synchronize()
    assert int(g["kv_cache"].abs().sum()) == 0
