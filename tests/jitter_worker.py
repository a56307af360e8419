"""Worker of tests/test_jitter_gpu.py: runs in its own process with MLAPO_LIB pointing at the instrumented build of the library
(-DMLAPO_INSTRUMENTED; the shipped library carries neither the jitter nor the delay hook).  Exit code 0 = every disturbed result
equals the undisturbed one."""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for _p in (HERE, ROOT, os.path.join(ROOT, "sgl-kernel-npu_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)
import torch  # noqa: E402

from oracle import mla_oracle as orc  # noqa: E402
import mla_harness as hz  # noqa: E402

BF, FP = torch.bfloat16, torch.float16
PT = "per_tensor_quant_asymm"
JITTER = 2097152
LATE_DRAIN = 1048576

SHAPES = [
    # N, H, D, dtype, cache mode, extra debug flags, what it walks
    (8, 16, 7168, BF, "krope_ctkv", 0),            # mm1<1,128> split-K 8, q_fused kSplit 4
    (128, 64, 7168, BF, "int8_nzcache", 0),        # q_fused kSplit 2, int8 drains
    (200, 128, 7168, FP, "nzcache", 0),            # mm1_pair<128>, q_fused one CTA per head, two token tiles
    (512, 16, 7168, BF, "krope_ctkv", 0),          # mm1_pair<256>, q_fused 128-token CTAs, kSplit 2
    (300, 8, 7168, BF, "nzcache", 8192),           # CTA-pair q kernel, one unit per pair
    (1300, 32, 6144, BF, "int8_nzcache", 8192),    # mm1_dequant + warp-per-row norm + pair kernel, 192 units on 74 pairs, int8
    (2400, 16, 7168, FP, "krope_ctkv", 8192),      # the same with act drains, 160 units
]


def _outputs(g, taps):
    out = {k: g[k].cpu() for k in ("q_out0", "q_out1", "kv_cache", "kv_cache_rope")}
    out.update({k: taps[k] for k in ("a0", "acc0", "a1")})
    return out


def check_shape(shape, cuda_device, disturb=JITTER, reps=4):
    n, h, d, dt, mode, extra = shape
    case = orc.gen_case(N=n, H=h, D=d, dtype=dt, cache_mode=mode, seed=11, num_blocks=max(12, n // 64), neg_slot_frac=0.05)
    case.t["kv_cache"].fill_(3)
    case.t["kv_cache_rope"].fill_(3)
    g, taps = hz.run_cuda(case, cuda_device, quant_mode=PT, debug="tap", extra_flags=extra)
    want = _outputs(g, taps)
    for rep in range(reps):
        g2, taps2 = hz.run_cuda(case, cuda_device, quant_mode=PT, debug="tap", extra_flags=extra | disturb)
        got = _outputs(g2, taps2)
        for k, v in want.items():
            assert torch.equal(got[k], v), f"{k} differs under disturbance {disturb} (repetition {rep})"


def check_chained(cuda_device):
    """Eight calls back to back on ONE workspace with the no-wait start (MLAPO_FLAG_INPUT_COMPLETE), alternating two inputs: the
    quantiser of call k+1 runs under the q kernel of call k.  Every result must equal the stand-alone one."""
    from sgl_kernel_npu import cabi
    n, h, d = 256, 32, 7168
    cases = [orc.gen_case(N=n, H=h, D=d, dtype=BF, cache_mode="krope_ctkv", seed=s, num_blocks=12, neg_slot_frac=0.0) for s in (5, 6)]
    alone = []
    for c in cases:
        g, _ = hz.run_cuda(c, cuda_device, quant_mode=PT, debug=False)
        alone.append((g["q_out0"].clone(), g["q_out1"].clone()))
    gs = [hz.to_device(c, cuda_device) for c in cases]
    params, outs = [], []
    for g, c in zip(gs, cases):
        g["wdqkv_km"] = cabi.repack_nz_weight(g["wdqkv"])
        g["wuq_km"] = cabi.repack_nz_weight(g["wuq"])
    old = cabi.set_debug_flags(JITTER)
    try:
        for k in range(8):
            g, c = gs[k & 1], cases[k & 1]
            o0, o1 = torch.zeros_like(g["q_out0"]), torch.zeros_like(g["q_out1"])
            outs.append((o0, o1))
            params.append(cabi.build_params(g["hiddenState"], g["wdqkv_km"], g["descale0"], g["gamma1"], g["beta1"], g["wuq_km"],
                                            g["descale1"], g["gamma2"], g["cos"], g["sin"], g["wuk"], g["kv_cache"], g["kv_cache_rope"],
                                            g["slotmapping"], g["quant_scale0"], g["quant_offset0"], g["bias0"], g["quant_scale1"],
                                            g["quant_offset1"], g["bias1"], g["ctkv_scale"], g["q_nope_scale"], c.cache_mode, PT, o0, o1,
                                            flags=cabi.FLAG_INPUT_COMPLETE))
        ws = torch.empty(cabi.workspace_layout(params[0]).total_bytes, dtype=torch.uint8, device=cuda_device)
        torch.cuda.synchronize(cuda_device)
        for p in params:
            cabi.launch(p, ws)
        torch.cuda.synchronize(cuda_device)
    finally:
        cabi.set_debug_flags(old)
    for k, (o0, o1) in enumerate(outs):
        assert torch.equal(o0, alone[k & 1][0]) and torch.equal(o1, alone[k & 1][1]), f"call {k} differs from the stand-alone result"


def main():
    from sgl_kernel_npu import cabi
    assert cabi.lib.mlapo_debug_is_instrumented() == 1, "this worker needs the -DMLAPO_INSTRUMENTED build (MLAPO_LIB)"
    dev = torch.device("cuda", 0)
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "chained"):
        check_chained(dev)
        print("chained ok", flush=True)
    for i, shape in enumerate(SHAPES):
        if which in ("all", str(i)):
            check_shape(shape, dev)
            print("shape", i, shape[:2], "ok", flush=True)
    if which in ("all", "late_drain"):
        for shape in SHAPES[-2:]:   # several units per pair, int8 and act drains
            check_shape(shape, dev, disturb=LATE_DRAIN, reps=1)
            print("late drain groups", shape[:2], "ok", flush=True)


if __name__ == "__main__":
    main()
