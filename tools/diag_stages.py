"""GPU diagnostic: per-stage mismatch counts of the CUDA path against the oracle (not a test)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-npu_b200"), os.path.join(ROOT, "tests")]
import torch

from oracle import mla_oracle as orc
import mla_harness as hz

CASES = [
    (1, 2, 256, "bf16", "krope_ctkv", "per_tensor_quant_asymm"),
    (5, 2, 7168, "bf16", "krope_ctkv", "per_tensor_quant_asymm"),
    (31, 32, 7168, "bf16", "nzcache", "per_tensor_quant_asymm"),
    (130, 4, 6144, "bf16", "int8_nzcache", "per_tensor_quant_asymm"),
    (16, 8, 7168, "fp16", "krope_ctkv", "per_tensor_quant_asymm"),
    (9, 3, 1024, "bf16", "krope_ctkv", "per_token_quant_symm"),
]
DT = {"bf16": torch.bfloat16, "fp16": torch.float16}

if __name__ == "__main__":
    dev = torch.device("cuda:0")
    sel = [int(a) for a in sys.argv[1:]] or range(len(CASES))
    for i in sel:
        n, h, d, dt, mode, qm = CASES[i]
        case = orc.gen_case(N=n, H=h, D=d, dtype=DT[dt], cache_mode=mode, seed=42 + i, num_blocks=8,
                            neg_slot_frac=0.2 if n > 4 else 0.0)
        t0 = time.time()
        try:
            g, taps = hz.run_cuda(case, dev, quant_mode=qm, cache_fill=3)
        except Exception as e:  # noqa
            print(f"CASE {CASES[i]}: LAUNCH FAILED: {e}", flush=True)
            break
        case.t["kv_cache"].fill_(3)
        case.t["kv_cache_rope"].fill_(3)
        res, _, _ = hz.compare(case, qm, g, taps)
        print(f"CASE {CASES[i]} splits0={taps['splits0']} ({time.time() - t0:.1f}s)")
        for k, v in res.items():
            print(f"   {k:14s} {v}", flush=True)
