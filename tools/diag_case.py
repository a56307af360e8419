"""GPU diagnostic: one case through the harness with chosen debug flags; prints per-stage mismatch stats."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-npu_b200"), os.path.join(ROOT, "tests")]
import torch
from oracle import mla_oracle as orc
import mla_harness as hz
from sgl_kernel_npu import cabi

n, h, d, dt, mode, qm, flags = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], sys.argv[5], sys.argv[6], int(sys.argv[7])
DT = {"bf16": torch.bfloat16, "fp16": torch.float16}
case = orc.gen_case(N=n, H=h, D=d, dtype=DT[dt], cache_mode=mode, seed=42, num_blocks=max(12, n // 64))
dev = torch.device("cuda:0")
orig = cabi.set_debug_flags
cabi.set_debug_flags = lambda f: orig(f | flags)   # keep the extra flags on while the harness sets its own
g, taps = hz.run_cuda(case, dev, quant_mode=qm)
cabi.set_debug_flags = orig
res, ocase, otaps = hz.compare(case, qm, g, taps)
print(sys.argv[1:], {k: v for k, v in res.items()})
bad = (g["q_out1"].cpu() != ocase.t["q_out1"]).nonzero()
print("q_out1 mismatches at (token, head, col):", bad[:8].tolist())
for key in ("kv_cache", "kv_cache_rope"):
    a, b = g[key].cpu().reshape(-1), ocase.t[key].reshape(-1)
    idx = (a != b).nonzero().reshape(-1)
    print(key, "mismatches:", [(int(i), float(a[i]), float(b[i])) for i in idx[:8]])
