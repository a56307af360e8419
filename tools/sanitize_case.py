"""Small calls of every launch shape, for compute-sanitizer (one tool per gpurun call):
  compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_case.py
Covers: the decode chain (C1 size), the one-token-tile and two-token-tile MM1 / q_fused shapes, the prefill shapes (dequantising MM1,
y0-reading norm), the CTA-pair q kernel, int8 outputs, per-token mode, calls back to back on one workspace under programmatic
dependent launch (MLAPO_FLAG_INPUT_COMPLETE), batch_matmul_transpose and one split_qkv shape.  Results are checked against the
oracle where it is cheap, so a sanitizer run is also a (slow) parity run."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-npu_b200"), os.path.join(ROOT, "tests")]
import torch

from oracle import mla_oracle as orc
import mla_harness as hz
from sgl_kernel_npu import cabi

dev = torch.device("cuda:0")
PT, TOK = "per_tensor_quant_asymm", "per_token_quant_symm"
CASES = [  # N, H, D, cache_mode, quant_mode, debug flags, params flags
    (8, 16, 7168, "krope_ctkv", PT, 0, 0),          # C1: decode chain, one token tile
    (8, 16, 7168, "krope_ctkv", PT, 0, 1),          # ... started under MLAPO_FLAG_INPUT_COMPLETE
    (130, 4, 2048, "int8_nzcache", PT, 0, 0),       # two token tiles, int8 outputs
    (130, 4, 2048, "nzcache", PT, 8192, 0),         # the CTA-pair q kernel
    (40, 3, 2048, "krope_ctkv", TOK, 0, 0),         # per-token mode
    (1100, 2, 2048, "krope_ctkv", PT, 0, 0),        # prefill shapes: dequantising MM1, norm on act rows
    (1100, 2, 2048, "int8_nzcache", PT, 8192, 0),   # ... with the CTA-pair q kernel
]
for n, h, d, mode, qm, dbg, fl in CASES:
    case = orc.gen_case(N=n, H=h, D=d, dtype=torch.bfloat16, cache_mode=mode, seed=5, num_blocks=max(12, n // 64), neg_slot_frac=0.1)
    g, taps = hz.run_cuda(case, dev, quant_mode=qm, debug="tap", extra_flags=dbg, flags=fl)
    res, _, _ = hz.compare(case, qm, g, taps)
    bad = {k: v for k, v in res.items() if k in ("a0", "acc0", "y0", "a1", "kv_cache_rope") and v[0] != 0}
    print(f"case N={n} H={h} D={d} {mode} {qm} dbg={dbg} flags={fl}: mismatching stages {bad or 'none'}; q_out0 relerr {res['q_out0_relerr']:.2e}", flush=True)

# back to back on one workspace, no synchronisation in between
case = orc.gen_case(N=64, H=4, D=2048, dtype=torch.bfloat16, cache_mode="krope_ctkv", seed=9, num_blocks=8)
g = hz.to_device(case, dev)
g["wdqkv_km"], g["wuq_km"] = cabi.repack_nz_weight(g["wdqkv"]), cabi.repack_nz_weight(g["wuq"])
p = cabi.build_params(g["hiddenState"], g["wdqkv_km"], g["descale0"], g["gamma1"], g["beta1"], g["wuq_km"], g["descale1"], g["gamma2"], g["cos"],
                      g["sin"], g["wuk"], g["kv_cache"], g["kv_cache_rope"], g["slotmapping"], g["quant_scale0"], g["quant_offset0"], g["bias0"],
                      g["quant_scale1"], g["quant_offset1"], g["bias1"], g["ctkv_scale"], g["q_nope_scale"], case.cache_mode, PT, g["q_out0"],
                      g["q_out1"], flags=1)
ws = torch.zeros(cabi.workspace_layout(p).total_bytes, dtype=torch.uint8, device=dev)
for _ in range(3):
    cabi.launch(p, ws)
torch.cuda.synchronize()
print("back to back: done", flush=True)

a = torch.randn(36, 8, 512, device=dev).to(torch.bfloat16)
b = torch.randn(8, 512, 128, device=dev).to(torch.bfloat16)
c = torch.empty(36, 8, 128, device=dev, dtype=torch.bfloat16)
cabi.batch_matmul_transpose(a, b, c)
torch.cuda.synchronize()
print("batch_matmul_transpose: max |c - bmm| =", float((c.float() - torch.bmm(a.transpose(0, 1).float(), b.float()).transpose(0, 1)).abs().max()), flush=True)

from sgl_kernel_npu.norm.split_qkv_rmsnorm_rope import split_qkv_rmsnorm_rope
qkv = torch.randn(12, 4096 + 2 * 1024, device=dev).to(torch.bfloat16)
sin, cos = torch.rand(12, 128, device=dev).to(torch.bfloat16), torch.rand(12, 128, device=dev).to(torch.bfloat16)
w = torch.randn(128, device=dev).to(torch.bfloat16)
try:
    out = split_qkv_rmsnorm_rope(qkv, sin, cos, 4096, 1024, 128, eps=1e-6, q_weight=w, k_weight=w)
    torch.cuda.synchronize()
    print("split_qkv_rmsnorm_rope: ok", [tuple(t.shape) for t in out], flush=True)
except TypeError as e:
    print("split_qkv_rmsnorm_rope: signature differs,", e)
print("sanitize_case: finished", flush=True)
