"""One shape of split_qkv_rmsnorm_rope, a few calls: the target of an ncu capture (tools/bench_split_qkv.py times it)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-npu_b200")]
import torch

from sgl_kernel_npu.norm.split_qkv_rmsnorm_rope import split_qkv_rmsnorm_rope

b, qh, kvh, hd, rd = (int(x) for x in (sys.argv[1:6] if len(sys.argv) > 5 else (8192, 6144, 1024, 128, 128)))
dev = torch.device("cuda:0")
dt = torch.bfloat16
qkv = torch.randn(b, qh + 2 * kvh, device=dev).to(dt)
sin, cos = torch.rand(b, rd, device=dev).to(dt), torch.rand(b, rd, device=dev).to(dt)
w = [torch.randn(hd, device=dev).to(dt) for _ in range(4)]
for _ in range(3):
    out = split_qkv_rmsnorm_rope(qkv, sin, cos, qh, kvh, hd, eps=1e-6, q_weight=w[0], k_weight=w[1], q_bias=w[2], k_bias=w[3])
torch.cuda.synchronize()
print("ok", out[0].float().abs().mean().item())
