"""One shape of split_qkv_tp_rmsnorm_rope on one rank (tp_world = 1), a few calls: the target of an ncu capture."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-npu_b200")]
import torch

from sgl_kernel_npu.norm.split_qkv_tp_rmsnorm_rope import split_qkv_tp_rmsnorm_rope

b, qh, kvh, hd, rd = 8192, 6144, 1024, 128, 128
dev = torch.device("cuda:0")
dt = torch.bfloat16
qkv = torch.randn(b, qh + 2 * kvh, device=dev).to(dt)
sin, cos = torch.rand(b, rd, device=dev).to(dt), torch.rand(b, rd, device=dev).to(dt)
qw, kw = torch.randn(qh, device=dev).to(dt), torch.randn(kvh, device=dev).to(dt)
for _ in range(3):
    out = split_qkv_tp_rmsnorm_rope(qkv, cos, sin, qh, kvh, hd, 1e-6, qw, kw, rd, 1, None)
torch.cuda.synchronize()
print("ok", out[0].float().abs().mean().item())
