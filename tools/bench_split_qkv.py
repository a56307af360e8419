"""GPU timing of split_qkv_rmsnorm_rope (not part of bench.py's contract; numbers go to profiles/README.md).
Algorithmic bytes per token: 2 * (q_hidden + 2 kv_hidden) * 2 (read the fused row, write q, k, v) + 2 * rope_dim * 2 (sin, cos)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-npu_b200")]
import torch

from sgl_kernel_npu.norm.split_qkv_rmsnorm_rope import split_qkv_rmsnorm_rope


def main():
    dev = torch.device("cuda:0")
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peak = json.load(open(pk))["hbm_gbs"] if os.path.exists(pk) else 6650.0
    dt = torch.bfloat16
    # (batch, q_hidden, kv_hidden, head_dim, rope_dim): the reference test's shape, then the same model at serving batch sizes
    for (b, qh, kvh, hd, rd) in [(12, 6144, 1024, 128, 128), (256, 6144, 1024, 128, 128), (2048, 6144, 1024, 128, 64),
                                 (8192, 6144, 1024, 128, 128), (16384, 4096, 1024, 128, 128), (8192, 8192, 1024, 256, 64)]:
        per_call = 2 * b * (qh + 2 * kvh) * 2 + 4 * b * rd
        sets = max(2, min(64, int(300e6 // per_call) + 1))  # rotate buffers over > 126 MB so that nothing is L2-resident
        qkv = [torch.randn(b, qh + 2 * kvh, device=dev).to(dt) for _ in range(sets)]
        sin, cos = torch.rand(b, rd, device=dev).to(dt), torch.rand(b, rd, device=dev).to(dt)
        w = [torch.randn(hd, device=dev).to(dt) for _ in range(4)]
        call = lambda i: split_qkv_rmsnorm_rope(qkv[i], sin, cos, qh, kvh, hd, eps=1e-6, q_weight=w[0], k_weight=w[1], q_bias=w[2], k_bias=w[3])
        for i in range(sets):
            call(i)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream(dev)
        with torch.cuda.stream(s):
            outs = []  # kept alive, so that every call of the graph writes its own output buffers
            with torch.cuda.graph(g, stream=s):
                for i in range(sets):
                    outs.append(call(i))
            g.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 20
            e0.record(s)
            for _ in range(reps):
                g.replay()
            e1.record(s)
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (reps * sets)
        print(json.dumps({"shape": [b, qh, kvh, hd, rd], "us": round(us, 2), "GBps": round(per_call / us / 1e3, 1),
                          "frac_hbm": round(per_call / us / 1e3 / peak, 3), "Mtok_s": round(b / us, 2)}))


if __name__ == "__main__":
    main()
