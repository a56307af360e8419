"""GPU timing of batch_matmul_transpose at MLA shapes (not part of bench.py's contract; numbers go to profiles/README.md)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-npu_b200")]
import torch

from sgl_kernel_npu import cabi


def main():
    dev = torch.device("cuda:0")
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    for (m, batch, k, n) in [(36, 128, 512, 128), (256, 128, 512, 128), (1024, 128, 512, 128), (8192, 128, 512, 128)]:
        sets = 6 if m <= 1024 else 2  # rotate weights + activations so that nothing is L2-resident between calls
        a = [torch.randn(m, batch, k, device=dev).bfloat16() for _ in range(sets)]
        w = [torch.randn(batch, k, n, device=dev).bfloat16() for _ in range(sets)]
        c = [torch.empty(m, batch, n, device=dev, dtype=torch.bfloat16) for _ in range(sets)]
        for i in range(sets):
            cabi.batch_matmul_transpose(a[i], w[i], c[i])
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream(dev)
        with torch.cuda.stream(s):
            with torch.cuda.graph(g, stream=s):
                for i in range(sets):
                    cabi.batch_matmul_transpose(a[i], w[i], c[i])
            g.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 50
            e0.record(s)
            for _ in range(reps):
                g.replay()
            e1.record(s)
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (reps * sets)
        byts = 2 * (m * batch * k + batch * k * n + m * batch * n)
        flops = 2.0 * m * batch * k * n
        ref = torch.einsum("mbk,bkn->mbn", a[0].float(), w[0].float())
        err = (c[0].float() - ref).abs().max().item()
        print(json.dumps({"shape": [m, batch, k, n], "us": round(us, 2), "GBps": round(byts / us / 1e3, 1), "frac_hbm": round(byts / us / 1e3 / peak, 3),
                          "TFLOPs": round(flops / us / 1e6, 1), "max_abs_err_vs_fp32": round(err, 4)}))


if __name__ == "__main__":
    main()
