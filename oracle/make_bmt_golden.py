"""Writes tests/golden/bmt_golden.npz: the reference test's own golden expression (torch.bmm into a transposed view,
tests/python/sgl_kernel_npu/test_batch_matmul_transpose.py:18-20) evaluated on this container's CPU for the test's boundary
shapes (:59-72), fixed seeds.  torch_npu is not installed here, so these pin the oracle to torch.bmm, not to the NPU."""
import os

import numpy as np
import torch

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "bmt_golden.npz")
SHAPES = [(1, 1, 1, 1), (1, 10, 1, 1), (10, 1, 1, 10), (5, 5, 1, 5), (2, 2, 2, 1), (100, 1, 1, 100), (1, 100, 100, 1), (2, 3, 4, 5),
          (10, 20, 30, 40), (36, 2, 512, 128), (8, 10, 64, 128)]  # (b=tokens, m=batch, k, n); the two target cases with fewer heads


def compute_golden(a, b, res1, m, n):  # the reference's expression, verbatim
    torch.bmm(a.transpose(0, 1), b, out=res1.view(-1, m, n).transpose(0, 1))


def u16(t):
    return t.contiguous().view(torch.int16).numpy().view(np.uint16)


def main():
    rec = {"shapes": np.array(SHAPES, dtype=np.int64)}
    for i, (b, m, k, n) in enumerate(SHAPES):
        for dt, nm in ((torch.bfloat16, "bf16"), (torch.float16, "fp16")):
            g = torch.Generator().manual_seed(1000 + i)
            a = torch.randn(b, m, k, generator=g).to(dt)
            w = torch.randn(m, k, n, generator=g).to(dt)
            res = torch.empty((b, m * n), dtype=dt)
            compute_golden(a, w, res, m, n)
            rec[f"a_{i}_{nm}"], rec[f"b_{i}_{nm}"], rec[f"c_{i}_{nm}"] = u16(a), u16(w), u16(res.view(b, m, n))
    np.savez_compressed(OUT, **rec)
    print("wrote", OUT, os.path.getsize(OUT))


if __name__ == "__main__":
    main()
