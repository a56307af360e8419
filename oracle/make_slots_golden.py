"""Writes tests/golden/slots_golden.npz from the reference tests' own golden functions (imported from /root/reference with the
NPU modules stubbed): alloc_extend_native (test_alloc_extend_slot.py:88-152) on that file's cases plus seeded multi-request
batches, assign_req_to_token_pool_golden (test_cache_assign.py:42-61) on a scaled-down copy of its __main__ case."""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference/tests/python/sgl_kernel_npu"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "slots_golden.npz")


def load(name):
    for m in ("torch_npu", "sgl_kernel_npu"):
        sys.modules.setdefault(m, types.ModuleType(m))
    spec = importlib.util.spec_from_file_location("ref_" + name, os.path.join(REF, name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    ae = load("test_alloc_extend_slot")
    ca = load("test_cache_assign")
    rec = {}
    # (pre, seq, last_loc, page_size): the reference's five single-request cases, then batches
    cases = [([0], [7], [-1], 128), ([7], [8], [134], 128), ([7], [9], [134], 128), ([0], [300], [-1], 128), ([100], [400], [227], 128)]
    g = torch.Generator().manual_seed(11)
    for bs, ps in ((17, 128), (64, 16), (33, 1)):
        pre = torch.randint(0, 5 * ps + 3, (bs,), generator=g)
        ext = torch.randint(0, 3 * ps + 2, (bs,), generator=g)
        ext[::7] = 0
        last = torch.where(pre > 0, torch.randint(1, 500, (bs,), generator=g) * ps + (pre - 1) % ps, torch.full((bs,), -1))
        cases.append((pre.tolist(), (pre + ext).tolist(), last.tolist(), ps))
    for bs in (5, 40):  # decode batches: every request extends by one token (the reference's fast path)
        pre = torch.randint(1, 700, (bs,), generator=g)
        last = torch.randint(1, 500, (bs,), generator=g) * 128 + (pre - 1) % 128
        cases.append((pre.tolist(), (pre + 1).tolist(), last.tolist(), 128))
    rec["ae_n"] = np.array(len(cases))
    for i, (pre, seq, last, ps) in enumerate(cases):
        pre_t, seq_t, last_t = (torch.tensor(x, dtype=torch.int64) for x in (pre, seq, last))
        free = torch.randperm(4000, generator=g)[:2000].to(torch.int64) + 1
        est = int((((seq_t + ps - 1) // ps) - ((pre_t + ps - 1) // ps)).sum())
        ae.page_size_range = None
        # int32 free pages: the reference's decode fast path index-puts them into an int32 tensor (dtype mismatch with int64)
        out, _ = ae.alloc_extend_native(pre_t, seq_t, last_t, free.to(torch.int32), ps, est)
        rec[f"ae_{i}_pre"], rec[f"ae_{i}_seq"], rec[f"ae_{i}_last"], rec[f"ae_{i}_free"] = pre_t.numpy(), seq_t.numpy(), last_t.numpy(), free.numpy()
        rec[f"ae_{i}_ps"], rec[f"ae_{i}_out"], rec[f"ae_{i}_pages"] = np.array(ps), out.numpy().astype(np.int64), np.array(est)
    # cache_loc_assign golden
    bs, row, pool = 60, 512, 80
    token_pool = torch.arange(row, dtype=torch.int32).repeat(pool, 1)
    start = torch.randint(2, row, (bs,), generator=g, dtype=torch.int64) - 2
    end = torch.minimum(start + torch.randint(1, 4, (bs,), generator=g, dtype=torch.int64), torch.tensor(row))
    req = torch.randperm(pool, generator=g)[:bs].to(torch.int64)
    cache_loc = torch.randint(0, 10000, (int((end - start).sum()),), generator=g, dtype=torch.int32)
    want = token_pool.clone()
    ca.assign_req_to_token_pool_golden(req, want, start, end, cache_loc, bs)
    rec.update(cl_pool=token_pool.numpy(), cl_req=req.numpy(), cl_start=start.numpy(), cl_end=end.numpy(), cl_loc=cache_loc.numpy(), cl_want=want.numpy())
    np.savez_compressed(OUT, **rec)
    print("wrote", OUT, os.path.getsize(OUT))


if __name__ == "__main__":
    main()
