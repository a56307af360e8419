"""Slot producers: alloc_extend, cache_loc_assign / cache_loc_update, assign_cache_op.  Integer work, so every
comparison is bit-exact.  CPU: the numpy oracle against fixtures made by the reference tests' own golden functions
(oracle/make_slots_golden.py).  GPU: the CUDA ops (C ABI through ctypes, and torch.ops.npu.*) against both."""
import os

import numpy as np
import pytest
import torch

from oracle import slots_oracle as so

GOLD = os.path.join(os.path.dirname(__file__), "golden", "slots_golden.npz")


def ae_cases():
    z = np.load(GOLD)
    for i in range(int(z["ae_n"])):
        yield {k: z[f"ae_{i}_{k}"] for k in ("pre", "seq", "last", "free", "ps", "out", "pages")}


# ------------------------------------------------------------------------------------------------ oracle vs reference (CPU)
def test_oracle_alloc_extend_matches_reference_golden():
    n = 0
    for c in ae_cases():
        out, pages = so.alloc_extend(c["pre"], c["seq"], c["last"], c["free"], int(c["ps"]))
        assert np.array_equal(out, c["out"]) and pages == int(c["pages"])
        n += 1
    assert n >= 10


def test_oracle_cache_loc_matches_reference_golden():
    z = np.load(GOLD)
    got = so.cache_loc_assign(z["cl_req"], z["cl_pool"], z["cl_start"], z["cl_end"], z["cl_loc"])
    assert np.array_equal(got, z["cl_want"])
    # update is the inverse copy: reading the assigned pool back returns the packed locations
    back = so.cache_loc_update(z["cl_req"], got, z["cl_start"], z["cl_end"], np.zeros_like(z["cl_loc"]))
    assert np.array_equal(back, z["cl_loc"])


def test_oracle_assign_cache():
    out = np.zeros((3, 10), dtype=np.int16)
    src = np.arange(100, 120, dtype=np.int16)
    got = so.assign_cache(out, src, [1, 0, 7], [4, 2, 10], [0, 3, 5], [3, 5, 8])
    assert got[0].tolist() == [0, 100, 101, 102, 0, 0, 0, 0, 0, 0]
    assert got[1].tolist() == [103, 104, 0, 0, 0, 0, 0, 0, 0, 0]
    assert got[2].tolist() == [0, 0, 0, 0, 0, 0, 0, 105, 106, 107]


# ------------------------------------------------------------------------------------------------ CUDA
@pytest.mark.gpu
@pytest.mark.parametrize("idt", [torch.int64, torch.int32])
@pytest.mark.parametrize("via", ["cabi", "torch_op"])
def test_alloc_extend_gpu(idt, via, cuda_device):
    import sgl_kernel_npu  # noqa: F401
    from sgl_kernel_npu import cabi
    fn = cabi.alloc_extend if via == "cabi" else torch.ops.npu.alloc_extend
    for c in ae_cases():
        t = lambda a: torch.from_numpy(a).to(idt).to(cuda_device)
        out = torch.full((len(c["out"]) + 3,), -7, dtype=idt, device=cuda_device)  # three guard elements behind the result
        values = torch.full((1,), -1, dtype=idt, device=cuda_device)
        fn(t(c["pre"]), t(c["seq"]), t(c["last"]), t(c["free"]), int(c["ps"]), out[:len(c["out"])], values)
        assert np.array_equal(out.cpu().numpy()[:len(c["out"])].astype(np.int64), c["out"])
        assert out[len(c["out"]):].tolist() == [-7, -7, -7]
        assert int(values.item()) == int(c["pages"])
    # a large ragged batch against the oracle (int32 slots out of int64 lengths when idt is int64)
    g = torch.Generator().manual_seed(5)
    bs, ps = 700, 64
    pre = torch.randint(0, 900, (bs,), generator=g)
    seq = pre + torch.randint(0, 300, (bs,), generator=g)
    last = torch.where(pre > 0, torch.randint(1, 3000, (bs,), generator=g) * ps + (pre - 1) % ps, torch.full((bs,), -1))
    free = torch.randperm(9000, generator=g)[:4000] + 1
    want, pages = so.alloc_extend(pre.numpy(), seq.numpy(), last.numpy(), free.numpy(), ps)
    out = torch.empty(len(want), dtype=torch.int32 if idt == torch.int64 else idt, device=cuda_device)
    values = torch.zeros(1, dtype=idt, device=cuda_device)
    fn(pre.to(idt).to(cuda_device), seq.to(idt).to(cuda_device), last.to(idt).to(cuda_device), free.to(idt).to(cuda_device), ps, out, values)
    assert np.array_equal(out.cpu().numpy().astype(np.int64), want) and int(values.item()) == pages


@pytest.mark.gpu
@pytest.mark.parametrize("rdt", [torch.int64, torch.int32])
def test_cache_loc_assign_and_update_gpu(rdt, cuda_device):
    import sgl_kernel_npu  # noqa: F401
    from sgl_kernel_npu import cabi
    z = np.load(GOLD)
    dev = lambda a, dt=None: torch.from_numpy(a).to(dt or torch.from_numpy(a).dtype).to(cuda_device)
    for assign, update in ((cabi.cache_loc_assign, cabi.cache_loc_update), (torch.ops.npu.cache_loc_assign, torch.ops.npu.cache_loc_update)):
        pool = dev(z["cl_pool"])
        ret = assign(dev(z["cl_req"], rdt), pool, dev(z["cl_start"]), dev(z["cl_end"]), dev(z["cl_loc"]))
        assert ret.data_ptr() == pool.data_ptr()
        assert np.array_equal(pool.cpu().numpy(), z["cl_want"])
        loc = torch.zeros_like(dev(z["cl_loc"]))
        ret = update(dev(z["cl_req"], rdt), pool, dev(z["cl_start"]), dev(z["cl_end"]), loc)
        assert ret.data_ptr() == loc.data_ptr()
        assert np.array_equal(loc.cpu().numpy(), z["cl_loc"])
    with pytest.raises(RuntimeError, match="Only support inputTensor combo1"):
        cabi.cache_loc_assign(dev(z["cl_req"], rdt), dev(z["cl_pool"]).long(), dev(z["cl_start"]), dev(z["cl_end"]), dev(z["cl_loc"]))


@pytest.mark.gpu
@pytest.mark.parametrize("dt", [torch.int8, torch.int16, torch.int32, torch.int64])
def test_assign_cache_op_gpu(dt, cuda_device):
    """The reference test's shapes (test_inplace_assign_cache.py:74-121): token gaps 1, 2 and 50000, every integer width."""
    import sgl_kernel_npu  # noqa: F401
    from sgl_kernel_npu import cabi
    g = torch.Generator().manual_seed(9)
    for gap, seq_len, bs in ((1, 1024, 48), (2, 1024, 48), (50000, 100000, 6)):
        start = torch.randint(0, seq_len - gap - 1, (bs,), generator=g)
        end = start + gap
        pool = torch.randint(0, 32, (bs, seq_len), generator=g).to(dt)
        src = torch.randint(0, 32, (bs * gap,), generator=g).to(dt)
        cum = torch.cumsum(end - start, 0)
        s0 = torch.cat((torch.zeros(1, dtype=torch.int64), cum))[:bs].contiguous()
        want = so.assign_cache(pool.numpy(), src.numpy(), start.numpy(), end.numpy(), s0.numpy(), cum.numpy())
        for fn in (cabi.assign_cache_op, torch.ops.npu.assign_cache_op):
            out = pool.clone().to(cuda_device)
            ok = fn(out, src.to(cuda_device), start.to(cuda_device), end.to(cuda_device), s0.to(cuda_device), cum.to(cuda_device))
            assert ok is True
            assert np.array_equal(out.cpu().numpy(), want)
    # malformed arguments are refused with False, as in the reference host (assign_cache.cpp:40-58)
    out = torch.zeros(4, 8, dtype=dt, device=cuda_device)
    idx = torch.zeros(3, dtype=torch.int64, device=cuda_device)
    assert cabi.assign_cache_op(out, out.reshape(-1), idx, idx, idx, idx) is False
