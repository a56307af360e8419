"""batch_matmul_transpose.  Bars: the oracle (float64 sum, one rounding) agrees with the reference test's own
golden expression (torch.bmm, fixtures made by oracle/make_bmt_golden.py) up to isolated one-ulp roundings; the CUDA op is within
one output ulp (+ the fp32 accumulation bound, see ulp_report) of the oracle everywhere and bit-identical on >= 99 % of the
elements — fp32 accumulation order is the only freedom (the reference test, :36-56, allows 1e-5 of the largest output against
torch.bmm on its own device)."""
import os

import numpy as np
import pytest
import torch

from oracle import bmt_oracle as bo

GOLD = os.path.join(os.path.dirname(__file__), "golden", "bmt_golden.npz")
DT = {"bf16": torch.bfloat16, "fp16": torch.float16}


def _t(arr, dt):
    return torch.from_numpy(arr.view(np.int16).copy()).view(dt)


def ulp_report(got, want, a=None, b=None):
    """(worst error in units of the allowance, fraction of bit-identical elements).  Allowance per element: one output ulp of the
    wanted value (a rounding that falls the other way) plus the fp32 accumulation bound 2^-21 * sum_k |a_k b_k| (outputs whose
    products nearly cancel cannot be held to an ulp of their own size by ANY fp32-accumulating implementation)."""
    if want.numel() == 0:
        return 0.0, 1.0
    mant = 7 if want.dtype == torch.bfloat16 else 10
    g, w = got.double(), want.double()
    allow = torch.pow(2.0, torch.floor(torch.log2(w.abs().clamp_min(2.0 ** -24))) - mant)
    if a is not None:
        allow = allow + torch.einsum("mbk,bkn->mbn", a.double().abs(), b.double().abs()) * 2.0 ** -21
    return float(((g - w).abs() / allow).max()), float((got == want).double().mean())


def test_oracle_matches_reference_golden_expression():
    z = np.load(GOLD)
    for i, (b, m, k, n) in enumerate(z["shapes"].tolist()):
        for nm, dt in DT.items():
            a, w, c = _t(z[f"a_{i}_{nm}"], dt), _t(z[f"b_{i}_{nm}"], dt), _t(z[f"c_{i}_{nm}"], dt)
            got = bo.batch_matmul_transpose(a, w)
            worst, same = ulp_report(got, c, a, w)
            assert worst <= 1.0 and same >= 0.99, (b, m, k, n, nm, worst, same)


def test_oracle_nz_round_trip():
    g = torch.Generator().manual_seed(0)
    w = torch.randn(3, 32, 48, generator=g).bfloat16()
    assert torch.equal(bo.nz_unpack_b(bo.nz_pack_b(w)), w)
    a = torch.randn(5, 3, 32, generator=g).bfloat16()
    assert torch.equal(bo.batch_matmul_transpose(a, bo.nz_pack_b(w), "NZ"), bo.batch_matmul_transpose(a, w, "ND"))


CASES = [  # (m = tokens, batch = heads, k, n): the reference's boundary list (:59-72), its target cases, DSV3 decode and a long batch
    (1, 1, 1, 1), (1, 10, 1, 1), (10, 1, 1, 10), (5, 5, 1, 5), (2, 2, 2, 1), (100, 1, 1, 100), (1, 100, 100, 1), (2, 3, 4, 5), (10, 20, 30, 40),
    (36, 128, 512, 128), (8, 160, 512, 128), (256, 128, 512, 128), (300, 4, 512, 128), (1500, 2, 128, 64), (129, 3, 64, 192), (77, 5, 333, 19),
]


@pytest.mark.gpu
@pytest.mark.parametrize("dt", list(DT))
@pytest.mark.parametrize("shape", CASES, ids=[f"m{a}-b{b}-k{c}-n{d}" for a, b, c, d in CASES])
def test_cuda_vs_oracle(shape, dt, cuda_device):
    from sgl_kernel_npu import cabi
    m, batch, k, n = shape
    g = torch.Generator().manual_seed(m * 7 + batch * 3 + k + n)
    a = torch.randn(m, batch, k, generator=g).to(DT[dt])
    w = torch.randn(batch, k, n, generator=g).to(DT[dt])
    want = bo.batch_matmul_transpose(a, w)
    c = torch.full((m, batch, n), float("nan"), dtype=DT[dt], device=cuda_device)
    cabi.batch_matmul_transpose(a.to(cuda_device), w.to(cuda_device), c)
    worst, same = ulp_report(c.cpu(), want, a, w)
    assert worst <= 1.0 and same >= 0.99, (worst, same)
    if k % 16 == 0 and n % 16 == 0:  # NZ weights [batch, n/16, k, 16]
        c2 = torch.full_like(c, float("nan"))
        cabi.batch_matmul_transpose(a.to(cuda_device), bo.nz_pack_b(w).to(cuda_device), c2, format_mode="NZ")
        worst, same = ulp_report(c2.cpu(), want, a, w)
        assert worst <= 1.0 and same >= 0.99, (worst, same)


@pytest.mark.gpu
def test_cuda_vs_golden_fixture_and_zero(cuda_device):
    """The committed torch.bmm fixtures themselves, and the reference's all-zero case (:119-135)."""
    from sgl_kernel_npu import cabi
    z = np.load(GOLD)
    for i, (b, m, k, n) in enumerate(z["shapes"].tolist()):
        for nm, dt in DT.items():
            a, w, want = _t(z[f"a_{i}_{nm}"], dt), _t(z[f"b_{i}_{nm}"], dt), _t(z[f"c_{i}_{nm}"], dt)
            c = torch.empty((b, m, n), dtype=dt, device=cuda_device)
            cabi.batch_matmul_transpose(a.to(cuda_device), w.to(cuda_device), c)
            worst, same = ulp_report(c.cpu(), want, a, w)
            assert worst <= 1.0 and same >= 0.98, (b, m, k, n, nm, worst, same)
    for dt in DT.values():
        c = torch.full((5, 4, 2), 3.0, dtype=dt, device=cuda_device)
        cabi.batch_matmul_transpose(torch.zeros(5, 4, 3, dtype=dt, device=cuda_device), torch.zeros(4, 3, 2, dtype=dt, device=cuda_device), c)
        assert torch.all(c == 0)


@pytest.mark.gpu
def test_errors_and_torch_op(cuda_device):
    """Same refusals as the reference host (batch_matmul_transpose.cpp:40-80), and the torch.ops.npu entry point."""
    import sgl_kernel_npu  # noqa: F401  (registers torch.ops.npu.*)
    from sgl_kernel_npu import cabi
    a = torch.randn(4, 3, 64, device=cuda_device).bfloat16()
    w = torch.randn(3, 64, 64, device=cuda_device).bfloat16()
    c = torch.empty(4, 3, 64, device=cuda_device).bfloat16()
    for call in (cabi.batch_matmul_transpose, torch.ops.npu.batch_matmul_transpose):
        with pytest.raises(RuntimeError, match="Unsupported mode value"):
            call(a, w, c, format_mode="FRACTAL")
        with pytest.raises(RuntimeError, match="Unsupported mode value"):
            call(a, w, c, quant_mode="per_tensor")
        with pytest.raises(RuntimeError, match="tensor type is not the same"):
            call(a, w.half(), c)
        with pytest.raises(RuntimeError, match="tensor shape is wrong"):
            call(a, w[:2].contiguous(), c)
        with pytest.raises(RuntimeError, match="dim4 in nz format"):
            call(a, w, c, format_mode="NZ")
    torch.ops.npu.batch_matmul_transpose(a, w, c)
    worst, same = ulp_report(c.cpu(), bo.batch_matmul_transpose(a.cpu(), w.cpu()), a.cpu(), w.cpu())
    assert worst <= 1.0 and same >= 0.99
