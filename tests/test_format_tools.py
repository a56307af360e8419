"""Format tools: oracle restatements pinned to the reference's own helpers (CPU), CUDA kernels vs both (GPU)."""
import os

import numpy as np
import pytest
import torch

from oracle import mla_oracle as orc

GOLD = os.path.join(os.path.dirname(__file__), "golden", "format_tools.npz")


def _gold():
    return np.load(GOLD)


# ------------------------------------------------------------------------------------------------ oracle vs reference (CPU)
def test_oracle_nz_pack_matches_reference_transdata():
    g = _gold()
    assert np.array_equal(orc.nz_pack(torch.from_numpy(g["td_i8_in"]), 32).numpy(), g["td_i8_out"])
    assert np.array_equal(orc.nz_pack(torch.from_numpy(g["td_16_in"]), 16).numpy(), g["td_16_out"])


def test_oracle_descale_packing_matches_reference():
    g = _gold()
    assert np.array_equal(orc.pack_descale_fp16_path(g["descale_in"]).numpy(), g["descale_out"])
    # and the carrier decodes to the truncated fp32 the fp16 path multiplies by (test_mla:101-103)
    dec = orc.unpack_descale(torch.from_numpy(g["descale_out"])).numpy()
    assert np.array_equal(dec.view(np.uint32), g["descale_in"].view(np.uint32) & np.uint32(0xFFFFE000))


def test_oracle_cache_reader_matches_reference():
    g = _gold()
    cache = torch.from_numpy(g["nz_cache"])
    for s, want in zip(g["nz_slots"], g["nz_rows"]):
        assert np.array_equal(orc.read_cache_row(cache, int(s), 64, 16, True).numpy(), want)
    cache8 = torch.from_numpy(g["nz_cache8"])
    for s, want in zip(g["nz_slots"][:5], g["nz_rows8"]):
        assert np.array_equal(orc.read_cache_row(cache8, int(s), 64, 32, True).numpy(), want)


# ------------------------------------------------------------------------------------------------ CUDA vs golden and oracle
@pytest.mark.gpu
def test_transdata_gpu(cuda_device):
    from sgl_kernel_npu import cabi
    g = _gold()
    for key, c0 in (("td_i8", 32), ("td_16", 16)):
        out = cabi.transdata_nd_to_nz(torch.from_numpy(g[key + "_in"]).to(cuda_device), c0)
        assert np.array_equal(out.cpu().numpy(), g[key + "_out"])
    # ragged shapes (rows padded to 16, columns to c0, zero filled) and the operator's real weight geometry, against the oracle
    gen = torch.Generator().manual_seed(3)
    for shape, dt, c0 in (((50, 100), torch.int8, 32), ((33, 40), torch.int16, 16), ((1, 1), torch.int8, 32), ((2112, 7168), torch.int8, 32)):
        x = torch.randint(-100, 100, shape, dtype=dt, generator=gen)
        out = cabi.transdata_nd_to_nz(x.to(cuda_device), c0)
        assert torch.equal(out.cpu(), orc.nz_pack(x, c0))
    # bf16 payloads travel bit-exactly and the result feeds the repack the operator uses
    w = torch.randint(-128, 128, (192 * 4, 1536), dtype=torch.int8, generator=gen)
    nz = cabi.transdata_nd_to_nz(w.to(cuda_device))
    assert torch.equal(cabi.repack_nz_weight(nz).cpu(), w)


@pytest.mark.gpu
def test_pack_descale_gpu(cuda_device):
    from sgl_kernel_npu import cabi
    g = _gold()
    out = cabi.pack_descale_int64(torch.from_numpy(g["descale_in"]).to(cuda_device))
    assert np.array_equal(out.cpu().numpy(), g["descale_out"])
    x = torch.randn(4097) * 1e-3
    assert torch.equal(cabi.pack_descale_int64(x.to(cuda_device)).cpu(), orc.pack_descale_fp16_path(x.numpy()))


@pytest.mark.gpu
def test_gather_nz_cache_rows_gpu(cuda_device):
    from sgl_kernel_npu import cabi
    g = _gold()
    rows = cabi.gather_nz_cache_rows(torch.from_numpy(g["nz_cache"]).to(cuda_device), torch.from_numpy(g["nz_slots"]).to(cuda_device))
    assert np.array_equal(rows.cpu().numpy(), g["nz_rows"])
    rows8 = cabi.gather_nz_cache_rows(torch.from_numpy(g["nz_cache8"]).to(cuda_device), torch.from_numpy(g["nz_slots"][:5].copy()).to(cuda_device))
    assert np.array_equal(rows8.cpu().numpy(), g["nz_rows8"])
    # the operator's own cache geometry, with skipped (-1) and out-of-range slots
    gen = torch.Generator().manual_seed(5)
    cache = torch.randint(-2 ** 15, 2 ** 15, (4, 512 // 16, 128, 16), dtype=torch.int16, generator=gen)
    slots = torch.tensor([5, -1, 511, 300, 4 * 128], dtype=torch.int32)
    got = cabi.gather_nz_cache_rows(cache.to(cuda_device), slots.to(cuda_device)).cpu()
    for i, s in enumerate(slots.tolist()):
        want = orc.read_cache_row(cache, s, 512, 16, True) if 0 <= s < 4 * 128 else torch.zeros(512, dtype=torch.int16)
        assert torch.equal(got[i], want)
