"""Pin the oracle (oracle/mla_oracle.py) to the reference's own golden.

Fixtures in tests/golden/ were produced by oracle/make_golden.py from the reference's
unmodified generator + golden2_pytorch (test_mla_preprocess.py:140-321, :407-484).
CPU-only; runs in `-m "not gpu"`.
"""
import glob
import hashlib
import os

import numpy as np
import pytest
import torch

from oracle import mla_oracle as orc

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "golden_*.npz")))
DT = {"bf16": torch.bfloat16, "fp16": torch.float16}


def sha(t):
    return hashlib.sha256(t.detach().contiguous().view(torch.uint8).numpy().tobytes()).hexdigest()


def from_np(a, dtype):
    if dtype in (torch.bfloat16, torch.float16):
        return torch.from_numpy(a.view(np.int16).copy()).view(dtype)
    return torch.from_numpy(a.copy())


def parse(path):
    base = os.path.basename(path)[len("golden_"):-len(".npz")]
    n, h, d, dt, mode = base.split("_", 4)
    return int(n[1:]), int(h[1:]), int(d[1:]), dt, mode


def extract_rows(case, width_nope=512):
    """Read back each token's cache rows through the layout reader (test_mla:580-609)."""
    t = case.t
    nz = case.cache_mode != "krope_ctkv"
    c0 = 32 if case.cache_mode == "int8_nzcache" else 16
    kn, kp = [], []
    for s in t["slotmapping"].tolist():
        kn.append(orc.read_cache_row(t["kv_cache"], s, 512, c0, nz))
        kp.append(orc.read_cache_row(t["kv_cache_rope"], s, 64, 16, nz))
    return torch.stack(kn).unsqueeze(1), torch.stack(kp).unsqueeze(1)


def test_fixtures_present():
    assert len(GOLDEN) >= 8


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_generator_reproduces_reference_inputs(path):
    n, h, d, dt, mode = parse(path)
    g = np.load(path)
    case = orc.gen_case(N=n, H=h, D=d, dtype=DT[dt], cache_mode=mode, seed=42)
    for key in g.files:
        if key.startswith("sha_"):
            assert sha(case.t[key[4:]]) == str(g[key]), key
    # NZ packing is a pure permutation of the ND weights
    assert torch.equal(orc.nz_unpack(case.t["wdqkv"]), case.t["wdqkv_nd"])
    assert torch.equal(orc.nz_unpack(case.t["wuq"]), case.t["wuq_nd"])
    assert tuple(case.t["wdqkv"].shape) == (1, d // 32, 2112, 32)


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_oracle_matches_reference_golden(path):
    n, h, d, dt, mode = parse(path)
    act = DT[dt]
    g = np.load(path)
    heads = g["heads"].tolist()
    q0_dt = torch.int8 if mode == "int8_nzcache" else act
    ref_q_nope = from_np(g["q_nope"], q0_dt)
    ref_q_pe = from_np(g["q_pe"], act)
    ref_k_nope = from_np(g["k_nope"], q0_dt)
    ref_k_pe = from_np(g["k_pe"], act)

    # (1) same reduction as the golden -> every elementwise stage must be bit-identical
    case = orc.gen_case(N=n, H=h, D=d, dtype=act, cache_mode=mode, seed=42)
    orc.run_case(case, var_mode="torch")
    k_nope, k_pe = extract_rows(case)
    assert sha(case.t["q_out1"]) == str(g["q_pe_sha"])          # full q_rope tensor, all heads
    assert torch.equal(case.t["q_out1"][:, heads], ref_q_pe)
    assert torch.equal(k_pe, ref_k_pe)
    assert torch.equal(k_nope, ref_k_nope)
    # BMM: oracle = correctly rounded dot product, golden = torch bf16 bmm -> within 1 ulp(act)
    q0 = case.t["q_out0"][:, heads]
    if q0_dt == torch.int8:
        diff = (q0.int() - ref_q_nope.int()).abs()
        assert diff.max().item() <= 1 and (diff != 0).float().mean().item() < 5e-3
    else:
        assert torch.allclose(q0.float(), ref_q_nope.float(), rtol=1e-2, atol=1e-3)
        assert (q0 != ref_q_nope).float().mean().item() < 0.05

    # (2) order-free reduction (the CUDA contract): inside the reference's own tolerance
    case2 = orc.gen_case(N=n, H=h, D=d, dtype=act, cache_mode=mode, seed=42)
    orc.run_case(case2, var_mode="exact")
    k_nope2, k_pe2 = extract_rows(case2)
    assert torch.equal(k_pe2, ref_k_pe)                          # no reduction involved
    assert torch.allclose(case2.t["q_out1"][:, heads].float(), ref_q_pe.float(), rtol=1e-3, atol=1e-3)
    if q0_dt == torch.int8:
        assert (k_nope2.int() - ref_k_nope.int()).abs().max().item() <= 1
    else:
        assert torch.allclose(k_nope2.float(), ref_k_nope.float(), rtol=1e-3, atol=1e-3)


def test_mode_strings_and_defaults():
    case = orc.gen_case(N=2, H=1, D=256, seed=1)
    with pytest.raises(RuntimeError, match="Unsupported cache_mode"):
        case.cache_mode = "bogus"
        orc.run_case(case)
    assert orc.DEFAULT_CACHE_MODE == "krope_ctkv" and orc.DEFAULT_QUANT_MODE == "per_token_quant_symm"


def test_slot_minus_one_is_skipped_and_other_rows_untouched():
    case = orc.gen_case(N=8, H=1, D=256, seed=3, num_blocks=4, neg_slot_frac=0.25)
    case.t["kv_cache"].fill_(7.0)
    case.t["kv_cache_rope"].fill_(7.0)
    taps = {}
    orc.run_case(case, taps=taps)
    slots = case.t["slotmapping"].tolist()
    assert slots.count(-1) == 2
    kv = case.t["kv_cache"].view(-1, 512)
    written = {s for s in slots if s >= 0}
    for row in range(kv.shape[0]):
        if row in written:
            assert torch.equal(kv[row], taps["k_nope"][slots.index(row)])
        else:
            assert bool((kv[row] == 7.0).all())


@pytest.mark.parametrize("mode", ["krope_ctkv", "int8_nzcache", "nzcache"])
def test_cache_layout_roundtrip(mode):
    case = orc.gen_case(N=5, H=1, D=256, seed=5, num_blocks=3, cache_mode=mode)
    taps = {}
    orc.run_case(case, taps=taps)
    k_nope, k_pe = extract_rows(case)
    assert torch.equal(k_nope.squeeze(1), taps["k_nope"])
    assert torch.equal(k_pe.squeeze(1), taps["k_pe"])
    # bytes written == N rows exactly
    assert int((case.t["kv_cache_rope"] != 0).sum()) <= 5 * 64


def test_per_token_mode_runs_and_is_close_to_per_tensor_free_math():
    """Unpinned mode: sanity only — dequantised MM1 output tracks the fp32 matmul."""
    case = orc.gen_case(N=4, H=2, D=512, seed=9)
    taps = {}
    orc.run_case(case, quant_mode="per_token_quant_symm", taps=taps)
    x = case.t["hiddenState"].float()
    ref = (x @ case.t["wdqkv_nd"].float().t()) * case.t["descale0"]
    got = taps["y0"].float()
    assert torch.allclose(got, ref, rtol=0.05, atol=0.05 * ref.abs().max().item())


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_int8_activations_do_not_depend_on_the_reduction_order(path):
    """north_star: int8 activations bit-exact against the golden.  a0 involves no reduction.  a1 = quant(RMSNorm(q_a)) sees the
    fp32 row sum, which `var_mode="torch"` takes in the golden's order and `var_mode="exact"` (the CUDA contract) as the
    exactly accumulated sum: on every fixture the two must give the same int8 a1 (and so the same acc1 / y1 / q_rope), i.e.
    the CUDA kernels, held bit-exact to "exact", are bit-exact to the golden's a1 as well.  Flips are counted, not hidden."""
    n, h, d, dt, mode = parse(path)
    taps_t, taps_e = {}, {}
    orc.run_case(orc.gen_case(N=n, H=h, D=d, dtype=DT[dt], cache_mode=mode, seed=42), var_mode="torch", taps=taps_t)
    orc.run_case(orc.gen_case(N=n, H=h, D=d, dtype=DT[dt], cache_mode=mode, seed=42), var_mode="exact", taps=taps_e)
    flips = {k: int((taps_t[k] != taps_e[k]).sum()) for k in ("a0", "acc0", "y0", "a1", "acc1", "y1", "k_pe")}
    assert flips == {k: 0 for k in flips}, flips
    # k_nope is rounded from the normalised fp32 row: a 1-ulp difference of the row sum may move an element across a rounding
    # boundary; the reference's own tolerance (allclose 1e-3, test_mla:712-744) covers that.  Count what happens here.
    # Observed: 0 on seven fixtures, 1 element of 15 872 on golden_N31_H64_D7168_fp16_krope_ctkv.
    kn_flips = int((taps_t["k_nope"] != taps_e["k_nope"]).sum())
    assert kn_flips <= 1, kn_flips
    assert torch.allclose(taps_t["k_nope"].float(), taps_e["k_nope"].float(), rtol=1e-3, atol=1e-3)
