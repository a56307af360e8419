"""split_qkv_rmsnorm_rope.  CPU: the numpy oracle against fixtures made with the reference test's own golden
helpers (oracle/make_split_qkv_golden.py).  GPU: the CUDA kernel, called through the reference's module path
sgl_kernel_npu.norm.split_qkv_rmsnorm_rope (ctypes over the C ABI), against the fixtures and against the oracle.
Tolerances: the reference test asserts atol 5e-2 (test_split_qkv_rmsnorm_rope.py:122-137).  Here: nearly all elements identical to
the fp32 result rounded once, and none further off than ONE bf16 / fp16 ulp of the magnitude the result was summed from
(|rot * sin| + |n * cos|: where the rotation cancels, an ulp of the small result is not a sound yardstick for fp32 arithmetic
that associates the sum of squares differently and takes 1/sqrt by a Newton step)."""
import os

import numpy as np
import pytest
import torch

from oracle import split_qkv_oracle as sq

GOLD = os.path.join(os.path.dirname(__file__), "golden", "split_qkv_golden.npz")


def bf(bits):
    return torch.from_numpy(bits.view(np.int16).copy()).view(torch.bfloat16)


def fixture(ci):
    z = np.load(GOLD)
    qh, kvh, hd, bsz = (int(v) for v in z["meta"])
    t = {k: bf(z[f"c{ci}_{k}"]) for k in ("qkv", "qw", "kw", "qb", "kb", "sin", "cos")}
    return qh, kvh, hd, bsz, int(z[f"c{ci}_rope_dim"]), t, z[f"c{ci}_q"], z[f"c{ci}_k"]


def ordinal(t):
    """16-bit float bit patterns mapped to integers that are monotone in the value, so that differences count ulps"""
    b = t.contiguous().view(torch.int16).to(torch.int32)
    return torch.where(b < 0, -(b & 0x7FFF), b)


def close(got, want_f32):
    """(fraction identical, worst ulp distance) of `got` against the float32 reference rounded once to got's dtype"""
    want = torch.from_numpy(np.ascontiguousarray(want_f32)).to(got.dtype) if isinstance(want_f32, np.ndarray) else want_f32
    d = (ordinal(got.cpu()) - ordinal(want.reshape(got.shape))).abs()
    return float((d == 0).float().mean()), int(d.max())


# ------------------------------------------------------------------------------------------------ oracle vs reference (CPU)
@pytest.mark.parametrize("ci", [0, 1])
def test_oracle_matches_reference_golden(ci):
    qh, kvh, hd, bsz, rd, t, gq, gk = fixture(ci)
    q, k, v = sq.split_qkv_rmsnorm_rope(t["qkv"], t["sin"].reshape(bsz, rd), t["cos"].reshape(bsz, rd), qh, kvh, hd, 1e-6,
                                        t["qw"], t["kw"], t["qb"], t["kb"])
    for got, want in ((q, gq), (k, gk)):
        same, worst = close(got, want)
        assert same >= 0.999 and worst <= 1, (same, worst)
        np.testing.assert_allclose(got.float().numpy(), want, atol=5e-2)
    assert torch.equal(v, t["qkv"][:, qh + kvh:])


def gated_fixture():
    z = np.load(GOLD)
    qh, kvh, hd, bsz, rd = (int(v) for v in z["g_meta"])
    t = {k: bf(z[f"g_{k}"]) for k in ("qkv", "qw", "kw", "sin", "cos", "v", "gate")}
    return qh, kvh, hd, bsz, rd, t, z["g_q"], z["g_k"]


def within_summed_ulp(got, want_f32, scale, mant=8):
    """|got - want| <= one ulp (mant bits of significand) of the magnitude the result was summed from"""
    err = np.abs(got.float().cpu().numpy() - want_f32)
    return bool((err <= np.exp2(np.floor(np.log2(np.maximum(scale, 1e-30))) - (mant - 1))).all())


def test_oracle_gated_matches_reference_golden():
    qh, kvh, hd, bsz, rd, t, gq, gk = gated_fixture()
    q, k, v, gate = sq.split_qkvgate_gemma_rmsnorm_rope(t["qkv"], t["sin"], t["cos"], qh, kvh, hd, rd, 1e-6, t["qw"], t["kw"])
    for got, want in ((q, gq), (k, gk)):
        same, worst = close(got, want)
        assert same >= 0.999 and worst <= 1, (same, worst)
        np.testing.assert_allclose(got.float().numpy(), want, atol=5e-2)
    assert torch.equal(v, t["v"].reshape(v.shape)) and torch.equal(gate, t["gate"].reshape(gate.shape))


def pos_fixture(ci):
    z = np.load(GOLD)
    bsz, qh, kvh, hd, rd, norm = (int(v) for v in z[f"p{ci}_meta"])
    t = {k: bf(z[f"p{ci}_{k}"]) for k in ("qkv", "qw", "kw", "q", "k", "v")}
    t["pos"], t["cache"] = torch.from_numpy(z[f"p{ci}_pos"]), torch.from_numpy(z[f"p{ci}_cache"])
    return bsz, qh, kvh, hd, rd, bool(norm), t


def assert_ref_tolerance(got, want):
    """the reference test's own bound (_assert_close_fp32, test_split_qkv_rmsnorm_rope_pos_cache_half_npu.py:128-136, :196-198): its
    golden computes op by op in bf16, so it is not closer than this to fp32 arithmetic"""
    torch.testing.assert_close(got.float().cpu(), want.float().reshape(got.shape), atol=5e-2, rtol=5e-3)


@pytest.mark.parametrize("ci", [0, 1, 2])
def test_oracle_pos_cache_matches_reference_golden(ci):
    bsz, qh, kvh, hd, rd, norm, t = pos_fixture(ci)
    q, k, v = sq.split_qkv_rmsnorm_rope_pos_cache_half(t["qkv"], t["pos"], t["cache"], qh, kvh, hd, 1e-6 if norm else None,
                                                       t["qw"] if norm else None, t["kw"] if norm else None, rope_dim=rd)
    assert_ref_tolerance(q, t["q"])
    assert_ref_tolerance(k, t["k"])
    assert torch.equal(v, t["v"].reshape(v.shape))
    # and most elements agree to the bit although the golden rounds after every operation
    assert close(q, t["q"].reshape(q.shape))[0] > 0.6


def mrope_fixture(ci):
    z = np.load(GOLD)
    ntok, nqh, nkh, hs, rd, inter, gate, bias, *section = (int(v) for v in z[f"m{ci}_meta"])
    names = ["qkv", "qw", "kw", "cs", "v"] + (["qb", "kb"] if bias else []) + (["gate"] if gate else [])
    t = {k: bf(z[f"m{ci}_{k}"]) for k in names}
    t["cs"] = t["cs"].reshape(3, ntok, rd)
    return ntok, nqh, nkh, hs, rd, bool(inter), bool(gate), section, t, z[f"m{ci}_q"], z[f"m{ci}_k"]


@pytest.mark.parametrize("ci", [0, 1])
def test_oracle_mrope_matches_reference_golden(ci):
    ntok, nqh, nkh, hs, rd, inter, gate, section, t, gq, gk = mrope_fixture(ci)
    q, k, v, g = sq.split_qkv_rmsnorm_mrope(t["qkv"], t["qw"], t["kw"], t["cs"], nqh, nkh, hs, 1e-6, section, inter, rd, t.get("qb"), t.get("kb"), gate)
    for got, want in ((q, gq), (k, gk)):
        same, worst = close(got, want)
        assert same >= 0.999 and worst <= 1, (same, worst)
    assert torch.equal(v, t["v"].reshape(v.shape))
    assert g.shape == (ntok, nqh * hs if gate else 0) and (not gate or torch.equal(g, t["gate"].reshape(g.shape)))


def tp_fixture():
    z = np.load(GOLD)
    qh, kvh, hd, bsz = (int(v) for v in z["tp_meta"])
    t = {k: bf(z[f"tp_{k}"]) for k in ("qkv", "qw", "kw", "sin", "cos")}
    t["sin"], t["cos"] = t["sin"].reshape(bsz, hd), t["cos"].reshape(bsz, hd)
    return qh, kvh, hd, bsz, t, z["tp_q"], z["tp_k"]


def test_oracle_tp_matches_reference_golden():
    qh, kvh, hd, bsz, t, gq, gk = tp_fixture()
    (q, k, v), = sq.split_qkv_tp_rmsnorm_rope([t["qkv"]], t["cos"], t["sin"], qh, kvh, hd, 1e-6, [t["qw"]], [t["kw"]], hd)
    for got, want in ((q, gq), (k, gk)):
        same, worst = close(got, want)
        assert same >= 0.995 and worst <= 2, (same, worst)   # the intermediate rounding can flip on an fp32 ulp of the norm
        np.testing.assert_allclose(got.float().numpy(), want, atol=5e-2)
    assert torch.equal(v, t["qkv"][:, qh + kvh:])


def tp_close(got, want, dt):
    """tp variant: two roundings, the first one can flip on an fp32 ulp of the norm -> nearly all identical, none beyond 2 ulp of the
    largest magnitude a rotation pair holds (bounded here by the row's maximum)"""
    same, _ = close(got, want)
    ulp = float(np.exp2(np.floor(np.log2(float(want.float().abs().max()))) - (7 if dt == torch.bfloat16 else 10)))
    err = float((got.float().cpu() - want.float()).abs().max())
    return same >= 0.99 and err <= 2 * ulp, (same, err, ulp)


# ------------------------------------------------------------------------------------------------ CUDA
@pytest.mark.gpu
def test_cuda_tp_single_rank_vs_golden_fixture(cuda_device):
    from sgl_kernel_npu.norm.split_qkv_tp_rmsnorm_rope import split_qkv_tp_rmsnorm_rope as fn
    qh, kvh, hd, bsz, t, gq, gk = tp_fixture()
    d = {k: v.to(cuda_device) for k, v in t.items()}
    q, k, v = fn(d["qkv"], d["cos"], d["sin"], qh, kvh, hd, 1e-6, d["qw"], d["kw"], hd, 1, None)
    for got, want in ((q, gq), (k, gk)):
        same, worst = close(got, want)
        assert same >= 0.99 and worst <= 2, (same, worst)
        np.testing.assert_allclose(got.float().cpu().numpy(), want, atol=5e-2)
    assert torch.equal(v.cpu(), t["qkv"][:, qh + kvh:])


# (batch, q_hidden, kv_hidden, head_dim, rotary_dim, dtype, ranks)
TP_CASES = [(12, 6144, 1024, 128, 128, torch.bfloat16, 1), (300, 2048, 512, 128, 64, torch.bfloat16, 1), (7, 512, 256, 64, 32, torch.float16, 1),
            (40, 1024, 256, 256, 64, torch.bfloat16, 2), (12, 6144, 1024, 128, 128, torch.bfloat16, 4), (5, 256, 64, 32, 16, torch.float16, 2)]


@pytest.mark.gpu
@pytest.mark.parametrize("case", TP_CASES, ids=lambda c: "b%d_q%d_kv%d_hd%d_r%d_%s_tp%d" % (c[:5] + (str(c[5])[6:], c[6])))
def test_cuda_tp_vs_oracle(case, cuda_device):
    """ranks > 1: the ranks' shards run one after the other on this GPU through the two-phase modes of the C ABI, the sum of the
    per-rank means (the all-reduce) taken in between; tests/test_split_qkv_tp_nccl.py does the same over NCCL when two GPUs are there"""
    from sgl_kernel_npu.norm import split_qkv_tp_rmsnorm_rope as m
    b, qh, kvh, hd, rd, dt, tp = case
    g = torch.Generator().manual_seed(b + qh + rd + tp)
    xs = [torch.randn(b, qh + 2 * kvh, generator=g).to(dt) * (r + 1) for r in range(tp)]
    qws, kws = [torch.randn(qh, generator=g).to(dt) for _ in range(tp)], [torch.randn(kvh, generator=g).to(dt) for _ in range(tp)]
    cos, sin = torch.rand(b, rd, generator=g).to(dt), torch.rand(b, rd, generator=g).to(dt)
    want = sq.split_qkv_tp_rmsnorm_rope(xs, cos, sin, qh, kvh, hd, 1e-6, qws, kws, rd)
    dev = lambda t: t.to(cuda_device)
    if tp == 1:
        got = [m.split_qkv_tp_rmsnorm_rope(dev(xs[0]), dev(cos), dev(sin), qh, kvh, hd, 1e-6, dev(qws[0]), dev(kws[0]), rd, 1, None)]
    else:
        new = lambda w: torch.empty(b, w, device=cuda_device, dtype=dt)
        bufs = [(dev(xs[r]), dev(cos), dev(sin), new(qh), new(kvh), new(kvh), dev(qws[r]), dev(kws[r])) for r in range(tp)]
        means = [torch.empty(b, 2, device=cuda_device) for _ in range(tp)]
        for r in range(tp):
            m._launch(m.MODE_SUMS, *bufs[r], means[r], qh, kvh, hd, rd, 1e-6, tp)
        total = torch.stack(means).sum(0)
        for r in range(tp):
            m._launch(m.MODE_APPLY, *bufs[r], total, qh, kvh, hd, rd, 1e-6, tp)
        got = [bufs[r][3:6] for r in range(tp)]
    for r in range(tp):
        for name, g_, w_ in zip("qk", got[r][:2], want[r][:2]):
            ok, info = tp_close(g_, w_, dt)
            assert ok, (r, name, info)
        assert torch.equal(got[r][2].cpu(), want[r][2])


@pytest.mark.gpu
@pytest.mark.parametrize("ci", [0, 1])
def test_cuda_mrope_vs_golden_fixture(ci, cuda_device):
    from sgl_kernel_npu.norm.split_qkv_rmsnorm_mrope import triton_split_qkv_rmsnorm_mrope as fn
    ntok, nqh, nkh, hs, rd, inter, gate, section, t, gq, gk = mrope_fixture(ci)
    d = {k: v.to(cuda_device) for k, v in t.items()}
    q, k, v, g = fn(d["qkv"], d["qw"], d["kw"], d["cs"], nqh, nkh, hs, 1e-6, section, inter, rope_dim=rd, q_bias=d.get("qb"), k_bias=d.get("kb"),
                    has_gate=gate)
    *_, mq, mk = sq.split_qkv_rmsnorm_mrope(t["qkv"], t["qw"], t["kw"], t["cs"], nqh, nkh, hs, 1e-6, section, inter, rd, t.get("qb"), t.get("kb"),
                                            gate, with_scale=True)
    for got, want, m in ((q, gq, mq), (k, gk, mk)):
        assert close(got, want)[0] >= 0.995 and within_summed_ulp(got, want, m)
    assert torch.equal(v.cpu(), t["v"].reshape(v.shape))
    assert g.shape == (ntok, nqh * hs if gate else 0) and (not gate or torch.equal(g.cpu(), t["gate"].reshape(g.shape)))


# (tokens, q heads, kv heads, head_size, rope_dim, section, interleaved, gate, bias, cos_sin dtype, dtype)
MROPE_CASES = [
    (300, 16, 2, 128, 128, [24, 20, 20], True, False, False, torch.bfloat16, torch.bfloat16),    # Qwen2.5-VL-like sections
    (300, 16, 2, 128, 128, [16, 24, 24], False, False, True, torch.float32, torch.bfloat16),
    (33, 8, 2, 256, 64, [11, 11, 10], True, True, False, torch.float16, torch.float16),          # partial rotary, gate (Qwen3.5-like)
    (5, 4, 1, 64, 20, [4, 3, 3], False, True, True, torch.bfloat16, torch.bfloat16),              # scalar kernel
]


@pytest.mark.gpu
@pytest.mark.parametrize("case", MROPE_CASES, ids=lambda c: "t%d_h%d_%d_hs%d_r%d_%s" % (c[:5] + ("".join(str(int(x)) for x in c[6:9]),)))
def test_cuda_mrope_vs_oracle(case, cuda_device):
    from sgl_kernel_npu.norm.split_qkv_rmsnorm_mrope import triton_split_qkv_rmsnorm_mrope as fn
    ntok, nqh, nkh, hs, rd, section, inter, gate, bias, cdt, dt = case
    g = torch.Generator().manual_seed(ntok + hs + rd)
    x = torch.randn(ntok, nqh * hs * (2 if gate else 1) + 2 * nkh * hs, generator=g).to(dt)
    cs = (torch.rand(3, ntok, rd, generator=g) * 2 - 1).to(cdt)
    rn = lambda: torch.randn(hs, generator=g).to(dt)
    qw, kw = rn(), rn()
    qb, kb = (rn(), rn()) if bias else (None, None)
    wq, wk, wv, wg, mq, mk = sq.split_qkv_rmsnorm_mrope(x, qw, kw, cs, nqh, nkh, hs, 1e-6, section, inter, rd, qb, kb, gate, with_scale=True)
    dev = lambda t: None if t is None else t.to(cuda_device)
    q, k, v, gt = fn(dev(x), dev(qw), dev(kw), dev(cs), nqh, nkh, hs, 1e-6, section, inter, rope_dim=rd, q_bias=dev(qb), k_bias=dev(kb), has_gate=gate)
    mant = 8 if dt == torch.bfloat16 else 11
    for got, want, m in ((q, wq, mq), (k, wk, mk)):
        assert close(got, want)[0] >= 0.99
        assert within_summed_ulp(got, want.float().numpy(), m, mant)
    assert torch.equal(v.cpu(), wv) and torch.equal(gt.cpu(), wg)


@pytest.mark.gpu
@pytest.mark.parametrize("ci", [0, 1, 2])
def test_cuda_pos_cache_vs_golden_fixture(ci, cuda_device):
    from sgl_kernel_npu.norm.split_qkv_rmsnorm_rope_pos_cache_half_npu import split_qkv_rmsnorm_rope_pos_cache_half_npu as fn
    bsz, qh, kvh, hd, rd, norm, t = pos_fixture(ci)
    d = {k: v.to(cuda_device) for k, v in t.items()}
    q, k, v = fn(d["qkv"], d["pos"], d["cache"], qh, kvh, hd, eps=1e-6 if norm else None, q_weight=d["qw"] if norm else None,
                 k_weight=d["kw"] if norm else None, rope_dim=rd)
    assert_ref_tolerance(q, t["q"])
    assert_ref_tolerance(k, t["k"])
    assert torch.equal(v.cpu(), t["v"].reshape(v.shape))


# (batch, q_hidden, kv_hidden, head_dim, rope_dim, norm, bias, cast, cache dtype, positions dtype, input dtype)
POS_CASES = [
    (12, 6144, 1024, 128, 128, True, False, True, torch.float32, torch.int64, torch.bfloat16),     # the reference test's shapes
    (8, 4096, 1024, 128, 64, True, False, True, torch.float32, torch.int64, torch.bfloat16),
    (300, 2048, 512, 128, 64, True, True, False, torch.bfloat16, torch.int32, torch.bfloat16),
    (9, 1024, 512, 512, 128, True, True, True, torch.float16, torch.int32, torch.float16),
    (7, 512, 256, 64, 64, False, False, True, torch.float32, torch.int64, torch.bfloat16),
    (5, 256, 128, 64, 20, True, True, True, torch.float32, torch.int64, torch.bfloat16),          # scalar kernel
]


@pytest.mark.gpu
@pytest.mark.parametrize("case", POS_CASES, ids=lambda c: "b%d_q%d_kv%d_hd%d_r%d_%d%d%d_%s_%s_%s" % (
    c[:5] + tuple(int(x) for x in c[5:8]) + tuple(str(x)[6:] for x in c[8:])))
def test_cuda_pos_cache_vs_oracle(case, cuda_device):
    from sgl_kernel_npu.norm.split_qkv_rmsnorm_rope_pos_cache_half_npu import split_qkv_rmsnorm_rope_pos_cache_half_npu as fn
    b, qh, kvh, hd, rd, norm, bias, cast, cdt, pdt, dt = case
    g = torch.Generator().manual_seed(b * 7 + hd + rd)
    x = torch.randn(b, qh + 2 * kvh, generator=g).to(dt)
    max_seq = 97
    cache = (torch.rand(max_seq, rd, generator=g) * 2 - 1).to(cdt)
    pos = torch.randint(-3, max_seq + 3, (b,), generator=g).to(pdt)   # a few positions outside the cache: clamped
    rn = lambda: torch.randn(hd, generator=g).to(dt)
    qw, kw = (rn(), rn()) if norm else (None, None)
    qb, kb = (rn(), rn()) if norm and bias else (None, None)
    eps = 1e-6 if norm else None
    wq, wk, wv, mq, mk = sq.split_qkv_rmsnorm_rope_pos_cache_half(x, pos, cache, qh, kvh, hd, eps, qw, kw, qb, kb, rd, cast, with_scale=True)
    dev = lambda t: None if t is None else t.to(cuda_device)
    q, k, v = fn(dev(x), dev(pos), dev(cache), qh, kvh, hd, eps=eps, q_weight=dev(qw), k_weight=dev(kw), q_bias=dev(qb), k_bias=dev(kb),
                 rope_dim=rd, cast_norm_to_bf16=cast)
    mant = 8 if dt == torch.bfloat16 else 11
    for got, want, m in ((q, wq, mq), (k, wk, mk)):
        # with cast_norm a 1-ulp difference of the fp32 norm can flip the intermediate rounding: fewer identical elements then
        assert close(got, want)[0] >= (0.97 if cast and norm else 0.99)
        lim = 2 if cast and norm else 1
        err = np.abs(got.float().cpu().numpy() - want.float().numpy())
        assert (err <= lim * np.exp2(np.floor(np.log2(np.maximum(m, 1e-30))) - (mant - 1))).all()
    assert torch.equal(v.cpu(), wv)
    with pytest.raises(ValueError):
        fn(dev(x), dev(pos), dev(cache), qh, kvh, hd, eps=1e-6, q_weight=dev(rn())[:hd // 2], k_weight=dev(rn()), rope_dim=rd)


@pytest.mark.gpu
def test_cuda_gated_vs_golden_fixture(cuda_device):
    from sgl_kernel_npu.norm.split_qkv_rmsnorm_rope import split_qkvgate_gemma_rmsnorm_rope
    qh, kvh, hd, bsz, rd, t, gq, gk = gated_fixture()
    d = {k: v.to(cuda_device) for k, v in t.items()}
    q, k, v, gate = split_qkvgate_gemma_rmsnorm_rope(d["qkv"], d["sin"], d["cos"], qh, kvh, hd, rd, 1e-6, d["qw"], d["kw"])
    *_, mq, mk = sq.split_qkvgate_gemma_rmsnorm_rope(t["qkv"], t["sin"], t["cos"], qh, kvh, hd, rd, 1e-6, t["qw"], t["kw"], with_scale=True)
    for got, want, m in ((q, gq, mq), (k, gk, mk)):
        assert close(got, want)[0] >= 0.995 and within_summed_ulp(got, want, m)
        np.testing.assert_allclose(got.float().cpu().numpy(), want, atol=5e-2)
    assert torch.equal(v.cpu(), t["v"].reshape(v.shape)) and torch.equal(gate.cpu(), t["gate"].reshape(gate.shape))


# (batch, q_hidden, kv_hidden, head_dim, rope_dim, dtype): Qwen3-Next-like (partial rotary 0.25), a ragged chunking, the scalar kernel
GATED_CASES = [(12, 512, 256, 256, 64, torch.bfloat16), (300, 4096, 512, 256, 64, torch.bfloat16), (33, 2048, 1024, 128, 128, torch.float16),
               (5, 256, 128, 64, 20, torch.bfloat16), (4, 64, 32, 16, 8, torch.bfloat16)]


@pytest.mark.gpu
@pytest.mark.parametrize("case", GATED_CASES, ids=lambda c: "b%d_q%d_kv%d_hd%d_r%d_%s" % (c[:5] + (str(c[5])[6:],)))
def test_cuda_gated_vs_oracle(case, cuda_device):
    from sgl_kernel_npu.norm.split_qkv_rmsnorm_rope import split_qkvgate_gemma_rmsnorm_rope
    b, qh, kvh, hd, rd, dt = case
    g = torch.Generator().manual_seed(b * 17 + hd + rd)
    x = torch.randn(b, 2 * qh + 2 * kvh, generator=g).to(dt)
    sin, cos = torch.rand(b, rd, generator=g).to(dt), torch.rand(b, rd, generator=g).to(dt)
    qw, kw = torch.randn(hd, generator=g).to(dt), torch.randn(hd, generator=g).to(dt)
    wq, wk, wv, wg, mq, mk = sq.split_qkvgate_gemma_rmsnorm_rope(x, sin, cos, qh, kvh, hd, rd, 1e-6, qw, kw, with_scale=True)
    dev = lambda t: t.to(cuda_device)
    q, k, v, gate = split_qkvgate_gemma_rmsnorm_rope(dev(x), dev(sin), dev(cos), qh, kvh, hd, rd, 1e-6, dev(qw), dev(kw))
    mant = 8 if dt == torch.bfloat16 else 11
    for got, want, m in ((q, wq, mq), (k, wk, mk)):
        assert close(got, want)[0] >= 0.99
        assert within_summed_ulp(got, want.float().numpy(), m, mant)
    assert torch.equal(v.cpu(), wv) and torch.equal(gate.cpu(), wg)


@pytest.mark.gpu
@pytest.mark.parametrize("ci", [0, 1])
def test_cuda_vs_golden_fixture(ci, cuda_device):
    from sgl_kernel_npu.norm.split_qkv_rmsnorm_rope import split_qkv_rmsnorm_rope
    qh, kvh, hd, bsz, rd, t, gq, gk = fixture(ci)
    d = {k: v.to(cuda_device) for k, v in t.items()}
    q, k, v = split_qkv_rmsnorm_rope(d["qkv"], d["sin"], d["cos"], qh, kvh, hd, eps=1e-6, q_weight=d["qw"], k_weight=d["kw"],
                                     q_bias=d["qb"], k_bias=d["kb"])
    *_, mq, mk = sq.split_qkv_rmsnorm_rope(t["qkv"], t["sin"].reshape(bsz, rd), t["cos"].reshape(bsz, rd), qh, kvh, hd, 1e-6,
                                           t["qw"], t["kw"], t["qb"], t["kb"], with_scale=True)
    for got, want, m in ((q, gq, mq), (k, gk, mk)):
        same, _ = close(got, want)
        assert same >= 0.995, same
        err = np.abs(got.float().cpu().numpy() - want)
        assert (err <= np.exp2(np.floor(np.log2(np.maximum(m, 1e-30))) - 7)).all()  # one bf16 ulp of the summed magnitude
        np.testing.assert_allclose(got.float().cpu().numpy(), want, atol=5e-2)
    assert torch.equal(v.cpu(), t["qkv"][:, qh + kvh:])


# (batch, q_hidden, kv_hidden, head_dim, rope_dim, norm, bias, neox, dtype); the first two are the reference test's own shapes
CASES = [
    (12, 6144, 1024, 128, 128, True, True, True, torch.bfloat16),
    (12, 6144, 1024, 128, 64, True, True, True, torch.bfloat16),
    (12, 6144, 1024, 128, 128, False, False, True, torch.bfloat16),   # test_split_qkv_rope: no norm
    (5, 512, 256, 64, 32, True, False, True, torch.float16),
    (7, 1024, 256, 128, 128, True, True, False, torch.bfloat16),
    (3, 256, 128, 32, 16, False, False, True, torch.bfloat16),
    (1, 1024, 1024, 256, 64, True, True, False, torch.float16),
    (257, 4096, 512, 128, 64, True, False, True, torch.bfloat16),
    (4, 512, 256, 128, 24, True, True, True, torch.bfloat16),      # rope_dim a multiple of 8 only
    (4, 512, 256, 128, 96, True, False, False, torch.bfloat16),
    (9, 1024, 512, 512, 128, True, True, True, torch.float16),    # 16 elements per lane
    (9, 1024, 512, 512, 64, True, True, False, torch.bfloat16),
    (3, 256, 128, 64, 20, True, True, True, torch.bfloat16),       # rope_dim the vector layout does not fit: scalar kernel
    (3, 64, 32, 16, 8, True, True, False, torch.float16),          # head_dim below 32: scalar kernel
]


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=lambda c: "b%d_q%d_kv%d_hd%d_r%d_%d%d%d_%s" % (c[:5] + tuple(int(x) for x in c[5:8]) + (str(c[8])[6:],)))
def test_cuda_vs_oracle(case, cuda_device):
    from sgl_kernel_npu.norm.split_qkv_rmsnorm_rope import split_qkv_rmsnorm_rope
    b, qh, kvh, hd, rd, norm, bias, neox, dt = case
    g = torch.Generator().manual_seed(b * 131 + hd + rd)
    rn = lambda *s: torch.randn(*s, generator=g).to(dt)
    qkv = rn(b, qh + 2 * kvh)
    sin, cos = torch.rand(b, rd, generator=g).to(dt), torch.rand(b, rd, generator=g).to(dt)
    qw, kw = (rn(hd), rn(hd)) if norm else (None, None)
    qb, kb = (rn(hd), rn(hd)) if norm and bias else (None, None)
    eps = 1e-6 if norm else None
    want = sq.split_qkv_rmsnorm_rope(qkv, sin, cos, qh, kvh, hd, eps, qw, kw, qb, kb, neox, with_scale=True)
    dev = lambda t: None if t is None else t.to(cuda_device)
    got = split_qkv_rmsnorm_rope(dev(qkv), dev(sin), dev(cos), qh, kvh, hd, eps=eps, q_weight=dev(qw), k_weight=dev(kw),
                                 q_bias=dev(qb), k_bias=dev(kb), is_neox_style=neox)
    mant = 8 if dt == torch.bfloat16 else 11
    for name, g_, w_, m_ in zip("qk", got[:2], want[:2], want[3:]):
        same, worst = close(g_, w_)
        assert same >= 0.99, (name, same, worst)
        ulp = np.exp2(np.floor(np.log2(np.maximum(m_, 1e-30))) - (mant - 1))   # one ulp of the summed magnitude
        err = np.abs(g_.float().cpu().numpy() - w_.float().numpy())
        assert (err <= ulp).all(), (name, float((err / ulp).max()))
    assert torch.equal(got[2].cpu(), want[2])
    if not norm:  # no reduction on this path: nothing is associated differently, so the result is bit-exact
        assert torch.equal(got[0].cpu(), want[0]) and torch.equal(got[1].cpu(), want[1])


@pytest.mark.gpu
def test_argument_checks(cuda_device):
    from sgl_kernel_npu.norm.split_qkv_rmsnorm_rope import split_qkv_rmsnorm_rope
    x = torch.zeros(2, 96 * 3, dtype=torch.bfloat16, device=cuda_device)
    s = torch.zeros(2, 32, dtype=torch.bfloat16, device=cuda_device)
    with pytest.raises(AssertionError):
        split_qkv_rmsnorm_rope(x, s, s, 96, 96, 96)                    # head_dim not a power of two
    x = torch.zeros(2, 128 + 2 * 96, dtype=torch.bfloat16, device=cuda_device)
    with pytest.raises(AssertionError):
        split_qkv_rmsnorm_rope(x, s, s, 128, 96, 32)                   # q_hidden not a multiple of kv_hidden
    e = torch.zeros(0, 128 * 3, dtype=torch.bfloat16, device=cuda_device)
    q, k, v = split_qkv_rmsnorm_rope(e, s[:0], s[:0], 128, 128, 64)    # empty batch: empty outputs, no launch
    assert q.shape == (0, 128) and k.shape == (0, 128) and v.shape == (0, 128)
