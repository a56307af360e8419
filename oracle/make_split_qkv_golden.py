"""Writes tests/golden/split_qkv_golden.npz with the reference TEST's own golden helpers rms_norm / custom_rope
(tests/python/sgl_kernel_npu/test_split_qkv_rmsnorm_rope.py:11-44), imported from /root/reference with the NPU modules stubbed,
applied the way that test applies them (:101-120), on a scaled-down copy of its two cases (rope_dim = head_dim and head_dim / 2).
Inputs are stored as bf16 bit patterns, the golden q / k as float32 (the helpers compute in float32 and do not round)."""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference/tests/python/sgl_kernel_npu/test_split_qkv_rmsnorm_rope.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "split_qkv_golden.npz")


def load():
    for m in ("torch_npu", "sgl_kernel_npu", "sgl_kernel_npu.norm"):
        sys.modules.setdefault(m, types.ModuleType(m))
    stub = types.ModuleType("sgl_kernel_npu.norm.split_qkv_rmsnorm_rope")
    stub.split_qkv_rmsnorm_rope = stub.split_qkvgate_gemma_rmsnorm_rope = None
    sys.modules["sgl_kernel_npu.norm.split_qkv_rmsnorm_rope"] = stub
    spec = importlib.util.spec_from_file_location("ref_test_split_qkv", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load_pos_cache_test():
    for m in ("sglang", "sglang.srt"):
        sys.modules.setdefault(m, types.ModuleType(m))
    u = types.ModuleType("sglang.srt.utils")
    u.is_npu = lambda: False
    sys.modules["sglang.srt.utils"] = u
    spec = importlib.util.spec_from_file_location("ref_test_pos_cache", os.path.join(os.path.dirname(REF), "test_split_qkv_rmsnorm_rope_pos_cache_half_npu.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load_named(test_file, module, attr):
    for m in ("torch_npu", "sgl_kernel_npu", "sgl_kernel_npu.norm"):
        sys.modules.setdefault(m, types.ModuleType(m))
    stub = types.ModuleType(module)
    setattr(stub, attr, None)
    sys.modules[module] = stub
    spec = importlib.util.spec_from_file_location("ref_" + test_file[:-3], os.path.join(os.path.dirname(REF), test_file))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def bits(t):
    return t.contiguous().view(torch.int16).numpy().view(np.uint16)


def main():
    ref = load()
    qh, kvh, hd, bsz, eps = 1024, 256, 128, 6, 1e-6
    rec = {"meta": np.array([qh, kvh, hd, bsz])}
    for ci, rope_dim in enumerate((128, 64)):
        g = torch.Generator().manual_seed(100 + ci)
        bf = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16)
        qkv, qw, kw, qb, kb = bf(bsz, qh + 2 * kvh), bf(hd), bf(hd), bf(hd), bf(hd)
        sin = torch.rand(bsz, 1, 1, rope_dim, generator=g).to(torch.bfloat16)
        cos = torch.rand(bsz, 1, 1, rope_dim, generator=g).to(torch.bfloat16)
        _q, _k, _v = qkv.split([qh, kvh, kvh], dim=-1)
        nq = ref.rms_norm(_q.reshape(-1, hd), qw, qb, eps).reshape(bsz, 1, -1, hd)
        nk = ref.rms_norm(_k.reshape(-1, hd), kw, kb, eps).reshape(bsz, 1, -1, hd)
        rq, rk = ref.custom_rope(nq[..., :rope_dim], nk[..., :rope_dim], sin, cos, half_rope_dim=rope_dim // 2)
        cq = np.concatenate((rq, nq[..., rope_dim:]), axis=-1).reshape(bsz, -1)
        ck = np.concatenate((rk, nk[..., rope_dim:]), axis=-1).reshape(bsz, -1)
        for name, t in (("qkv", qkv), ("qw", qw), ("kw", kw), ("qb", qb), ("kb", kb), ("sin", sin), ("cos", cos)):
            rec[f"c{ci}_{name}"] = bits(t)
        rec[f"c{ci}_q"], rec[f"c{ci}_k"] = cq.astype(np.float32), ck.astype(np.float32)
        rec[f"c{ci}_rope_dim"] = np.array(rope_dim)
    # the gated gemma variant through the reference test's golden_ref_split_qkvgate_gemma_rmsnorm_rope (:221-266), its own shape
    qh, kvh, hd, bsz, rope_dim = 512, 256, 256, 12, 64
    g = torch.Generator().manual_seed(200)
    bf = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16)
    x, qw, kw = bf(bsz, 2 * qh + 2 * kvh), bf(hd), bf(hd)
    sin = torch.rand(bsz, 1, 1, rope_dim, generator=g).to(torch.bfloat16)
    cos = torch.rand(bsz, 1, 1, rope_dim, generator=g).to(torch.bfloat16)
    gq, gk, gv, gg = ref.golden_ref_split_qkvgate_gemma_rmsnorm_rope(x, sin, cos, qh, kvh, hd, rope_dim, eps, qw, kw)
    for name, t in (("qkv", x), ("qw", qw), ("kw", kw), ("sin", sin), ("cos", cos), ("v", gv), ("gate", gg)):
        rec[f"g_{name}"] = bits(t)
    rec["g_q"], rec["g_k"] = gq.astype(np.float32), gk.astype(np.float32)
    rec["g_meta"] = np.array([qh, kvh, hd, bsz, rope_dim])
    # the position-cache variant through the reference test's golden_split_qkv_rmsnorm_rope_pos_cache_half
    # (test_split_qkv_rmsnorm_rope_pos_cache_half_npu.py:73-125; it computes in bf16, op by op) and its cache builder (:14-33)
    pc = load_pos_cache_test()
    for ci, (bsz, qh, kvh, hd, rope_dim, norm) in enumerate(((12, 1024, 256, 128, 128, True), (8, 1024, 256, 128, 64, True), (8, 512, 256, 64, 64, False))):
        g = torch.Generator().manual_seed(300 + ci)
        max_pos = 128
        x = torch.randn(bsz, qh + 2 * kvh, generator=g).to(torch.bfloat16)
        pos = torch.randint(0, max_pos, (bsz,), generator=g, dtype=torch.int64)
        cache = pc._make_cos_sin_cache(max_pos, rope_dim, 10000.0, torch.float32, torch.device("cpu"))
        qw, kw = torch.randn(hd, generator=g).to(torch.bfloat16), torch.randn(hd, generator=g).to(torch.bfloat16)
        gq, gk, gv = pc.golden_split_qkv_rmsnorm_rope_pos_cache_half(x, pos, cache, qh, kvh, hd, rope_dim, eps if norm else None,
                                                                     qw if norm else None, kw if norm else None, use_qk_norm=norm)
        for name, t in (("qkv", x), ("qw", qw), ("kw", kw), ("q", gq), ("k", gk), ("v", gv)):
            rec[f"p{ci}_{name}"] = bits(t)
        rec[f"p{ci}_pos"], rec[f"p{ci}_cache"] = pos.numpy(), cache.numpy()
        rec[f"p{ci}_meta"] = np.array([bsz, qh, kvh, hd, rope_dim, int(norm)])
    # the mrope variant through the reference test's _golden (test_split_qkv_rmsnorm_mrope.py:74-110), its two parametrisations
    mr = load_named("test_split_qkv_rmsnorm_mrope.py", "sgl_kernel_npu.norm.split_qkv_rmsnorm_mrope", "triton_split_qkv_rmsnorm_mrope")
    for ci, (ntok, rope_dim, section, inter, gate, bias) in enumerate(((7, 256, [32, 48, 48], True, True, False), (65, 128, [48, 40, 40], False, False, True))):
        g = torch.Generator().manual_seed(400 + ci)
        nqh, nkh, hs = 2, 1, 256
        bf = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16)
        x = bf(ntok, nqh * hs * (2 if gate else 1) + 2 * nkh * hs)
        qw, kw = bf(hs), bf(hs)
        qb, kb = (bf(hs), bf(hs)) if bias else (None, None)
        cs = (torch.rand(3, ntok, rope_dim, generator=g) * 2 - 1).to(torch.bfloat16)
        gq, gk, gv, gg = mr._golden(x, qw, kw, cs, nqh, nkh, hs, eps, section, inter, rope_dim, qb, kb, gate)
        for name, t in (("qkv", x), ("qw", qw), ("kw", kw), ("cs", cs), ("v", gv)) + ((("qb", qb), ("kb", kb)) if bias else ()) + ((("gate", gg),) if gate else ()):
            rec[f"m{ci}_{name}"] = bits(t)
        rec[f"m{ci}_q"], rec[f"m{ci}_k"] = gq.numpy().astype(np.float32), gk.numpy().astype(np.float32)
        rec[f"m{ci}_meta"] = np.array([ntok, nqh, nkh, hs, rope_dim, int(inter), int(gate), int(bias)] + section)
    # the tensor-parallel variant at tp_world = 1 through the reference test's rms_norm_tp / custom_rope
    # (test_split_qkv_tp_rmsnorm_rope.py:7-45), applied as that test does (:78-86), on a scaled-down copy of its case
    tp = load_named("test_split_qkv_tp_rmsnorm_rope.py", "sgl_kernel_npu.norm.split_qkv_tp_rmsnorm_rope", "split_qkv_tp_rmsnorm_rope")
    qh, kvh, hd, bsz = 1024, 256, 128, 6
    g = torch.Generator().manual_seed(500)
    bf = lambda *s: torch.randn(*s, generator=g).to(torch.bfloat16)
    x, qw, kw = bf(bsz, qh + 2 * kvh), bf(qh), bf(kvh)
    sin = torch.rand(bsz, 1, 1, hd, generator=g).to(torch.bfloat16)
    cos = torch.rand(bsz, 1, 1, hd, generator=g).to(torch.bfloat16)
    _q, _k, _v = x.split([qh, kvh, kvh], dim=-1)
    nq = tp.rms_norm_tp(_q, qw, eps, 1).reshape(bsz, 1, -1, hd)
    nk = tp.rms_norm_tp(_k, kw, eps, 1).reshape(bsz, 1, -1, hd)
    cq, ck = tp.custom_rope(nq, nk, sin, cos)
    for name, t in (("qkv", x), ("qw", qw), ("kw", kw), ("sin", sin), ("cos", cos)):
        rec[f"tp_{name}"] = bits(t)
    rec["tp_q"], rec["tp_k"] = cq.reshape(bsz, -1).astype(np.float32), ck.reshape(bsz, -1).astype(np.float32)
    rec["tp_meta"] = np.array([qh, kvh, hd, bsz])
    np.savez_compressed(OUT, **rec)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
