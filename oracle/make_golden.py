"""Generate ``tests/golden/*.npz`` from the REFERENCE's own generator and golden.

TEST INFRASTRUCTURE ONLY (see oracle/mla_oracle.py header).  Runs only in the build
container, where ``/root/reference`` exists; the fixtures it writes are committed so
that the GPU box (which has no reference tree) can still check against them.

It imports ``tests/python/sgl_kernel_npu/test_mla_preprocess.py`` unmodified from
``/root/reference`` with three import stubs (``torch_npu``, ``sgl_kernel_npu``,
``utils`` — none of them is touched by ``gen_random_tensors`` / ``golden2_pytorch``),
calls ``TestMLAPO.gen_random_tensors(device="cpu")`` and
``TestMLAPO.golden2_pytorch(dev="cpu")`` (test_mla:140-321, :407-484) and stores

* sha256 of every generated input tensor (so tests can prove that
  ``oracle.mla_oracle.gen_case`` reproduces the reference inputs bit for bit), and
* the golden outputs ``q_nope, q_pe, k_nope, k_pe`` (q outputs restricted to a few
  heads to keep fixtures small; ``heads`` is recorded).

Usage:  python oracle/make_golden.py            (rewrites tests/golden/)
"""
import hashlib
import importlib.util
import os
import sys
import types

import numpy as np
import torch

REF_TEST = "/root/reference/tests/python/sgl_kernel_npu/test_mla_preprocess.py"
OUT_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

# (N, H, D, dtype, cache_mode): the reference's shapes (test_mla:487-498) x its 4 dtype/mode cases (:746-808)
CASES = [
    (1, 32, 7168, "bf16", "krope_ctkv"),
    (16, 32, 7168, "bf16", "krope_ctkv"),
    (16, 64, 7168, "bf16", "int8_nzcache"),
    (31, 32, 7168, "bf16", "nzcache"),
    (31, 128, 7168, "bf16", "krope_ctkv"),
    (31, 128, 6144, "bf16", "int8_nzcache"),
    (16, 32, 7168, "fp16", "krope_ctkv"),
    (31, 64, 7168, "fp16", "krope_ctkv"),
]
DT = {"bf16": torch.bfloat16, "fp16": torch.float16}

# reference dataDict key -> op-surface name (SURVEY §9.1 naming map; test_mla:547-578)
NAME_MAP = {
    "hidden": "hiddenState", "gamma1": "gamma0", "beta1": "beta0", "qscale1": "quant_scale0",
    "qoffset1": "quant_offset0", "wdqkv": "wdqkv_nd", "bias1": "bias0", "descale1": "descale0",
    "gamma2": "gamma1", "beta2": "beta1", "gamma3": "gamma2", "qscale2": "quant_scale1",
    "qoffset2": "quant_offset1", "wuq": "wuq_nd", "bias2": "bias1", "descale2": "descale1",
    "wuk": "wuk", "sin": "sin", "cos": "cos", "slotMapping": "slotmapping",
    "ctkv_scale": "ctkv_scale", "qnope_scale": "q_nope_scale",
}


def tensor_sha(t: torch.Tensor) -> str:
    t = t.detach().cpu().contiguous()
    return hashlib.sha256(t.view(torch.uint8).numpy().tobytes()).hexdigest()


def to_np(t: torch.Tensor) -> np.ndarray:
    """bf16 has no numpy dtype: store 16-bit floats as their uint16 bit pattern."""
    t = t.detach().cpu().contiguous()
    if t.dtype in (torch.bfloat16, torch.float16):
        return t.view(torch.int16).numpy().view(np.uint16)
    return t.numpy()


def load_reference_test():
    def stub(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    import unittest
    stub("torch_npu")
    stub("torch_npu.testing")
    stub("torch_npu.testing.testcase", TestCase=unittest.TestCase, run_tests=lambda: None)
    stub("sgl_kernel_npu")
    stub("utils", require_npu_op=lambda name: None)
    if not hasattr(torch, "npu"):
        torch.npu = types.SimpleNamespace(config=types.SimpleNamespace(allow_internal_format=False))
    spec = importlib.util.spec_from_file_location("ref_test_mla_preprocess", REF_TEST)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def pick_heads(h):
    return sorted({0, 1, h // 2, h - 1})


def main():
    ref = load_reference_test()
    os.makedirs(OUT_DIR, exist_ok=True)
    for (n, h, d, dt, mode) in CASES:
        tc = ref.TestMLAPO()
        tc.gen_random_tensors(N=n, headNum=h, hiddenDim=d, dtype=DT[dt], cache_mode=mode, device="cpu", seed=42)
        out = tc.golden2_pytorch(N=n, headNum=h, dtype=DT[dt], cache_mode=mode, dev="cpu")
        heads = pick_heads(h)
        rec = {
            "meta": np.array([n, h, d], dtype=np.int64),
            "heads": np.array(heads, dtype=np.int64),
            "q_nope": to_np(out["q_nope"][:, heads, :]),
            "q_pe": to_np(out["q_pe"][:, heads, :]),
            "k_nope": to_np(out["k_nope"]),
            "k_pe": to_np(out["k_pe"]),
            "q_nope_sha": np.array(tensor_sha(out["q_nope"])),
            "q_pe_sha": np.array(tensor_sha(out["q_pe"])),
        }
        for ref_key, op_key in NAME_MAP.items():
            rec["sha_" + op_key] = np.array(tensor_sha(tc.dataDict[ref_key]))
        name = f"golden_N{n}_H{h}_D{d}_{dt}_{mode}.npz"
        np.savez_compressed(os.path.join(OUT_DIR, name), **rec)
        print("wrote", name, {k: v.shape for k, v in rec.items() if hasattr(v, "shape") and v.ndim > 0})


def format_fixtures(ref):
    """Known answers of the reference's own format helpers (test_mla:38-54, 110-135) on small seeded inputs."""
    g = torch.Generator().manual_seed(7)
    rec = {}
    # aligned shapes only: the reference helper hands F.pad its row/column paddings in swapped order (test_mla:42-43),
    # so it fails on ragged inputs; padding is checked against oracle.nz_pack instead
    w8 = torch.randint(-128, 128, (64, 128), dtype=torch.int8, generator=g)
    rec["td_i8_in"] = w8.numpy()
    rec["td_i8_out"] = ref.transdata(w8, (16, 32)).numpy()
    w16 = torch.randint(-2 ** 15, 2 ** 15, (32, 48), dtype=torch.int16, generator=g)  # 16-bit payload, block (16, 16)
    rec["td_16_in"] = w16.numpy()
    rec["td_16_out"] = ref.transdata(w16, (16, 16)).numpy()
    d = (torch.rand(257, generator=g) / 1000).numpy().astype(np.float32)
    rec["descale_in"] = d.copy()
    rec["descale_out"] = ref.trans_descale_param(d.copy()).numpy()
    cache = torch.randint(-2 ** 15, 2 ** 15, (3, 64 // 16, 128, 16), dtype=torch.int16, generator=g)  # rope-cache geometry
    slots = np.array([0, 1, 127, 128, 200, 383], dtype=np.int32)
    rows = [ref.extract_from_nzcache(cache.view(3, -1, 1, 64), int(s) // 128, int(s) % 128)[0].numpy() for s in slots]
    rec["nz_cache"] = cache.numpy()
    rec["nz_slots"] = slots
    rec["nz_rows"] = np.stack(rows)
    cache8 = torch.randint(-128, 128, (2, 64 // 32, 128, 32), dtype=torch.int8, generator=g)
    rows8 = [ref.extract_from_nzcache(cache8.view(2, -1, 1, 64), int(s) // 128, int(s) % 128, 32)[0].numpy() for s in slots[:5]]
    rec["nz_cache8"] = cache8.numpy()
    rec["nz_rows8"] = np.stack(rows8)
    np.savez_compressed(os.path.join(OUT_DIR, "format_tools.npz"), **rec)
    print("wrote format_tools.npz")


if __name__ == "__main__":
    if "--formats-only" not in sys.argv:
        main()
    format_fixtures(load_reference_test())
