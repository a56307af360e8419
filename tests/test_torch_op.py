"""The drop-in operator: torch.ops.npu.mla_preprocess with the reference's schema, called the way the reference's own
test calls it (tests/python/sgl_kernel_npu/test_mla_preprocess.py:547-578: FRACTAL_NZ weights, keyword arguments,
caches passed both as inputs and as kv_cache_out*)."""
import pytest
import torch

from oracle import mla_oracle as orc

# csrc/pytorch_extensions.cpp:105-115, whitespace-normalised by torch's schema printer
REFERENCE_SCHEMA = (
    "npu::mla_preprocess(Tensor hiddenState, Tensor gamma0, Tensor beta0, Tensor wdqkv, Tensor descale0, Tensor gamma1, "
    "Tensor beta1, Tensor wuq, Tensor descale1, Tensor gamma2, Tensor cos, Tensor sin, Tensor wuk, Tensor kv_cache, "
    "Tensor kv_cache_rope, Tensor slotmapping, Tensor quant_scale0, Tensor quant_offset0, Tensor bias0, Tensor quant_scale1, "
    "Tensor quant_offset1, Tensor bias1, *, Tensor? ctkv_scale=None, Tensor? q_nope_scale=None, str? cache_mode=None, "
    "str? quant_mode=None, Tensor(a!) q_out0, Tensor(b!) kv_cache_out0, Tensor(c!) q_out1, Tensor(d!) kv_cache_out1) "
    "-> (Tensor(a!), Tensor(b!), Tensor(c!), Tensor(d!))")


def load():
    import sgl_kernel_npu
    assert sgl_kernel_npu.OP_LIBRARY_LOADED, "operator library not built (python -c 'import __graft_entry__ as g; g.build()')"
    return torch.ops.npu


def test_import_registers_the_reference_schema():
    npu = load()
    assert str(npu.mla_preprocess.default._schema) == REFERENCE_SCHEMA
    assert "sm_100a" in npu.sgl_kernel_npu_version()


def test_no_cpu_fallback():
    npu = load()
    c = orc.gen_case(N=1, H=1, D=128, seed=0, num_blocks=1).t
    with pytest.raises((NotImplementedError, RuntimeError)):
        npu.mla_preprocess(c["hiddenState"], c["gamma0"], c["beta0"], c["wdqkv"], c["descale0"], c["gamma1"], c["beta1"], c["wuq"],
                           c["descale1"], c["gamma2"], c["cos"], c["sin"], c["wuk"], c["kv_cache"], c["kv_cache_rope"], c["slotmapping"],
                           c["quant_scale0"], c["quant_offset0"], c["bias0"], c["quant_scale1"], c["quant_offset1"], c["bias1"],
                           q_out0=c["q_out0"], kv_cache_out0=c["kv_cache"], q_out1=c["q_out1"], kv_cache_out1=c["kv_cache_rope"])


def call_op(npu, g, cache_mode, quant_mode):
    return npu.mla_preprocess(
        g["hiddenState"], g["gamma0"], g["beta0"], g["wdqkv"], g["descale0"], g["gamma1"], g["beta1"], g["wuq"], g["descale1"],
        g["gamma2"], g["cos"], g["sin"], g["wuk"], g["kv_cache"], g["kv_cache_rope"], g["slotmapping"],
        quant_scale0=g["quant_scale0"], quant_offset0=g["quant_offset0"], bias0=g["bias0"], quant_scale1=g["quant_scale1"],
        quant_offset1=g["quant_offset1"], bias1=g["bias1"], ctkv_scale=g["ctkv_scale"], q_nope_scale=g["q_nope_scale"],
        cache_mode=cache_mode, quant_mode=quant_mode, q_out0=g["q_out0"], kv_cache_out0=g["kv_cache"], q_out1=g["q_out1"],
        kv_cache_out1=g["kv_cache_rope"])


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [(16, 32, 7168, torch.bfloat16, "krope_ctkv"), (31, 32, 7168, torch.bfloat16, "int8_nzcache"),
                                 (16, 8, 7168, torch.float16, "krope_ctkv"), (140, 16, 6144, torch.bfloat16, "nzcache")],
                         ids=lambda c: f"N{c[0]}-H{c[1]}-{c[4]}-{str(c[3])[6:]}")
def test_op_matches_oracle(cfg, cuda_device):
    npu = load()
    import mla_harness as hz
    n, h, d, dt, mode = cfg
    qm = "per_tensor_quant_asymm"
    case = orc.gen_case(N=n, H=h, D=d, dtype=dt, cache_mode=mode, seed=42, num_blocks=12, neg_slot_frac=0.1)
    g = hz.to_device(case, cuda_device)
    out = call_op(npu, g, mode, qm)
    torch.cuda.synchronize()
    # returns the four out tensors themselves (op_host/mla_preprocess.cpp:703)
    assert out[0].data_ptr() == g["q_out0"].data_ptr() and out[1].data_ptr() == g["kv_cache"].data_ptr()
    assert out[2].data_ptr() == g["q_out1"].data_ptr() and out[3].data_ptr() == g["kv_cache_rope"].data_ptr()
    orc.run_case(case, quant_mode=qm)
    assert torch.equal(g["q_out1"].cpu(), case.t["q_out1"])
    assert torch.equal(g["kv_cache"].cpu(), case.t["kv_cache"])
    assert torch.equal(g["kv_cache_rope"].cpu(), case.t["kv_cache_rope"])
    if mode == "int8_nzcache":
        assert (g["q_out0"].cpu().int() - case.t["q_out0"].int()).abs().max().item() <= 1
    else:
        ref = case.t["q_out0"].float()
        assert ((g["q_out0"].cpu().float() - ref).abs() / ref.abs().clamp_min(1e-3)).max().item() <= 1e-2
    # second call reuses the cached repack of the same weight tensors and gives identical bytes
    q0_first = g["q_out0"].clone()
    call_op(npu, g, mode, qm)
    torch.cuda.synchronize()
    assert torch.equal(g["q_out0"], q0_first)


@pytest.mark.gpu
def test_op_errors_like_the_reference(cuda_device):
    npu = load()
    import mla_harness as hz
    case = orc.gen_case(N=2, H=2, D=256, seed=1, num_blocks=2)
    g = hz.to_device(case, cuda_device)
    with pytest.raises(RuntimeError, match="Unsupported cache_mode value: 'bogus'"):   # op_host/mla_preprocess.cpp:619
        call_op(npu, g, "bogus", "per_tensor_quant_asymm")
    with pytest.raises(RuntimeError, match="Unsupported quant_mode value: 'nope'"):
        call_op(npu, g, "krope_ctkv", "nope")
    bad = dict(g)
    bad["kv_cache"] = g["kv_cache"].clone()   # kv_cache_out0 no longer aliases kv_cache... passes clone as both -> fine
    bad["slotmapping"] = g["slotmapping"].to(torch.int64)
    with pytest.raises(RuntimeError, match="slotmapping must be"):
        call_op(npu, bad, "krope_ctkv", "per_tensor_quant_asymm")
    # default quant_mode is per_token_quant_symm (op_host/mla_preprocess.cpp:635): runs, and differs from per-tensor
    out_tok = call_op(npu, g, None, None)[0].clone()
    out_ten = call_op(npu, g, None, "per_tensor_quant_asymm")[0].clone()
    torch.cuda.synchronize()
    assert not torch.equal(out_tok, out_ten)


@pytest.mark.gpu
def test_explicit_repack_op(cuda_device):
    npu = load()
    w = torch.randint(-128, 128, (192, 1536), dtype=torch.int8)
    km = npu.mla_preprocess_repack_weight(orc.nz_pack(w).to(cuda_device))
    assert torch.equal(km.cpu(), w)
