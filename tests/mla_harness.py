"""Shared GPU-parity harness: run one generated case through the oracle (CPU) and the CUDA
library (through the C ABI, via sgl_kernel_npu.cabi) and compare stage by stage."""
import torch

from oracle import mla_oracle as orc

TO_DEVICE_KEYS = ["hiddenState", "gamma0", "beta0", "wdqkv", "descale0", "gamma1", "beta1", "wuq", "descale1", "gamma2",
                  "cos", "sin", "wuk", "kv_cache", "kv_cache_rope", "slotmapping", "quant_scale0", "quant_offset0",
                  "bias0", "quant_scale1", "quant_offset1", "bias1", "ctkv_scale", "q_nope_scale", "q_out0", "q_out1"]


def to_device(case, device):
    return {k: case.t[k].to(device) for k in TO_DEVICE_KEYS}


def run_cuda(case, device, quant_mode="per_tensor_quant_asymm", debug=True, cache_fill=None, extra_flags=0, flags=0):
    """Returns (gpu tensors dict, taps dict on CPU).
    extra_flags: further mlapo_set_debug_flags bits (8192 = CTA-pair q kernel wherever n > 128); flags: MlapoParams.flags.
    debug=True: synchronise + check after every kernel (which also serialises them) and write the y0 tap; debug="tap": only the
    y0 tap, the kernels stay chained by programmatic dependent launch exactly as in production; debug=False: nothing."""
    from sgl_kernel_npu import cabi
    g = to_device(case, device)
    if cache_fill is not None:
        g["kv_cache"].fill_(cache_fill)
        g["kv_cache_rope"].fill_(cache_fill)
    g["wdqkv_km"] = cabi.repack_nz_weight(g["wdqkv"])
    g["wuq_km"] = cabi.repack_nz_weight(g["wuq"])
    old = cabi.set_debug_flags((2 if debug == "tap" else 3 if debug else 0) | extra_flags)
    try:
        p = cabi.build_params(g["hiddenState"], g["wdqkv_km"], g["descale0"], g["gamma1"], g["beta1"], g["wuq_km"],
                              g["descale1"], g["gamma2"], g["cos"], g["sin"], g["wuk"], g["kv_cache"], g["kv_cache_rope"],
                              g["slotmapping"], g["quant_scale0"], g["quant_offset0"], g["bias0"], g["quant_scale1"],
                              g["quant_offset1"], g["bias1"], g["ctkv_scale"], g["q_nope_scale"], case.cache_mode,
                              quant_mode, g["q_out0"], g["q_out1"], flags=flags)
        lay = cabi.workspace_layout(p)
        ws = torch.zeros(lay.total_bytes, dtype=torch.uint8, device=device)
        cabi.launch(p, ws)
        torch.cuda.synchronize(device)
    finally:
        cabi.set_debug_flags(old)
    n, h, d = case.N, case.H, case.D
    taps = {}
    wsc = ws.cpu()

    def view(off, nbytes, dtype, shape):
        return wsc[off:off + nbytes].view(dtype).reshape(shape)

    taps["a0"] = view(lay.a0, n * d, torch.int8, (n, d))
    part = view(lay.partial0, lay.splits0 * n * 2112 * 4, torch.int32, (lay.splits0, n, 2112))
    taps["acc0"] = part.sum(dim=0, dtype=torch.int32)
    taps["a1"] = view(lay.a1, n * 1536, torch.int8, (n, 1536))
    if debug:
        taps["y0"] = view(lay.y0_tap, n * 2112 * 2, case.dtype, (n, 2112))
    taps["splits0"] = lay.splits0
    return g, taps


def extract_cache_rows(case, kv_cache, kv_cache_rope):
    nz = case.cache_mode != "krope_ctkv"
    c0 = 32 if case.cache_mode == "int8_nzcache" else 16
    rows_n, rows_p = [], []
    for s in case.t["slotmapping"].tolist():
        if s < 0:
            rows_n.append(None)
            rows_p.append(None)
            continue
        rows_n.append(orc.read_cache_row(kv_cache, s, 512, c0, nz))
        rows_p.append(orc.read_cache_row(kv_cache_rope, s, 64, 16, nz))
    return rows_n, rows_p


def compare(case, quant_mode, g, taps):
    """Run the oracle and return {stage: (mismatches, total, max_abs_diff)}."""
    otaps = {}
    ocase = orc.MlaCase(case.N, case.H, case.D, case.dtype, case.cache_mode, {k: v.clone() for k, v in case.t.items()})
    orc.run_case(ocase, quant_mode=quant_mode, var_mode="exact", taps=otaps)
    res = {}

    def stat(name, got, want):
        got, want = got.cpu(), want.cpu()
        if got.dtype in (torch.int8, torch.int32):
            diff = (got.to(torch.int64) - want.to(torch.int64)).abs()
        else:
            diff = (got.float() - want.float()).abs()
        bad = (got != want) if got.dtype != torch.float32 else (diff != 0)
        res[name] = (int(bad.sum()), got.numel(), float(diff.max()) if diff.numel() else 0.0)

    stat("a0", taps["a0"], otaps["a0"])
    stat("acc0", taps["acc0"], otaps["acc0"])
    if "y0" in taps:
        stat("y0", taps["y0"], otaps["y0"])
    stat("a1", taps["a1"], otaps["a1"])
    stat("q_out1", g["q_out1"], ocase.t["q_out1"])
    stat("q_out0", g["q_out0"], ocase.t["q_out0"])
    # caches: whole-tensor equality (also proves untouched rows stay untouched)
    stat("kv_cache", g["kv_cache"], ocase.t["kv_cache"])
    stat("kv_cache_rope", g["kv_cache_rope"], ocase.t["kv_cache_rope"])
    q0g, q0o = g["q_out0"].cpu().float(), ocase.t["q_out0"].float()
    denom = q0o.abs().clamp_min(1e-3 if case.cache_mode != "int8_nzcache" else 1.0)
    res["q_out0_relerr"] = float(((q0g - q0o).abs() / denom).max())
    return res, ocase, otaps
