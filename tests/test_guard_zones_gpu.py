"""Out-of-bounds writes: every buffer the operator WRITES (both outputs, both caches, the workspace) is carved out of a larger
allocation with 64 KiB of 0xA5 on either side; after calls of every launch shape the guard bytes must be untouched.
(compute-sanitizer is closed on the GPU pool this repository is developed on — profiles/r02/sanitizer_closed.txt — so writes
past a buffer are looked for this way; TMA stores are clipped by their tensor maps, the plain stores — quantised activations,
a1, the paged-cache scatter, the split q_rope stores — are what this guards.)"""
import pytest
import torch

from oracle import mla_oracle as orc
import mla_harness as hz

pytestmark = pytest.mark.gpu
GUARD = 65536
PT, TOK = "per_tensor_quant_asymm", "per_token_quant_symm"


def guarded(t, device):
    nbytes = t.numel() * t.element_size()
    buf = torch.full((nbytes + 2 * GUARD,), 0xA5, dtype=torch.uint8, device=device)
    view = buf[GUARD:GUARD + nbytes].view(t.dtype).reshape(t.shape)
    view.copy_(t.to(device))
    return view, buf, nbytes


def intact(buf, nbytes):
    return bool((buf[:GUARD] == 0xA5).all()) and bool((buf[GUARD + nbytes:] == 0xA5).all())


SHAPES = [  # N, H, D, cache_mode, quant_mode, debug flags
    (8, 16, 7168, "krope_ctkv", PT, 0),            # 4 CTAs per (head, tile) unit
    (16, 2, 2048, "int8_nzcache", PT, 0),
    (1, 32, 7168, "nzcache", PT, 0),
    (127, 64, 7168, "int8_nzcache", PT, 0),        # 2 CTAs per unit, ragged last rows
    (129, 16, 7168, "nzcache", TOK, 0),            # one row in the second token tile
    (257, 5, 2048, "int8_nzcache", PT, 0),
    (300, 128, 7168, "krope_ctkv", PT, 0),         # two token tiles per CTA, partial second super-tile
    (300, 8, 2048, "krope_ctkv", PT, 8192),        # CTA-pair q kernel
    (1100, 3, 2048, "nzcache", PT, 0),             # prefill shapes (dequantising MM1)
    (1100, 3, 2048, "int8_nzcache", TOK, 8192),
]


@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "N%d-H%d-D%d-%s-%s-f%d" % (s[0], s[1], s[2], s[3], s[4][:9], s[5]))
def test_no_write_outside_the_callers_buffers(shape, cuda_device):
    from sgl_kernel_npu import cabi
    n, h, d, mode, qm, dbg = shape
    case = orc.gen_case(N=n, H=h, D=d, dtype=torch.bfloat16, cache_mode=mode, seed=13, num_blocks=max(4, n // 100 + 2), neg_slot_frac=0.1)
    g = hz.to_device(case, cuda_device)
    bufs = {}
    for k in ("q_out0", "q_out1", "kv_cache", "kv_cache_rope"):
        g[k], buf, nb = guarded(case.t[k], cuda_device)
        bufs[k] = (buf, nb)
    g["wdqkv_km"], g["wuq_km"] = cabi.repack_nz_weight(g["wdqkv"]), cabi.repack_nz_weight(g["wuq"])
    old = cabi.set_debug_flags(dbg)
    try:
        p = cabi.build_params(g["hiddenState"], g["wdqkv_km"], g["descale0"], g["gamma1"], g["beta1"], g["wuq_km"], g["descale1"],
                              g["gamma2"], g["cos"], g["sin"], g["wuk"], g["kv_cache"], g["kv_cache_rope"], g["slotmapping"],
                              g["quant_scale0"], g["quant_offset0"], g["bias0"], g["quant_scale1"], g["quant_offset1"], g["bias1"],
                              g["ctkv_scale"], g["q_nope_scale"], mode, qm, g["q_out0"], g["q_out1"])
        ws_bytes = cabi.workspace_layout(p).total_bytes
        ws_buf = torch.full((ws_bytes + 2 * GUARD,), 0xA5, dtype=torch.uint8, device=cuda_device)
        ws = ws_buf[GUARD:GUARD + ws_bytes]
        for _ in range(2):
            cabi.launch(p, ws)
        torch.cuda.synchronize(cuda_device)
    finally:
        cabi.set_debug_flags(old)
    bufs["workspace"] = (ws_buf, ws_bytes)
    for k, (buf, nb) in bufs.items():
        assert intact(buf, nb), f"bytes outside {k} were written"
    # and the call did its work inside them
    assert int((g["q_out1"].float() != 0).sum()) > 0
