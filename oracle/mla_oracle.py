"""CPU oracle for the ``mla_preprocess`` hot path (DeepSeek-V3 MLA attention prologue).

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product: the
only permitted importers are ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``.  The product path is
the CUDA library behind ``include/mla_preprocess_b200.h`` and it fails loudly when
that library is missing; it never routes through this file.

What this restates
------------------
The reference publishes its arithmetic contract as a pure-PyTorch golden,
``golden2_pytorch`` in ``tests/python/sgl_kernel_npu/test_mla_preprocess.py:407-484``
(helpers ``:57-107``), and checks the AscendC kernel against it with
``allclose(atol=rtol=1e-3)`` (``:712-744``).  This module is a from-scratch
restatement of that contract, organised as explicit stages with taps so that the
int8 activations, int32 accumulators and cache placement can be compared bit for bit.

Pinning
-------
* ``per_tensor_quant_asymm`` is pinned: ``oracle/make_golden.py`` imports the
  reference's own generator + ``golden2_pytorch`` in the build container and writes
  ``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks this module against
  those fixtures (``var_mode="torch"`` must reproduce them bit for bit).
* ``per_token_quant_symm`` is "parity unpinned": the reference never instantiates
  that branch (``csrc/mla_preprocess/op_kernel/mla_preprocess_kernel.cpp:140-237``
  has no tiling key for it) and no reference test exercises it.  It is restated from
  the uninstantiated template branches cited on each function below.

The one order-dependent quantity
--------------------------------
Everything in the contract is element-wise IEEE arithmetic or exact integer
arithmetic except the two RMSNorm row sums (``mean(x^2)``), whose fp32 value depends
on the summation order of whatever computes it (torch's CPU reduction is itself
ISA-dependent).  ``var_mode="exact"`` (default) defines the sum as the exactly
accumulated (fp64) sum of squares rounded once to fp32; that definition is
order-free, so the CUDA kernels can be held to bit-exactness against it.
``var_mode="torch"`` calls ``Tensor.pow(2).mean()`` exactly as the golden does.
"""
from __future__ import annotations

import dataclasses
from typing import Dict, Optional

import numpy as np
import torch

# Shape constants of the path: csrc/mla_preprocess/op_kernel/mla_preprocess.h:55-64,
# tests/python/sgl_kernel_npu/test_mla_preprocess.py:18-28.
MM1_OUT = 2112          # [k_nope 512 | k_pe 64 | q_a 1536]  (kv first: test_mla:343-345)
KV_LORA = 512
ROPE = 64
Q_LORA = 1536
Q_NOPE = 128
Q_HEAD = Q_NOPE + ROPE  # 192, per head [nope 128 | rope 64] (test_mla:370-372)
EPS = 1e-6
NZ_BLOCK = 128          # cache block size in the NZ modes (bf16.hpp:2549-2550)

CACHE_MODES = {"krope_ctkv": 1, "int8_nzcache": 2, "nzcache": 3}      # op_host/mla_preprocess.cpp:605-606
QUANT_MODES = {"per_tensor_quant_asymm": 0, "per_token_quant_symm": 1}  # op_host/mla_preprocess.cpp:608-611
DEFAULT_CACHE_MODE = "krope_ctkv"                                      # op_host/mla_preprocess.cpp:634
DEFAULT_QUANT_MODE = "per_token_quant_symm"                            # op_host/mla_preprocess.cpp:635


# --------------------------------------------------------------------------- formats

def nz_pack(w_nd: torch.Tensor, c0: int = 32) -> torch.Tensor:
    """ND ``[Nout, K]`` -> FRACTAL_NZ ``[1, K/c0, Nout, c0]``.

    Element (n, k) lands at ``((k // c0) * Nout + n) * c0 + k % c0``
    (test_mla:38-54 ``transdata`` with block (16, 32); kernel side ``GetOffsetB``
    mla_preprocess_mix_bf16.hpp:1581-1588).  Nout is padded to 16, K to c0.
    """
    n, k = w_nd.shape
    n_pad, k_pad = -(-n // 16) * 16, -(-k // c0) * c0
    if (n_pad, k_pad) != (n, k):
        w_nd = torch.nn.functional.pad(w_nd, (0, k_pad - k, 0, n_pad - n))
    return w_nd.reshape(n_pad, k_pad // c0, c0).permute(1, 0, 2).reshape(1, k_pad // c0, n_pad, c0).contiguous()


def nz_unpack(w_nz: torch.Tensor) -> torch.Tensor:
    """Inverse of :func:`nz_pack`: ``[1, K/c0, Nout, c0]`` -> ``[Nout, K]``."""
    _, kb, n, c0 = w_nz.shape
    return w_nz.reshape(kb, n, c0).permute(1, 0, 2).reshape(n, kb * c0).contiguous()


def pack_descale_fp16_path(descale_f32: np.ndarray) -> torch.Tensor:
    """fp16 path: descale travels as int64 = (fp32 bits & 0xFFFFE000) | 1<<46 (test_mla:110-119)."""
    bits = np.ascontiguousarray(descale_f32, dtype=np.float32).view(np.uint32) & np.uint32(0xFFFFE000)
    packed = bits.astype(np.uint64) | np.uint64(1 << 46)
    return torch.from_numpy(packed.view(np.int64).copy())


def unpack_descale(descale: torch.Tensor) -> torch.Tensor:
    """Return fp32 descale; int64 carriers decode to the fp32 held in their low 32 bits (test_mla:101-103)."""
    if descale.dtype == torch.int64:
        return descale.to(torch.int32).view(torch.float32)
    return descale.to(torch.float32)


def cache_shapes(cache_mode: str, num_blocks: int, block_size: int = NZ_BLOCK):
    """Cache tensor shapes per mode (csrc/mla_preprocess/README.md:27-28, 44-47)."""
    if cache_mode == "krope_ctkv":
        return (num_blocks, block_size, 1, KV_LORA), (num_blocks, block_size, 1, ROPE)
    if cache_mode == "int8_nzcache":
        return (num_blocks, KV_LORA // 32, block_size, 32), (num_blocks, ROPE // 16, block_size, 16)
    if cache_mode == "nzcache":
        return (num_blocks, KV_LORA // 16, block_size, 16), (num_blocks, ROPE // 16, block_size, 16)
    raise ValueError(f"Unsupported cache_mode value: '{cache_mode}'")


def cache_row_index(slot: int, width: int, c0: int, nz: bool, block_size: int = NZ_BLOCK) -> np.ndarray:
    """Flat element offsets of one token's ``width`` cache entries.

    ND (mode 1): ``slot*width + c`` (bf16.hpp:2546-2547, 2677-2681).
    NZ (modes 2/3): block = slot // 128, r = slot % 128,
    ``block*128*width + (c // c0)*128*c0 + r*c0 + c % c0`` (bf16.hpp:2644-2675;
    reader test_mla:122-135).
    """
    c = np.arange(width, dtype=np.int64)
    if not nz:
        return int(slot) * width + c
    blk, r = divmod(int(slot), block_size)
    return blk * block_size * width + (c // c0) * block_size * c0 + r * c0 + c % c0


def read_cache_row(cache: torch.Tensor, slot: int, width: int, c0: int, nz: bool) -> torch.Tensor:
    idx = torch.from_numpy(cache_row_index(slot, width, c0, nz))
    return cache.reshape(-1)[idx]


# ------------------------------------------------------------------------ generator

@dataclasses.dataclass
class MlaCase:
    N: int
    H: int
    D: int
    dtype: torch.dtype
    cache_mode: str
    t: Dict[str, torch.Tensor]  # op-surface names (csrc/pytorch_extensions.cpp:105-115)


def gen_case(N=1, H=16, D=7168, dtype=torch.bfloat16, cache_mode="krope_ctkv", seed=42,
             num_blocks=1078, block_size=NZ_BLOCK, neg_slot_frac=0.0, full_range_weights=False,
             slot_runs=False) -> MlaCase:
    """Synthetic inputs with the reference generator's distributions and draw order.

    Follows test_mla:140-321 draw by draw (same ``np.random.seed``, same sequence of
    ``np.random`` calls, same float64->dtype conversions) so that, for equal
    arguments, the tensors are bit-identical to the reference's ``dataDict`` —
    ``tests/test_oracle_golden.py`` checks this against hashes taken from the
    reference generator itself.  Weights are returned in ND layout (``wdqkv_nd``,
    ``wuq_nd``) and in the FRACTAL_NZ layout the op consumes (``wdqkv``, ``wuq``;
    test_mla:537-544).  Extensions beyond the reference (applied after all reference
    draws, so they do not disturb them): ``neg_slot_frac`` marks a fraction of tokens
    with slot -1; ``full_range_weights`` redraws int8 weights over [-128, 127];
    ``slot_runs`` replaces the unique random slots by per-request contiguous runs at
    random block offsets (a decode batch of requests of ragged length, SURVEY.md §8d "C5").
    """
    torch.manual_seed(seed)
    np.random.seed(seed)
    rnd = np.random

    def as_t(a, dt):
        return torch.from_numpy(a).to(dt)

    t: Dict[str, torch.Tensor] = {}
    t["hiddenState"] = as_t(rnd.uniform(-2.0, 2.0, size=(N, D)), dtype)
    t["gamma0"] = torch.ones([D], dtype=dtype)            # accepted, never read
    t["beta0"] = torch.zeros([D], dtype=dtype)
    t["quant_scale0"] = as_t(rnd.uniform(-2.0, 2.0, size=(1)), dtype)
    t["quant_offset0"] = as_t(rnd.uniform(-128.0, 127.0, size=(1)), torch.int8)
    t["wdqkv_nd"] = as_t(rnd.uniform(-2.0, 2.0, size=(MM1_OUT, D)), torch.int8)
    t["bias0"] = as_t(rnd.randint(-10, 10, (MM1_OUT)).astype(np.int32), torch.int32)
    descale0 = (rnd.rand(MM1_OUT) / 1000).astype(np.float32)
    t["gamma1"] = as_t(rnd.uniform(-1.0, 1.0, size=(Q_LORA)), dtype)
    t["beta1"] = as_t(rnd.randint(-2, 2, (Q_LORA)).astype(np.float16), dtype)
    t["gamma2"] = as_t(rnd.uniform(-1.0, 1.0, size=(KV_LORA)), dtype)
    t["quant_scale1"] = as_t(rnd.uniform(-2.0, 2.0, size=(1)), dtype)
    t["quant_offset1"] = as_t(rnd.uniform(-128.0, 127.0, size=(1)), torch.int8)
    t["wuq_nd"] = as_t(rnd.uniform(-2.0, 2.0, size=(H * Q_HEAD, Q_LORA)), torch.int8)
    t["bias1"] = as_t(rnd.randint(-10, 10, (H * Q_HEAD)).astype(np.int32), torch.int32)
    descale1 = (rnd.rand(H * Q_HEAD) / 1000).astype(np.float32)
    t["wuk"] = as_t(rnd.uniform(-2.0, 2.0, size=(H, Q_NOPE, KV_LORA)), dtype)
    t["sin"] = as_t(rnd.uniform(-1.0, 1.0, size=(N, ROPE)), dtype)
    t["cos"] = as_t(rnd.uniform(-1.0, 1.0, size=(N, ROPE)), dtype)
    slots = rnd.choice(num_blocks * block_size, N, replace=False).astype(np.int32)
    t["ctkv_scale"] = as_t(rnd.uniform(-2.0, 2.0, size=(1)), dtype)
    t["q_nope_scale"] = as_t(rnd.uniform(-1.0, 1.0, size=(H)), dtype)

    if dtype == torch.float16:
        t["descale0"] = pack_descale_fp16_path(descale0)
        t["descale1"] = pack_descale_fp16_path(descale1)
    else:
        t["descale0"] = torch.from_numpy(descale0)
        t["descale1"] = torch.from_numpy(descale1)

    # ---- extensions (after every reference draw) ----
    if full_range_weights:
        t["wdqkv_nd"] = as_t(rnd.randint(-128, 128, (MM1_OUT, D)).astype(np.int8), torch.int8)
        t["wuq_nd"] = as_t(rnd.randint(-128, 128, (H * Q_HEAD, Q_LORA)).astype(np.int8), torch.int8)
    if slot_runs:
        # requests of 1..32 tokens; each takes a run of consecutive slots that starts at a random offset of a block of
        # its own (so runs never collide and may straddle a block boundary into the request's second block)
        lens, left = [], N
        while left > 0:
            n_req = int(min(left, rnd.randint(1, 33)))
            lens.append(n_req)
            left -= n_req
        if 2 * len(lens) > num_blocks:
            raise ValueError(f"slot_runs: {len(lens)} requests need {2 * len(lens)} cache blocks, have {num_blocks}")
        first_blocks = rnd.permutation(num_blocks // 2)[:len(lens)] * 2
        pos = 0
        for n_req, blk in zip(lens, first_blocks):
            start = int(blk) * block_size + int(rnd.randint(0, block_size))
            slots[pos:pos + n_req] = np.arange(start, start + n_req, dtype=np.int32)
            pos += n_req
    if neg_slot_frac > 0 and N > 1:
        k = max(1, int(round(N * neg_slot_frac)))
        slots[rnd.choice(N, k, replace=False)] = -1
    t["slotmapping"] = torch.from_numpy(slots)

    t["wdqkv"] = nz_pack(t["wdqkv_nd"])
    t["wuq"] = nz_pack(t["wuq_nd"])

    s0, s1 = cache_shapes(cache_mode, num_blocks, block_size)
    q0_dtype = torch.int8 if cache_mode == "int8_nzcache" else dtype
    t["kv_cache"] = torch.zeros(s0, dtype=q0_dtype)
    t["kv_cache_rope"] = torch.zeros(s1, dtype=dtype)
    t["q_out0"] = torch.zeros((N, H, KV_LORA), dtype=q0_dtype)
    t["q_out1"] = torch.zeros((N, H, ROPE), dtype=dtype)
    return MlaCase(N=N, H=H, D=D, dtype=dtype, cache_mode=cache_mode, t=t)


# ------------------------------------------------------------------ arithmetic stages

def _f16_saturate_round_to_i8(x_f32: torch.Tensor) -> torch.Tensor:
    """fp32 -> fp16 (RNE) -> clamp[-128,127] -> round-half-even -> int8 (test_mla:81-83; kernel_utils.h:62-88)."""
    h = x_f32.to(torch.float16)
    return torch.round(torch.clamp(h, -128, 127)).to(torch.int8)


def quant_per_tensor(x: torch.Tensor, scale: torch.Tensor, offset: torch.Tensor) -> torch.Tensor:
    """S1/S4 quantiser: ``q = sat_i8(fp16(fp32(x) / fp32(scale) + fp32(offset)))``.

    Golden divides (test_mla:77-83); the AscendC kernel multiplies by 1/scale
    (bf16.hpp:368, 431-435).  The golden is the parity target, so: division.
    """
    y = x.to(torch.float32) / scale.to(torch.float32) + offset.to(torch.float32)
    return _f16_saturate_round_to_i8(y)


def quant_mul(x: torch.Tensor, scale: torch.Tensor) -> torch.Tensor:
    """S9 q_nope re-quantiser (cache_mode 2): multiply by the per-head scale (test_mla:86-92; bf16.hpp:943-956)."""
    return _f16_saturate_round_to_i8(x.to(torch.float32) * scale.to(torch.float32))


def quant_per_token(x_f32: torch.Tensor):
    """Per-token symmetric quantiser (UNPINNED): ``s = max|x| / 127; q = sat_i8(fp16(x * (1/s)))``.

    bf16.hpp:437-475 (Quant) and :739-777 (RmsNormQuant) — the kernel forms the
    reciprocal once per row and multiplies.  Returns (int8, fp32 scale per row).
    """
    amax = x_f32.abs().amax(dim=-1, keepdim=True)
    s = amax / 127.0
    inv = 1.0 / s
    return _f16_saturate_round_to_i8(x_f32 * inv), s.squeeze(-1)


def int8_gemm_i32(a_i8: torch.Tensor, w_i8: torch.Tensor) -> torch.Tensor:
    """Exact ``int32 = a[M,K] . w[N,K]^T`` (test_mla:96-98; MMAD int8 bf16.hpp:1838-1947).

    Evaluated through a float BLAS whose accumulator provably holds every partial
    sum exactly: fp32 if K*max|a|*max|w| < 2^24, else fp64 (< 2^53 always).
    """
    k = a_i8.shape[-1]
    bound = k * int(a_i8.abs().max().item() if a_i8.numel() else 0) * int(w_i8.abs().max().item() if w_i8.numel() else 0)
    ft = torch.float32 if bound < (1 << 24) else torch.float64
    out = torch.empty((a_i8.shape[0], w_i8.shape[0]), dtype=torch.int32)
    step = 4096
    a = a_i8.to(ft)
    for n0 in range(0, w_i8.shape[0], step):
        out[:, n0:n0 + step] = (a @ w_i8[n0:n0 + step].to(ft).t()).to(torch.int32)
    return out


def dequant(acc_i32, bias_i32, descale_f32, out_dtype, tok_scale=None):
    """``act_rne(fp32(acc + bias) * descale[n])`` (test_mla:100-107; bf16.hpp:2204-2330).

    Per-token mode: no bias, extra ``* tokScale[m]`` (bf16.hpp:2281-2287).
    """
    if tok_scale is None:
        y = (acc_i32 + bias_i32).to(torch.float32) * descale_f32
    else:
        y = acc_i32.to(torch.float32) * descale_f32 * tok_scale.unsqueeze(-1)
    return y.to(out_dtype)


def sum_squares_f32(x: torch.Tensor, var_mode: str) -> torch.Tensor:
    """fp32 value of ``mean(x^2)`` over the last dim (test_mla:71; bf16.hpp:697-712)."""
    n = x.shape[-1]
    xf = x.to(torch.float32)
    if var_mode == "torch":
        return xf.pow(2).mean(dim=-1, keepdim=True)
    s = xf.to(torch.float64).pow(2).sum(dim=-1, keepdim=True).to(torch.float32)
    return s / float(n)


def rms_norm_f32(x, gamma, var_mode="exact", beta=None):
    """``fp32(x) * rsqrt(mean(x^2)+eps) * fp32(gamma) [+ fp32(beta)]`` in fp32 (test_mla:70-74, 432-433).

    ``rsqrt`` is evaluated as two correctly rounded operations, ``1 / sqrt(.)``
    (what torch's CPU ``rsqrt`` returns on the golden's inputs), mirrored on the
    CUDA path by ``__fdiv_rn(1, __fsqrt_rn(.))``.
    """
    var = sum_squares_f32(x, var_mode)
    # numpy's float32 sqrt and divide are the correctly rounded IEEE operations on every host; torch.sqrt on CPU is not
    # (observed: 0x40257d15 -> 0x3fcdd3ea instead of 0x3fcdd3eb), which flips one bf16 rounding in ~10^6 outputs
    v = (var + EPS).numpy().astype(np.float32)
    r = torch.from_numpy(np.float32(1.0) / np.sqrt(v, dtype=np.float32))
    y = x.to(torch.float32) * r * gamma.to(torch.float32)
    if beta is not None:
        y = y + beta.to(torch.float32)
    return y


def rope_half(x, cos, sin):
    """NeoX rotate-half over 64 lanes: ``x*cos + cat(-x[32:], x[:32])*sin`` in fp32, RNE to act (test_mla:57-67)."""
    xf = x.to(torch.float32)
    h = xf.shape[-1] // 2
    rot = torch.cat([-xf[..., h:], xf[..., :h]], dim=-1)
    c, s = cos.to(torch.float32), sin.to(torch.float32)
    while c.dim() < xf.dim():
        c, s = c.unsqueeze(1), s.unsqueeze(1)
    return (xf * c + rot * s).to(x.dtype)


def bmm_absorb(q_nope: torch.Tensor, wuk: torch.Tensor) -> torch.Tensor:
    """``q_out0[:,h,:] = act_rne(q_nope[:,h,:] . wuk[h])`` (test_mla:453-455; bf16.hpp:1127-1354).

    Accumulated in fp64 so the value rounded to the activation type is the correctly
    rounded dot product; torch's and the tensor core's fp32 accumulations are each
    within 1 ulp(act) of it, which is inside the stated 1e-2 tolerance.
    """
    o = torch.bmm(q_nope.transpose(0, 1).to(torch.float64), wuk.to(torch.float64))
    return o.transpose(0, 1).to(torch.float32).to(q_nope.dtype).contiguous()


# --------------------------------------------------------------------- the operator

def mla_preprocess(hiddenState, gamma0, beta0, wdqkv, descale0, gamma1, beta1, wuq, descale1, gamma2,
                   cos, sin, wuk, kv_cache, kv_cache_rope, slotmapping,
                   quant_scale0, quant_offset0, bias0, quant_scale1, quant_offset1, bias1, *,
                   ctkv_scale=None, q_nope_scale=None, cache_mode=None, quant_mode=None,
                   q_out0, kv_cache_out0, q_out1, kv_cache_out1,
                   var_mode: str = "exact", taps: Optional[dict] = None):
    """Oracle with the operator's own argument list (csrc/pytorch_extensions.cpp:105-115).

    Outputs are written in place into ``q_out0``, ``kv_cache_out0``, ``q_out1``,
    ``kv_cache_out1`` and returned as a 4-tuple, as the op does
    (op_host/mla_preprocess.cpp:703).  ``wdqkv``/``wuq`` are FRACTAL_NZ.
    ``gamma0``/``beta0`` are accepted and ignored: step 1 is quantisation only
    (bf16.hpp:2762, 424-486; golden test_mla:416-418).  ``taps`` (if a dict)
    receives every intermediate: a0, acc0, y0, a1, acc1, y1, k_nope, k_pe.
    """
    cache_mode = DEFAULT_CACHE_MODE if cache_mode is None else cache_mode
    quant_mode = DEFAULT_QUANT_MODE if quant_mode is None else quant_mode
    if cache_mode not in CACHE_MODES:
        raise RuntimeError(f"Unsupported cache_mode value: '{cache_mode}'")   # op_host/mla_preprocess.cpp:619
    if quant_mode not in QUANT_MODES:
        raise RuntimeError(f"Unsupported quant_mode value: '{quant_mode}'")
    per_token = QUANT_MODES[quant_mode] == 1
    act = hiddenState.dtype
    N, H = hiddenState.shape[0], wuk.shape[0]
    taps = {} if taps is None else taps

    w0 = nz_unpack(wdqkv)[:MM1_OUT, :hiddenState.shape[1]]
    w1 = nz_unpack(wuq)[:H * Q_HEAD, :Q_LORA]
    d0, d1 = unpack_descale(descale0), unpack_descale(descale1)

    # S1 + S2: quantise hidden (no RMSNorm), MM1, dequant (test_mla:416-425)
    if per_token:
        a0, s0 = quant_per_token(hiddenState.to(torch.float32))
    else:
        a0, s0 = quant_per_tensor(hiddenState, quant_scale0, quant_offset0), None
    acc0 = int8_gemm_i32(a0, w0)
    y0 = dequant(acc0, bias0, d0, act, tok_scale=s0)
    taps.update(a0=a0, acc0=acc0, y0=y0)

    # S3 split: kv first (test_mla:427-429)
    k_nope_in, k_pe_in, q_a = y0[:, :KV_LORA], y0[:, KV_LORA:KV_LORA + ROPE], y0[:, KV_LORA + ROPE:]

    # S4 + S5: q_a RMSNorm(+beta) -> quant -> MM2 -> dequant (test_mla:432, 436-446)
    q_n = rms_norm_f32(q_a, gamma1, var_mode, beta=beta1)
    if per_token:
        a1, s1 = quant_per_token(q_n)
    else:
        a1, s1 = quant_per_tensor(q_n, quant_scale1, quant_offset1), None
    acc1 = int8_gemm_i32(a1, w1)
    y1 = dequant(acc1, bias1, d1, act, tok_scale=s1).view(N, H, Q_HEAD)
    taps.update(a1=a1, acc1=acc1, y1=y1)

    # S6 + S7 (test_mla:448-463)
    qo0 = bmm_absorb(y1[..., :Q_NOPE], wuk)
    qo1 = rope_half(y1[..., Q_NOPE:], cos, sin)
    k_pe = rope_half(k_pe_in, cos, sin)

    # S8 + S9 (test_mla:433, 465-476)
    k_nope_f32 = rms_norm_f32(k_nope_in, gamma2, var_mode)
    if cache_mode == "int8_nzcache":
        if ctkv_scale is None or q_nope_scale is None:
            raise RuntimeError("ctkv_scale and q_nope_scale are required for cache_mode 'int8_nzcache'")
        qo0 = quant_mul(qo0, q_nope_scale.reshape(1, H, 1))
        k_nope = quant_per_tensor(k_nope_f32, ctkv_scale, torch.zeros(1))
    else:
        k_nope = k_nope_f32.to(act)
    taps.update(k_nope=k_nope, k_pe=k_pe)

    q_out0.copy_(qo0)
    q_out1.copy_(qo1)

    # S10: ReshapeAndCache scatter; slot == -1 rows are skipped (bf16.hpp:2528-2531)
    nz = cache_mode != "krope_ctkv"
    c0_nope = 32 if cache_mode == "int8_nzcache" else 16
    flat0, flat1 = kv_cache_out0.reshape(-1), kv_cache_out1.reshape(-1)
    assert flat0.data_ptr() == kv_cache_out0.data_ptr() and flat1.data_ptr() == kv_cache_out1.data_ptr()
    slots = slotmapping.to(torch.int64).tolist()
    for tkn, slot in enumerate(slots):
        if slot == -1:
            continue
        flat0[torch.from_numpy(cache_row_index(slot, KV_LORA, c0_nope, nz))] = k_nope[tkn]
        flat1[torch.from_numpy(cache_row_index(slot, ROPE, 16, nz))] = k_pe[tkn]
    return q_out0, kv_cache_out0, q_out1, kv_cache_out1


def run_case(case: MlaCase, quant_mode="per_tensor_quant_asymm", var_mode="exact", taps=None):
    """Convenience: run the oracle on a generated case (mutates the case's out tensors)."""
    t = case.t
    return mla_preprocess(
        t["hiddenState"], t["gamma0"], t["beta0"], t["wdqkv"], t["descale0"], t["gamma1"], t["beta1"],
        t["wuq"], t["descale1"], t["gamma2"], t["cos"], t["sin"], t["wuk"], t["kv_cache"], t["kv_cache_rope"],
        t["slotmapping"], t["quant_scale0"], t["quant_offset0"], t["bias0"], t["quant_scale1"],
        t["quant_offset1"], t["bias1"], ctkv_scale=t["ctkv_scale"], q_nope_scale=t["q_nope_scale"],
        cache_mode=case.cache_mode, quant_mode=quant_mode, q_out0=t["q_out0"], kv_cache_out0=t["kv_cache"],
        q_out1=t["q_out1"], kv_cache_out1=t["kv_cache_rope"], var_mode=var_mode, taps=taps)
