"""CPU oracle for split_qkv_rmsnorm_rope.  TEST INFRASTRUCTURE ONLY — imported by tests/ and nothing else.

numpy restatement of the contract of the reference's Triton kernel (python/sgl_kernel_npu/sgl_kernel_npu/norm/
split_qkv_rmsnorm_rope.py:9-372): fp32 throughout, one rounding to the tensor dtype at the store.  Pinned by
tests/test_split_qkv.py against fixtures produced with the reference TEST's own golden helpers rms_norm / custom_rope
(tests/python/sgl_kernel_npu/test_split_qkv_rmsnorm_rope.py:11-44, script oracle/make_split_qkv_golden.py)."""
import numpy as np
import torch


def split_qkv_rmsnorm_rope(qkv, sin, cos, q_hidden, kv_hidden, head_dim, eps=None, q_weight=None, k_weight=None, q_bias=None,
                           k_bias=None, is_neox_style=True, with_scale=False, gemma=False, cast_norm=False):
    """-> (q, k, v); with_scale adds (mq, mk), float32 arrays |rot * sin| + |n * cos| (|n| outside the rotated dims): the magnitude
    the result was summed from, which bounds a sound tolerance where the rotation cancels."""
    dt = qkv.dtype
    f = lambda t: None if t is None else t.to(torch.float32).numpy().astype(np.float32)
    x = f(qkv)
    b = x.shape[0]
    rd = sin.shape[-1]
    s, c = f(sin).reshape(b, 1, rd), f(cos).reshape(b, 1, rd)
    h = rd // 2

    def one(part, w, bias):
        n = part.reshape(b, -1, head_dim)
        if eps is not None:
            rstd = np.float32(1) / np.sqrt(np.mean(n * n, axis=-1, keepdims=True, dtype=np.float32) + np.float32(eps), dtype=np.float32)
            n = n * rstd * (f(w) + np.float32(1) if gemma else f(w))
            if bias is not None:
                n = n + f(bias)
            if cast_norm:  # the rotation starts from the value as the tensor dtype holds it
                n = torch.from_numpy(n).to(dt).to(torch.float32).numpy()
        rot = n[..., :rd]
        if is_neox_style:
            cat = np.concatenate((-rot[..., h:], rot[..., :h]), axis=-1)
            ss, cc = s, c
        else:
            cat = np.empty_like(rot)
            cat[..., 0::2] = -rot[..., 1::2]
            cat[..., 1::2] = rot[..., 0::2]
            ss, cc = np.repeat(s[..., :h], 2, axis=-1), np.repeat(c[..., :h], 2, axis=-1)
        out = np.concatenate((cat * ss + rot * cc, n[..., rd:]), axis=-1).astype(np.float32)
        scale = np.concatenate((np.abs(cat * ss) + np.abs(rot * cc), np.abs(n[..., rd:])), axis=-1).astype(np.float32)
        return torch.from_numpy(out.reshape(b, -1)).to(dt), scale.reshape(b, -1)

    q, mq = one(x[:, :q_hidden], q_weight, q_bias)
    k, mk = one(x[:, q_hidden:q_hidden + kv_hidden], k_weight, k_bias)
    v = qkv[:, q_hidden + kv_hidden:].clone()
    return (q, k, v, mq, mk) if with_scale else (q, k, v)


def split_qkvgate_gemma_rmsnorm_rope(qkvgate, sin, cos, q_hidden, kv_hidden, head_dim, rope_dim, eps, q_weight, k_weight, with_scale=False):
    """The gated variant (reference split_qkv_rmsnorm_rope.py:441-747, golden test_split_qkv_rmsnorm_rope.py:221-266): every q head
    is followed by its gate head; q, k normalised with (1 + weight); -> (q, k, v, gate[, mq, mk])."""
    b = qkvgate.shape[0]
    qg = qkvgate[:, :2 * q_hidden].reshape(b, q_hidden // head_dim, 2, head_dim)
    q_in = qg[:, :, 0].reshape(b, q_hidden)
    gate = qg[:, :, 1].reshape(b, q_hidden).clone()
    plain = torch.cat((q_in, qkvgate[:, 2 * q_hidden:]), dim=1).contiguous()
    sin, cos = sin.reshape(b, rope_dim), cos.reshape(b, rope_dim)
    out = split_qkv_rmsnorm_rope(plain, sin, cos, q_hidden, kv_hidden, head_dim, eps, q_weight, k_weight, None, None, True, with_scale, gemma=True)
    return out[:3] + (gate,) + out[3:]


def split_qkv_rmsnorm_rope_pos_cache_half(qkv, positions, cos_sin_cache, q_hidden, kv_hidden, head_dim, eps=None, q_weight=None, k_weight=None,
                                          q_bias=None, k_bias=None, rope_dim=None, cast_norm=True, with_scale=False):
    """The position-cache variant (reference split_qkv_rmsnorm_rope_pos_cache_half_npu.py:25-230): cos / sin are row
    clamp(positions[t], 0, max_seq - 1) of the cache [max_seq, rope_dim] = [cos(rope_dim/2) | sin(rope_dim/2)], one value per neox pair;
    with cast_norm the normalised value is rounded to the tensor dtype before the rotation; fp32 arithmetic, one rounding at the store."""
    rd = head_dim if rope_dim is None else rope_dim
    p = positions.to(torch.int64).clamp(0, cos_sin_cache.shape[0] - 1)
    rows = cos_sin_cache.to(torch.float32)[p][:, :rd]
    cos, sin = rows[:, :rd // 2], rows[:, rd // 2:]
    # the same restatement as above with cos / sin repeated over both halves; they stay float32 (a float32 cache is not rounded)
    return split_qkv_rmsnorm_rope(qkv, _F32(torch.cat((sin, sin), 1)), _F32(torch.cat((cos, cos), 1)), q_hidden, kv_hidden, head_dim, eps,
                                  q_weight, k_weight, q_bias, k_bias, True, with_scale, cast_norm=cast_norm and eps is not None)


class _F32:
    """wraps a float32 tensor so that the restatement's `.to(float32).numpy()` and `.shape` see it unchanged"""
    def __init__(self, t):
        self.t, self.shape = t, t.shape

    def to(self, _):
        return self.t


def mrope_select(cos_sin, mrope_section, is_interleaved):
    """[3, tokens, rope_dim] -> [tokens, rope_dim]: pair i takes its cos / sin from the t, h or w row by section
    (the reference test's _select_mrope_cos_sin, test_split_qkv_rmsnorm_mrope.py:8-50)"""
    rd = cos_sin.shape[-1]
    i = np.arange(rd // 2)
    if is_interleaved:
        sec = np.where((i % 3 == 1) & (i <= 3 * mrope_section[1]), 1, np.where((i % 3 == 2) & (i <= 3 * mrope_section[2]), 2, 0))
    else:
        sec = np.where(i < mrope_section[0], 0, np.where(i < mrope_section[0] + mrope_section[1], 1, 2))
    sec = torch.from_numpy(np.concatenate((sec, sec)))
    return cos_sin.permute(1, 2, 0)[:, torch.arange(rd), sec]


def split_qkv_rmsnorm_mrope(qkv, q_weight, k_weight, cos_sin, num_q_heads, num_kv_heads, head_size, eps, mrope_section, is_interleaved,
                            rope_dim=None, q_bias=None, k_bias=None, has_gate=False, with_scale=False):
    """reference split_qkv_rmsnorm_mrope.py:57-333 / golden test_split_qkv_rmsnorm_mrope.py:74-110 -> (q, k, v, gate[, mq, mk])"""
    b = qkv.shape[0]
    qh, kvh = num_q_heads * head_size, num_kv_heads * head_size
    rd = head_size if rope_dim is None else rope_dim
    sel = mrope_select(cos_sin, mrope_section, is_interleaved).to(torch.float32)
    cos, sin = sel[:, :rd // 2], sel[:, rd // 2:]
    if has_gate:
        qg = qkv[:, :2 * qh].reshape(b, num_q_heads, 2, head_size)
        gate = qg[:, :, 1].reshape(b, qh).clone()
        plain = torch.cat((qg[:, :, 0].reshape(b, qh), qkv[:, 2 * qh:]), dim=1).contiguous()
    else:
        gate, plain = qkv.new_empty((b, 0)), qkv
    out = split_qkv_rmsnorm_rope(plain, _F32(torch.cat((sin, sin), 1)), _F32(torch.cat((cos, cos), 1)), qh, kvh, head_size, eps, q_weight, k_weight,
                                 q_bias, k_bias, True, with_scale)
    return out[:3] + (gate,) + out[3:]


def split_qkv_tp_rmsnorm_rope(inputs, cos, sin, q_hidden, kv_hidden, head_dim, eps, q_weights, k_weights, rotary_dim):
    """The tensor-parallel variant (reference split_qkv_tp_rmsnorm_rope.py:8-170; golden helpers rms_norm_tp / custom_rope,
    test_split_qkv_tp_rmsnorm_rope.py:7-45).  `inputs`, `q_weights`, `k_weights`: one tensor per rank (a single rank: lists of one).
    var = sum over ranks of mean_local(x^2) / tp_world; y = act(x / sqrt(var + eps) * w[col]) rounded to the tensor dtype, then the
    first rotary_dim dims of every head rotated with cos[t, :half], sin[t, :half], rounded again.  -> [(q, k, v)] per rank."""
    tp = len(inputs)
    dt = inputs[0].dtype
    f = lambda t: t.to(torch.float32).numpy().astype(np.float32)
    b = inputs[0].shape[0]
    half = rotary_dim // 2
    c, s = f(cos).reshape(b, -1)[:, None, :half], f(sin).reshape(b, -1)[:, None, :half]
    var = [sum(np.mean(f(x)[:, lo:hi] ** 2, axis=-1, keepdims=True, dtype=np.float32) for x in inputs) * np.float32(1.0 / tp)
           for lo, hi in ((0, q_hidden), (q_hidden, q_hidden + kv_hidden))]
    rnd = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(dt)
    outs = []
    for x, qw, kw in zip(inputs, q_weights, k_weights):
        res = []
        for (lo, hi), v_, w in (((0, q_hidden), var[0], qw), ((q_hidden, q_hidden + kv_hidden), var[1], kw)):
            y = f(x)[:, lo:hi] * (np.float32(1) / np.sqrt(v_ + np.float32(eps), dtype=np.float32)) * f(w)
            y = rnd(y).to(torch.float32).numpy().reshape(b, -1, head_dim)
            x1, x2 = y[..., :half], y[..., half:2 * half]
            y = np.concatenate((x1 * c - x2 * s, x2 * c + x1 * s, y[..., 2 * half:]), axis=-1)
            res.append(rnd(y.reshape(b, -1)))
        outs.append((res[0], res[1], x[:, q_hidden + kv_hidden:].clone()))
    return outs
