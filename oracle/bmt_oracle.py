"""CPU oracle for batch_matmul_transpose.  TEST INFRASTRUCTURE ONLY — imported by tests/ and nothing else.

Restates the operator contract of the reference (csrc/batch_matmul_transpose/README.md:1-16, host
op_host/batch_matmul_transpose.cpp:46-114): tensor_c[m, b, :] = tensor_a[m, b, :] @ tensor_b[b] for every batch b, inputs and
output fp16 or bf16.  The reference test's golden is ``torch.bmm(a.transpose(0, 1), b, out=c.view(-1, m, n).transpose(0, 1))``
(tests/python/sgl_kernel_npu/test_batch_matmul_transpose.py:18-20); here the sum is carried in float64 and rounded ONCE to the
output type, i.e. the correctly rounded result that every fp32-accumulating implementation is within one output ulp of.
Pinned by tests/test_batch_matmul_transpose.py against fixtures produced by that torch.bmm expression on this container's CPU
(oracle/make_bmt_golden.py; torch_npu itself is not available here).
"""
import torch


def nz_pack_b(b_nd: torch.Tensor) -> torch.Tensor:
    """ND [batch, k, n] -> NZ [batch, n/16, k, 16] (README.md:8); k and n must be multiples of 16."""
    batch, k, n = b_nd.shape
    assert k % 16 == 0 and n % 16 == 0
    return b_nd.reshape(batch, k, n // 16, 16).permute(0, 2, 1, 3).contiguous()


def nz_unpack_b(b_nz: torch.Tensor) -> torch.Tensor:
    batch, nb, k, c0 = b_nz.shape
    return b_nz.permute(0, 2, 1, 3).reshape(batch, k, nb * c0).contiguous()


def batch_matmul_transpose(tensor_a: torch.Tensor, tensor_b: torch.Tensor, format_mode: str = "ND") -> torch.Tensor:
    """a [m, batch, k], b ND [batch, k, n] or NZ [batch, n/16, k, 16] -> c [m, batch, n] in a's dtype."""
    if format_mode not in ("ND", "NZ"):
        raise RuntimeError(f"format_mode: Unsupported mode value {format_mode}")
    b = nz_unpack_b(tensor_b) if format_mode == "NZ" else tensor_b
    c = torch.einsum("mbk,bkn->mbn", tensor_a.double(), b.double())
    return c.to(tensor_a.dtype)
