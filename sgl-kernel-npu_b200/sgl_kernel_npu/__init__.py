"""B200 drop-in for the `mla_preprocess` path of sgl-kernel-npu.

Mirrors the reference loader (python/sgl_kernel_npu/sgl_kernel_npu/__init__.py:9-15): importing the
package loads the operator library so that `torch.ops.npu.mla_preprocess` exists.  The operator
library is optional at import time only in the sense that the C-ABI binding (`sgl_kernel_npu.cabi`)
works without it; neither has a CPU fallback.
"""
import os
import pathlib

import torch

_LIB_DIR = pathlib.Path(__file__).parent / "lib"
OP_LIBRARY = str(_LIB_DIR / "libsgl_kernel_npu.so")


def _load_sgl_kernel_npu():
    if os.path.exists(OP_LIBRARY):
        torch.ops.load_library(OP_LIBRARY)
        return True
    return False


OP_LIBRARY_LOADED = _load_sgl_kernel_npu()
