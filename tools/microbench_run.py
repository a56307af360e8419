"""Runs the measurement helpers of tools/microbench/libmlapo_microbench.so on cuda:0 and prints one JSON object."""
import ctypes
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = ctypes.CDLL(os.path.join(ROOT, "tools", "microbench", "libmlapo_microbench.so"))
lib.mlapo_mb_i8_peak.restype = ctypes.c_double
lib.mlapo_mb_i8_peak.argtypes = [ctypes.c_double, ctypes.POINTER(ctypes.c_double)]
lib.mlapo_mb_dsmem.restype = ctypes.c_double
lib.mlapo_mb_dsmem.argtypes = [ctypes.c_int, ctypes.c_int]

out = {"int8_peak_tops": {}, "dsmem_bytes_per_clk_per_cta": {}}
for label, ms in (("5ms", 5.0), ("30ms", 30.0), ("1s", 1000.0)):
    took = ctypes.c_double(0.0)
    tops = lib.mlapo_mb_i8_peak(ms, ctypes.byref(took))
    out["int8_peak_tops"][label] = {"tops": round(tops, 1), "ms": round(took.value, 2)}
for mode, name in ((0, "st.shared::cluster.v4"), (1, "cp.async.bulk smem->peer smem"), (2, "cp.reduce.async.bulk add.s32 -> peer smem")):
    out["dsmem_bytes_per_clk_per_cta"][name] = round(lib.mlapo_mb_dsmem(mode, 64), 2)
print(json.dumps(out))
