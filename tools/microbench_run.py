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

# MMA shapes of the op on CTA pairs: fraction of the nominal dense peak and SM cycles per instruction (separate line, optional)
if hasattr(lib, "mlapo_mb_pair_shape"):
    lib.mlapo_mb_pair_shape.restype = ctypes.c_double
    lib.mlapo_mb_pair_shape.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
    shapes = {}
    for kind, kname in ((0, "i8"), (1, "bf16")):
        for n in (64, 128, 192, 256):
            for chains in (1, 2):
                cyc = ctypes.c_double(0.0)
                frac = lib.mlapo_mb_pair_shape(kind, n, chains, ctypes.byref(cyc))
                shapes[f"{kname} M256 N{n} K{32 if kind == 0 else 16} chains={chains}"] = {"frac_of_nominal_peak": round(frac, 3),
                                                                                          "cycles_per_mma": round(cyc.value, 1)}
    print(json.dumps({"pair_mma_shapes": shapes}))
if hasattr(lib, "mlapo_mb_single_shape"):
    lib.mlapo_mb_single_shape.restype = ctypes.c_double
    lib.mlapo_mb_single_shape.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
    shapes = {}
    for kind, kname in ((0, "i8"), (1, "bf16")):
        for n in (64, 128, 192, 256):
            cyc = ctypes.c_double(0.0)
            frac = lib.mlapo_mb_single_shape(kind, n, 2, ctypes.byref(cyc))
            shapes[f"{kname} M128 N{n} K{32 if kind == 0 else 16}"] = {"frac_of_nominal_peak": round(frac, 3), "cycles_per_mma": round(cyc.value, 1)}
    if hasattr(lib, "mlapo_mb_single_shape_m"):
        lib.mlapo_mb_single_shape_m.restype = ctypes.c_double
        lib.mlapo_mb_single_shape_m.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_double)]
        for kind, kname in ((0, "i8"), (1, "bf16")):
            for n in (128, 256):
                cyc = ctypes.c_double(0.0)
                frac = lib.mlapo_mb_single_shape_m(kind, 64, n, 2, ctypes.byref(cyc))
                shapes[f"{kname} M64 N{n} K{32 if kind == 0 else 16}"] = {"frac_of_nominal_peak": round(frac, 3), "cycles_per_mma": round(cyc.value, 1)}
        for n in (128, 256):   # two issuing threads, each with its own accumulator (chains = 3 selects that form)
            cyc = ctypes.c_double(0.0)
            frac = lib.mlapo_mb_single_shape_m(0, 128, n, 3, ctypes.byref(cyc))
            shapes[f"i8 M128 N{n} K32, two issuing warps"] = {"frac_of_nominal_peak": round(frac, 3), "cycles_per_mma": round(cyc.value, 1)}
        for n in (128, 192, 256):   # the whole warp runs the issue loop, the asm block elects the lane (chains = 4 selects that form)
            cyc = ctypes.c_double(0.0)
            frac = lib.mlapo_mb_single_shape_m(0, 128, n, 4, ctypes.byref(cyc))
            shapes[f"i8 M128 N{n} K32, convergent issue loop with elect inside"] = {"frac_of_nominal_peak": round(frac, 3), "cycles_per_mma": round(cyc.value, 1)}
    print(json.dumps({"single_cta_mma_shapes": shapes}))
