"""The quantise kernel divides by the tensor scale with q1 = fma(fma(-q0, s, x), r, q0), r = RN(1/s), q0 = RN(x r) instead of the IEEE
division sequence (csrc/mlapo_quant_norm.cuh, QuantScale / quant_vectors).  That is only legitimate if q1 == RN(x / s) whenever s
carries <= 11 significant bits (a bf16 / fp16 value).  Checked here in exact rational arithmetic for ARBITRARY fp32 numerators
(a superset of the kernel's 11-bit ones), random and adversarial; tools/check_short_division.py is the same check at 60 000
samples.  (The norm kernels' second quantiser could use the same identity — its numerator is a full fp32 value — but there it
measured slower than the compiler's division by a loop-invariant divisor: 29.95 vs 28.65 us at C3, so they keep `/`.)"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def test_short_scale_division_is_correctly_rounded():
    env = dict(os.environ, N="6000")
    r = subprocess.run([sys.executable, os.path.join(os.path.dirname(HERE), "tools", "check_short_division.py")], env=env, capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0 and "mismatches 0" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
