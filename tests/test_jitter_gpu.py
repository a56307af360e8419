"""Race hunting without a sanitizer (compute-sanitizer is closed on this pool: profiles/r02/sanitizer_closed.txt).

Debug bit21 — in a separate build of the library with -DMLAPO_WAIT_JITTER, loaded by a worker process through MLAPO_LIB; the
shipped library carries no code for it — puts a random idle period of up to 3000 cycles in front of every fourth mbarrier wait of every kernel (chosen by a
hash of the clock, so every launch perturbs a different set of warps).  The kernels' results do not depend on timing — integer
sums are associative, every floating-point value is computed by one thread in a fixed order — so a run under jitter must equal
the undisturbed run of the same inputs BIT FOR BIT, outputs, caches and workspace taps alike; a protocol that only holds at the
usual timing (a barrier phase consumed late, a buffer released early) shows up as a mismatch or as a stuck wait (the bounded
waits turn that into an error naming the kernel).  The shapes walk every kernel and launch form of the path: single-CTA and
pair MM1 with 128- and 256-column tiles, the dequantising MM1, q_fused with 1 / 2 / 4 CTAs per unit, the CTA-pair q kernel with
several units per pair, all cache modes, both activation types, and chained calls with the no-wait start.
"""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
JITTER_LIB = os.path.join(os.path.dirname(HERE), "sgl-kernel-npu_b200", "sgl_kernel_npu", "lib", "libmla_preprocess_b200_jitter.so")


def _run_worker(which):
    assert os.path.exists(JITTER_LIB), f"{JITTER_LIB} missing: __graft_entry__.build() makes it"
    env = dict(os.environ, MLAPO_LIB=JITTER_LIB)
    r = subprocess.run([sys.executable, os.path.join(HERE, "jitter_worker.py"), which], env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, f"jitter worker failed:\n{r.stdout[-2000:]}\n{r.stderr[-4000:]}"


def test_results_do_not_depend_on_timing(cuda_device):
    """Every shape of jitter_worker.SHAPES (one undisturbed run, four runs under jitter: all outputs / caches / taps equal) and
    eight chained calls with the no-wait start on one workspace, in one worker process (one torch import)."""
    _run_worker("all")
