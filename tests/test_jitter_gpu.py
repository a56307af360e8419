"""Race hunting without a sanitizer (compute-sanitizer is closed on this pool: profiles/r02/sanitizer_closed.txt).

Debug bit21 — in the instrumented build of the library (-DMLAPO_INSTRUMENTED), loaded by a worker process through MLAPO_LIB; the
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
JITTER_LIB = os.path.join(os.path.dirname(HERE), "sgl-kernel-npu_b200", "sgl_kernel_npu", "lib", "libmla_preprocess_b200_instr.so")


def _run_worker(which):
    assert os.path.exists(JITTER_LIB), f"{JITTER_LIB} missing: __graft_entry__.build() makes it"
    env = dict(os.environ, MLAPO_LIB=JITTER_LIB)
    r = subprocess.run([sys.executable, os.path.join(HERE, "jitter_worker.py"), which], env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, f"jitter worker failed:\n{r.stdout[-2000:]}\n{r.stderr[-4000:]}"


def test_pair_kernel_late_drain_group(cuda_device):
    """Regression (160 / 192 units on 74 pairs): a pair streams several (head, 256-token) units, and the absorb of unit i+1 needs
    nothing from the drains of unit i.  With one `unit_full` barrier per TMEM buffer (shared by both TMEM regions) a drain group
    that reached its wait for chunk 2 / 3 of unit i late found the barrier two completions on and waited for ever (seen once in
    ~20 000 launches at C3, after a faster dequant made unit i+1's first chunks arrive sooner).  Debug bit20 makes every second
    CTA's drain groups 6000 cycles late before those chunks: with the barriers per (region, buffer) the call completes with the
    undisturbed result; with the shared barriers it ran into the bounded wait every time (a trap after ~2 s)."""
    _run_worker("late_drain")


def test_other_operators_under_jitter(cuda_device):
    """The parity tests of `batch_matmul_transpose` and of the `split_qkv_rmsnorm_rope` family (their kernels use the same bounded
    barrier waits), run as they are in a child pytest with the instrumented library and the jitter on from the first call."""
    assert os.path.exists(JITTER_LIB), f"{JITTER_LIB} missing: __graft_entry__.build() makes it"
    env = dict(os.environ, MLAPO_LIB=JITTER_LIB, MLAPO_DEBUG_FLAGS="2097152")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(HERE, "test_batch_matmul_transpose.py"), os.path.join(HERE, "test_split_qkv.py"),
                        "-m", "gpu", "-q", "-x", "-p", "no:cacheprovider", "-k", "cuda"], env=env, capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0 and " passed" in r.stdout, f"child pytest under jitter failed:\n{r.stdout[-3000:]}\n{r.stderr[-2000:]}"


def test_results_do_not_depend_on_timing(cuda_device):
    """Every shape of jitter_worker.SHAPES (one undisturbed run, four runs under jitter: all outputs / caches / taps equal) and
    eight chained calls with the no-wait start on one workspace, in one worker process (one torch import)."""
    _run_worker("all")
