#!/usr/bin/env python
"""bench.py — mla_preprocess throughput on B200 (BASELINE.json metric: MLA-preprocess tokens/s/GPU at DSV3 shapes;
% of HBM / INT8-TC roofline).

One "step" = one pass of the hot path (one `mla_preprocess` call) over one batch of synthetic DeepSeek-V3-shaped input.
Default workload = BASELINE.json configs[1] ("C2": decode, 256 tokens, 128 heads, D=7168, bf16 activations, W8A8, paged KV
block 128, cache_mode krope_ctkv, per-tensor quant).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c1|c2|c3|c4|c5] [--impl ours|reference]

The ONE JSON line carries, for the default workload:
* value / ms_per_step   tokens/s of the whole job with inputs resident in HBM, CUDA events on the launch stream.  The K-step block
             is repeated until at least 50 ms have been timed (a single block of K decode steps lasts well under a millisecond);
             ms_per_step is the MEDIAN block (ms_per_step_min / _max beside it), max over ranks per block.  Between steps the
             weight set rotates over R sets whose total size exceeds L2, so weights stream from HBM as they do when 61 layers
             cycle in serving (`config.l2`).
* e2e        the same metric through the reference-facing operator, `torch.ops.npu.mla_preprocess`, with FRACTAL_NZ weights (the
             drop-in call of the reference's test, test_mla_preprocess.py:547-578) and HOST buffers: per step H2D of
             hidden/cos/sin/slot_mapping from pinned memory and D2H of q_out0/q_out1, with the achieved PCIe GB/s per rank;
             `e2e.ctypes` is the same through the ctypes mirror of the C ABI.
* roofline   dominant kernel: algorithmic bytes per launch / CUDA-event duration of that kernel (serialised pass), against
             MEASURED_PEAKS.json, with the kernel's in-chain span from the per-CTA timeline marks beside it.
* prefill    sub-record: BASELINE configs[2] (C3, 8192 tokens x 128 heads) — tokens/s and the int8 tensor roofline against the
             nominal 4.5 POPS AND the peak measured on this GPU by tools/microbench (tcgen05.mma.kind::i8, resident tiles).
* c5         sub-record: BASELINE configs[4] (Kimi-K2 rank: 128 tokens x 64 heads, ragged slots) in every cache mode.
* tp         sub-record at WORLD_SIZE > 1: BASELINE configs[3] (C4: batch 512, the 128 heads split over the ranks).
* cpu_baseline  the oracle port (oracle/mla_oracle.py, torch CPU) timed on the host cores on the full batch (rank 0, N=1).
--impl reference times the reference's CPU implementation of the path.  The reference's device code is AscendC (cannot build
here) and its CPU golden lives in a test file that does not travel to the GPU box, so this arm runs the oracle port of that
golden (kind "port") with all host threads.
Multi-GPU: the path has no collective; ranks are independent DP-attention replicas (weak scaling) or head shards (TP, strong
scaling); barrier only to align the timing window, time = max over ranks.
"""
import argparse
import ctypes
import json
import math
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "sgl-kernel-npu_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

WORKLOADS = {
    # name: (tokens, heads, hidden, cache_mode, description)
    "c1": (8, 16, 7168, "krope_ctkv", "DeepSeek-V3 MLA preprocess decode, batch 8, 16 heads (TP8 slice), W8A8, paged block 128"),
    "c2": (256, 128, 7168, "krope_ctkv", "DeepSeek-V3 MLA preprocess decode, batch 256, 128 heads, single B200, paged KV block 128"),
    "c3": (8192, 128, 7168, "krope_ctkv", "DeepSeek-V3 MLA preprocess prefill, 8192 tokens, 128 heads, single B200"),
    "c4": (512, 128, 7168, "krope_ctkv", "DeepSeek-V3 head-sharded TP=n_gpus decode batch 512, 128 heads split over ranks, W_UQ/W_UK column-sharded, no collective"),
    "c5": (128, 64, 7168, "int8_nzcache", "Kimi-K2 (64 heads) DP-attention rank, batch 128, ragged slot_mapping, int8_nzcache"),
}
METRIC = "MLA-preprocess tokens/s/GPU (DSV3 shapes); % of HBM / INT8-TC roofline"
TP_WORKLOADS = {"c4"}   # heads sharded over ranks (strong scaling); every other workload runs one replica per rank (weak)
QUANT_MODE = "per_tensor_quant_asymm"
MIN_TIMED_MS = 50.0     # the K-step block is repeated until this much has been timed
INT8_NOMINAL_TOPS = 4500.0
L2_MB = 126


# ----------------------------------------------------------------------------------------------- multi-rank plumbing
def rank_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def shard_heads(h_total, world, rank):
    """TP: rank r owns heads [r*H/tp, (r+1)*H/tp); H must divide evenly."""
    if h_total % world:
        raise ValueError(f"{h_total} heads do not split over {world} ranks")
    per = h_total // world
    return rank * per, (rank + 1) * per


def job_tokens(n_tokens, world, sharding):
    """Tokens the whole job processes per step: replicas each take their own batch; TP ranks share one batch."""
    return n_tokens if sharding == "tp" else n_tokens * world


class Ranks:
    """Barrier + max-over-ranks; the data path itself has no collective."""

    def __init__(self, backend="nccl", device=None):
        self.rank, self.world, self.local = rank_env()
        self.dist, self.device = None, device
        if self.world > 1:
            import torch.distributed as dist
            kw = {"device_id": device} if (backend == "nccl" and device is not None) else {}
            dist.init_process_group(backend, **kw)
            self.dist = dist

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()

    def max(self, values):
        if self.dist is None:
            return [float(v) for v in values]
        import torch
        t = torch.tensor(list(values), dtype=torch.float64, device=self.device if self.device is not None else "cpu")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(v) for v in t]

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


def algorithmic_bytes(n, h, d, cache_mode):
    """compulsory bytes per call (weights once + per-token inputs/outputs/cache rows)."""
    w = 2112 * d + 1536 * 192 * h + 2 * 128 * 512 * h + (8 * 2112 + 8 * 192 * h + 4 * 1536 + 2 * 512)
    if cache_mode == "int8_nzcache":
        a = 2 * d + 256 + 4 + 512 * h + 128 * h + 640
    else:
        a = 2 * d + 256 + 4 + 1024 * h + 128 * h + 1152
    return w + n * a


def algorithmic_ops(n, h, d):
    """(int8 ops, bf16 flops) per call."""
    return 2.0 * n * (2112 * d + 1536 * 192 * h), 2.0 * n * 128 * 512 * h


def kernel_algorithmic_bytes(name, n, h, d, cache_mode):
    """Per-kernel compulsory bytes (each kernel's own inputs + outputs), stated in DESIGN.md §4."""
    q0 = 1 if cache_mode == "int8_nzcache" else 2
    qb = 1536 * 192 * h + 2 * 128 * 512 * h + n * 1536 + n * h * 64 * 2 + n * h * 512 * q0 + 8 * 192 * h + n * 256
    return {
        "quant_hidden": n * d * 2 + n * d,
        "mm1_splitk": 2112 * d + n * d + n * 2112 * 4,
        "mm1_pair": 2112 * d + n * d + n * 2112 * 4,
        "mm1_dequant": 2112 * d + n * d + n * 2112 * 2,
        "norm_rope_cache": n * 2112 * 4 + n * 1536 + n * (512 * q0 + 128) + n * 260 + 8 * 2112 + 4 * 1536 + 1024,
        "norm_rope_cache_y0": n * 2112 * 2 + n * 1536 + n * (512 * q0 + 128) + n * 260 + 4 * 1536 + 1024,
        "q_fused": qb,
        "q_pair": qb,
    }.get(name, 0)


def _peaks_file():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def tensor_peak_bf16():
    try:
        return float(_peaks_file()["bf16_tflops"])
    except Exception:
        return 1590.0  # B200_PROFILING.md fallback


def load_peaks():
    try:
        return float(_peaks_file()["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def measured_int8_peak(target_ms=30.0):
    """tcgen05.mma.kind::i8 peak of THIS GPU from tools/microbench (resident tiles, every SM); None if the helper is not built."""
    path = os.path.join(ROOT, "tools", "microbench", "libmlapo_microbench.so")
    if not os.path.exists(path):
        return None
    try:
        lib = ctypes.CDLL(path)
        lib.mlapo_mb_i8_peak.restype = ctypes.c_double
        lib.mlapo_mb_i8_peak.argtypes = [ctypes.c_double, ctypes.POINTER(ctypes.c_double)]
        out = {}
        for label, ms in (("burst", target_ms), ("sustained", 1000.0)):
            took = ctypes.c_double(0.0)
            tops = lib.mlapo_mb_i8_peak(ms, ctypes.byref(took))
            if tops <= 0:
                return None
            out[label] = {"tops": round(tops, 1), "ms": round(took.value, 2)}
        return out
    except Exception as e:  # noqa
        print(f"[bench] int8 peak microbench unavailable: {type(e).__name__}: {e}", file=sys.stderr)
        return None


class ClockSampler:
    """Samples SM clock + throttle reasons during the timed regions (pynvml).  NVML is initialised and the first sample is
    taken by `start()` itself, on the calling thread, so that even a short region has at least one sample."""

    def __init__(self, index):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._thr = None
        self._nv = self._hdl = self._get_reasons = None
        self._names = {}

    def _init_nvml(self):
        import pynvml as nv
        nv.nvmlInit()
        self._nv = nv
        self._hdl = nv.nvmlDeviceGetHandleByIndex(self.index)
        self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(self._hdl, nv.NVML_CLOCK_SM)
        self._names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake_slowdown",
        }
        self._get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons

    def _sample(self):
        self.samples.append(self._nv.nvmlDeviceGetClockInfo(self._hdl, self._nv.NVML_CLOCK_SM))
        mask = self._get_reasons(self._hdl)
        for bit, name in self._names.items():
            if bit and (mask & bit):
                self.reasons.add(name)

    def _run(self):
        try:
            while not self._stop.is_set():
                self._sample()
                time.sleep(0.002)
        except Exception as e:  # noqa
            self.reasons.add(f"clock_sampling_failed:{type(e).__name__}")

    def start(self):
        try:
            if self._nv is None:
                self._init_nvml()
        except Exception as e:  # noqa
            self.reasons.add(f"clock_sampling_unavailable:{type(e).__name__}")
            return
        self._stop.clear()
        self._thr = threading.Thread(target=self._run, daemon=True)
        self._thr.start()

    def pause(self):
        """Stop sampling (between timed regions); `start()` resumes into the same record."""
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=2)
            self._thr = None

    def sample_now(self):
        try:
            if self._nv is not None:
                self._sample()
        except Exception:
            pass

    def summary(self, since=0):
        s = sorted(self.samples[since:])
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


# ----------------------------------------------------------------------------------------------- synthetic inputs
def make_weight_set(torch, h, d, dev, gen, nz=False):
    """One layer's weights in the layouts the C ABI consumes (K-major int8; reference value distribution {-1,0,1}).
    nz=True adds the FRACTAL_NZ copies the operator schema takes ([1, K/32, n_out, 32], test_mla:537-544)."""
    ri = lambda lo, hi, shape, dt: torch.randint(lo, hi, shape, device=dev, generator=gen, dtype=torch.int32).to(dt)
    bf = torch.bfloat16
    w = dict(
        wdqkv_km=ri(-1, 2, (2112, d), torch.int8), wuq_km=ri(-1, 2, (192 * h, 1536), torch.int8),
        wuk=(torch.rand((h, 128, 512), device=dev, generator=gen) * 4 - 2).to(bf),
        descale0=torch.rand(2112, device=dev, generator=gen) / 1000, descale1=torch.rand(192 * h, device=dev, generator=gen) / 1000,
        bias0=ri(-10, 10, (2112,), torch.int32), bias1=ri(-10, 10, (192 * h,), torch.int32),
        gamma1=(torch.rand(1536, device=dev, generator=gen) * 2 - 1).to(bf), beta1=ri(-2, 2, (1536,), bf),
        gamma2=(torch.rand(512, device=dev, generator=gen) * 2 - 1).to(bf),
        quant_scale0=torch.tensor([0.0371], device=dev).to(bf), quant_offset0=torch.tensor([3], device=dev, dtype=torch.int8),
        quant_scale1=torch.tensor([0.0213], device=dev).to(bf), quant_offset1=torch.tensor([-5], device=dev, dtype=torch.int8),
        ctkv_scale=torch.tensor([0.043], device=dev).to(bf),
        q_nope_scale=(torch.rand(h, device=dev, generator=gen) * 0.1 + 0.01).to(bf),
    )
    if nz:
        to_nz = lambda km: km.reshape(km.shape[0], km.shape[1] // 32, 32).permute(1, 0, 2).reshape(1, km.shape[1] // 32, km.shape[0], 32).contiguous()
        w["wdqkv_nz"], w["wuq_nz"] = to_nz(w["wdqkv_km"]), to_nz(w["wuq_km"])
        w["gamma0"], w["beta0"] = torch.ones(d, dtype=bf, device=dev), torch.zeros(d, dtype=bf, device=dev)
    return w


def make_slots(torch, n, num_slots, g, ragged):
    """Unique random slots with 1/16 of the rows at -1 (padding rows of a ragged batch); `ragged` (C5): per-request contiguous
    runs at random block offsets with 10 % of the rows at -1."""
    if not ragged:
        slots = torch.randperm(num_slots, generator=g)[:n].to(torch.int32)
        if n >= 16:
            slots[torch.randperm(n, generator=g)[: n // 16]] = -1
        return slots
    slots = torch.empty(n, dtype=torch.int32)
    pos, req = 0, 0
    blocks = torch.randperm(num_slots // 256, generator=g)
    while pos < n:
        ln = int(min(n - pos, int(torch.randint(1, 33, (1,), generator=g))))
        start = int(blocks[req]) * 256 + int(torch.randint(0, 128, (1,), generator=g))
        slots[pos:pos + ln] = torch.arange(start, start + ln, dtype=torch.int32)
        pos, req = pos + ln, req + 1
    slots[torch.randperm(n, generator=g)[: max(1, n // 10)]] = -1
    return slots


def make_host_batch(torch, n, d, num_slots, seed, ragged=False):
    g = torch.Generator().manual_seed(seed)
    bf = torch.bfloat16
    hidden = (torch.rand((n, d), generator=g) * 4 - 2).to(bf).pin_memory()
    cos = (torch.rand((n, 64), generator=g) * 2 - 1).to(bf).pin_memory()
    sin = (torch.rand((n, 64), generator=g) * 2 - 1).to(bf).pin_memory()
    return hidden, cos, sin, make_slots(torch, n, num_slots, g, ragged).pin_memory()


def cache_tensors(torch, cache_mode, num_blocks, dev):
    bf = torch.bfloat16
    if cache_mode == "krope_ctkv":
        return torch.zeros((num_blocks, 128, 1, 512), dtype=bf, device=dev), torch.zeros((num_blocks, 128, 1, 64), dtype=bf, device=dev)
    if cache_mode == "int8_nzcache":
        return torch.zeros((num_blocks, 16, 128, 32), dtype=torch.int8, device=dev), torch.zeros((num_blocks, 4, 128, 16), dtype=bf, device=dev)
    return torch.zeros((num_blocks, 32, 128, 16), dtype=bf, device=dev), torch.zeros((num_blocks, 4, 128, 16), dtype=bf, device=dev)


def weight_set_count(h, d):
    set_bytes = 2112 * d + 1536 * 192 * h + 2 * 128 * 512 * h
    return max(2, min(8, -(-300_000_000 // set_bytes))), set_bytes   # total weights > 2 x L2


def workload_config(wid, world, cache_mode=None):
    """The `config` object of the JSON line; identical in both arms (the reference arm describes the same workload)."""
    n, h, d, cm, desc = WORKLOADS[wid]
    cm = cache_mode or cm
    sharding = "tp" if wid in TP_WORKLOADS else "replica"
    hp = h // world if sharding == "tp" else h
    n_sets, set_bytes = weight_set_count(hp, d)
    return {"workload": desc, "id": wid, "tokens_per_gpu": n, "heads_per_gpu": hp, "hidden": d, "cache_mode": cm, "quant_mode": QUANT_MODE,
            "sharding": ("single GPU" if world == 1 else "heads split over ranks (TP), no collective" if sharding == "tp"
                         else "replicas (DP-attention), no collective"),
            "l2": f"weights rotate over {n_sets} sets = {n_sets * set_bytes / 1e6:.0f} MB > {L2_MB} MB L2; no explicit flush"}


# ----------------------------------------------------------------------------------------------- one workload on this rank
class Workload:
    """Device-resident inputs, parameter blocks and (optionally) a CUDA graph of one weight rotation for one workload."""

    def __init__(self, torch, cabi, dev, rank, world, wid, cache_mode=None, flags=0, use_graph=True, nz=False, n_sets=None):
        self.torch, self.cabi, self.dev, self.wid = torch, cabi, dev, wid
        n, h, d, cm, self.desc = WORKLOADS[wid]
        self.cache_mode = cache_mode or cm
        self.sharding = "tp" if wid in TP_WORKLOADS else "replica"
        if self.sharding == "tp":
            h0, h1 = shard_heads(h, world, rank)
            h = h1 - h0           # synthetic weights: each rank simply draws its own head slice
        self.n, self.h, self.d = n, h, d
        bf = torch.bfloat16
        self.q0_dt = torch.int8 if self.cache_mode == "int8_nzcache" else bf
        self.num_blocks = max(16, (n * 4 + 127) // 128 + 8)
        if wid == "c5":
            self.num_blocks = max(self.num_blocks, 2 * n + 8)   # every request of the ragged batch gets two blocks of its own
        gen = torch.Generator(device=dev).manual_seed(1234 + rank)
        auto_sets, self.set_bytes = weight_set_count(h, d)
        self.n_sets = n_sets or auto_sets
        self.sets = [make_weight_set(torch, h, d, dev, gen, nz=nz) for _ in range(self.n_sets)]
        self.host = make_host_batch(torch, n, d, self.num_blocks * 128, 99 + rank, ragged=wid == "c5")
        self.hidden, self.cos, self.sin, self.slots = (t.to(dev) for t in self.host)
        self.kv, self.kvr = cache_tensors(torch, self.cache_mode, self.num_blocks, dev)
        self.outs = [(torch.zeros((n, h, 512), dtype=self.q0_dt, device=dev), torch.zeros((n, h, 64), dtype=bf, device=dev))
                     for _ in range(self.n_sets)]
        self.flags = flags
        self.plist = [self.params_for(i) for i in range(self.n_sets)]
        self.lay = cabi.workspace_layout(self.plist[0])
        self.ws = torch.empty(self.lay.total_bytes, dtype=torch.uint8, device=dev)
        self.launches_per_call = cabi.lib.mlapo_launches_per_call(self.plist[0])
        self.stream = torch.cuda.current_stream(dev)
        self.graph, self.launch_mode = None, "eager"
        for i in range(3):
            self.step(i)
        torch.cuda.synchronize(dev)
        if use_graph:
            self._capture()

    def params_for(self, i):
        w = self.sets[i]
        return self.cabi.build_params(self.hidden, w["wdqkv_km"], w["descale0"], w["gamma1"], w["beta1"], w["wuq_km"], w["descale1"],
                                      w["gamma2"], self.cos, self.sin, w["wuk"], self.kv, self.kvr, self.slots, w["quant_scale0"],
                                      w["quant_offset0"], w["bias0"], w["quant_scale1"], w["quant_offset1"], w["bias1"], w["ctkv_scale"],
                                      w["q_nope_scale"], self.cache_mode, QUANT_MODE, self.outs[i][0], self.outs[i][1], flags=self.flags)

    def step(self, i):
        self.cabi.launch(self.plist[i % self.n_sets], self.ws, self.stream.cuda_stream)

    def _capture(self):
        torch, dev = self.torch, self.dev
        try:
            gs = torch.cuda.Stream(dev)
            with torch.cuda.stream(gs):
                g_ = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g_, stream=gs):
                    for i in range(self.n_sets):
                        self.cabi.launch(self.plist[i], self.ws, gs.cuda_stream)
                g_.replay()
            torch.cuda.synchronize(dev)
            self.graph, self.launch_mode = g_, f"cuda_graph({self.n_sets} steps/replay)"
        except Exception as e:  # noqa
            print(f"[bench] graph capture unavailable ({type(e).__name__}: {e}); timing eager launches", file=sys.stderr)
            self.graph = None
            torch.cuda.synchronize(dev)

    def run_steps(self, k):
        if self.graph is not None:
            full, rem = divmod(k, self.n_sets)
            for _ in range(full):
                self.graph.replay()
            for i in range(rem):
                self.step(i)
        else:
            for i in range(k):
                self.step(i)

    def time_blocks(self, ranks, steps, warmup, sampler=None, min_ms=MIN_TIMED_MS, max_blocks=400):
        """Times blocks of `steps` steps until `min_ms` have been timed; returns the per-block ms (max over ranks)."""
        torch, dev = self.torch, self.dev
        self.run_steps(max(warmup, 3))
        torch.cuda.synchronize(dev)
        # calibration block (untimed for the record): how many blocks make min_ms; every rank must agree on the count
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record(self.stream)
        self.run_steps(steps)
        c1.record(self.stream)
        torch.cuda.synchronize(dev)
        est = ranks.max([c0.elapsed_time(c1)])[0]
        blocks = int(max(1, min(max_blocks, math.ceil(min_ms / max(est, 1e-3)))))
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(blocks + 1)]
        ranks.barrier()
        torch.cuda.synchronize(dev)
        if sampler is not None:
            sampler.start()
        evs[0].record(self.stream)
        for b in range(blocks):
            self.run_steps(steps)
            evs[b + 1].record(self.stream)
        torch.cuda.synchronize(dev)
        if sampler is not None:
            sampler.sample_now()
            sampler.pause()
        ranks.barrier()
        return ranks.max([evs[b].elapsed_time(evs[b + 1]) for b in range(blocks)])

    def kernel_times_us(self, dbg=0):
        """Per-kernel durations from a separate SERIALISED pass (CUDA events recorded by the library around each kernel; no
        programmatic overlap between the kernels, warm L2 for the intermediates)."""
        cabi = self.cabi
        cabi.set_debug_flags(4 | dbg)
        ksum, kcount = None, 0
        try:
            for i in range(40):
                self.step(i)
                buf = (ctypes.c_float * 8)()
                cnt = cabi.lib.mlapo_get_kernel_times_ms(buf, 8)
                if i >= 8 and cnt > 0:
                    vals = [buf[j] for j in range(cnt)]
                    ksum = vals if ksum is None else [a + b for a, b in zip(ksum, vals)]
                    kcount += 1
        finally:
            cabi.set_debug_flags(dbg)
        cabi.lib.mlapo_kernel_name.restype = ctypes.c_char_p
        out = {}
        if ksum:
            for j, v in enumerate(ksum):
                out[cabi.lib.mlapo_kernel_name(j).decode()] = v / kcount * 1e3
        return out

    def chain_spans_us(self, dbg=0):
        """Where each kernel sits inside a call when the calls run chained (programmatic dependent launch, back to back): from the
        per-CTA global-timer marks (debug bit3), relative to the first CTA of the call's first kernel:
        {kernel: [first CTA start, last CTA end]}.  None when the marks are not available for this launch shape.
        The marks exist only in the instrumented build of the library; run_ours() gets them from a child process that loads it
        (chain_spans_from_instrumented_build), so the timed legs of this process never run instrumented kernels."""
        torch, cabi, dev = self.torch, self.cabi, self.dev
        if not cabi.lib.mlapo_debug_is_instrumented():
            return None
        try:
            cabi.set_debug_flags(8 | dbg)
            lay = cabi.workspace_layout(self.plist[0])
            ws = torch.zeros(lay.total_bytes, dtype=torch.uint8, device=dev)
            for i in range(2 * self.n_sets):
                cabi.launch(self.plist[i % self.n_sets], ws, self.stream.cuda_stream)
            torch.cuda.synchronize(dev)
            tl = ws[lay.timeline:lay.timeline + 4 * 4096 * 16 * 8].view(torch.int64).reshape(4, 4096, 16).cpu()
            spans = {}
            for k, name in ((2, "quant_hidden"), (0, "mm1"), (3, "norm_rope_cache"), (1, "q_path")):
                a = tl[k]
                a = a[(a[:, 0] != 0) & (a[:, 15] != 0)]
                if a.shape[0]:
                    spans[name] = (int(a[:, 0].min()), int(a[:, 15].max()))
            if not spans:
                return None
            t0 = min(v[0] for v in spans.values())
            return {k: [round((a - t0) / 1e3, 2), round((b - t0) / 1e3, 2)] for k, (a, b) in spans.items()}
        except Exception as e:  # noqa
            print(f"[bench] chain spans unavailable: {type(e).__name__}: {e}", file=sys.stderr)
            return None
        finally:
            cabi.set_debug_flags(dbg)


def chain_spans_from_instrumented_build(wid, dbg, local, flags):
    """chain_spans_us() of workload `wid`, measured in a child process that loads the INSTRUMENTED library (per-CTA clock marks; its
    kernels are 3-4 % slower than the shipped ones, which is why nothing timed runs it).  Rank-local, no process group."""
    import subprocess
    from sgl_kernel_npu import cabi
    if not os.path.exists(cabi.INSTRUMENTED_LIB):
        print(f"[bench] chain spans unavailable: {cabi.INSTRUMENTED_LIB} not built", file=sys.stderr)
        return None
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT")}
    env["MLAPO_LIB"] = cabi.INSTRUMENTED_LIB
    cmd = [sys.executable, os.path.abspath(__file__), "--chain-spans-only", "--workload", wid, "--debug-flags", str(dbg), "--device-index",
           str(local)] + ([] if flags else ["--safe-start"])
    try:
        r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
        lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
        if r.returncode != 0 or not lines:
            print(f"[bench] chain spans unavailable: child rc={r.returncode}: {r.stderr[-300:]}", file=sys.stderr)
            return None
        return json.loads(lines[-1]).get("chain_spans_us")
    except Exception as e:  # noqa: BLE001
        print(f"[bench] chain spans unavailable: {type(e).__name__}: {e}", file=sys.stderr)
        return None


def run_chain_spans_only(args):
    """Child of chain_spans_from_instrumented_build(): one JSON line {"chain_spans_us": ...} from the instrumented library."""
    import torch
    from sgl_kernel_npu import cabi
    torch.cuda.set_device(args.device_index)
    dev = torch.device("cuda", args.device_index)
    flags = 0 if args.safe_start else cabi.FLAG_INPUT_COMPLETE
    wl = Workload(torch, cabi, dev, 0, 1, args.workload, flags=flags, use_graph=False, nz=False, n_sets=2)
    print(json.dumps({"chain_spans_us": wl.chain_spans_us(args.debug_flags), "instrumented": bool(cabi.lib.mlapo_debug_is_instrumented())}))


def block_stats(block_ms, steps):
    s = sorted(block_ms)
    med = s[len(s) // 2]
    return med / steps, s[0] / steps, s[-1] / steps


# ----------------------------------------------------------------------------------------------- e2e (host buffers)
def time_e2e(torch, wl, ranks, steps, call, label):
    """The metric end to end: every step copies its inputs from pinned host memory and reads its results back to pinned host
    memory.  The three parts run on three streams (copy-in, the operator, copy-out) ordered by events, with two input buffers,
    so that step i's read-back (the long pole) overlaps step i+1's copy-in and compute, as an asynchronous caller would."""
    dev, stream = wl.dev, wl.stream
    hidden_h, cos_h, sin_h, slots_h = wl.host
    q0_h = torch.empty((wl.n, wl.h, 512), dtype=wl.q0_dt).pin_memory()
    q1_h = torch.empty((wl.n, wl.h, 64), dtype=torch.bfloat16).pin_memory()
    in_bufs = [tuple(torch.empty_like(t, device=dev) for t in wl.host) for _ in range(2)]
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    ev_in = [torch.cuda.Event() for _ in range(2)]
    ev_free = [torch.cuda.Event() for _ in range(2)]          # the operator has finished reading input buffer b
    ev_out = [torch.cuda.Event() for _ in range(wl.n_sets)]   # the read-back of output set k has finished
    ev_comp = [torch.cuda.Event() for _ in range(wl.n_sets)]
    for e in ev_free + ev_out:
        e.record(stream)

    def e2e_step(i):
        b, k = i % 2, i % wl.n_sets
        hid_d, cos_d, sin_d, sl_d = in_bufs[b]
        with torch.cuda.stream(s_in):
            s_in.wait_event(ev_free[b])
            hid_d.copy_(hidden_h, non_blocking=True)
            cos_d.copy_(cos_h, non_blocking=True)
            sin_d.copy_(sin_h, non_blocking=True)
            sl_d.copy_(slots_h, non_blocking=True)
            ev_in[b].record(s_in)
        stream.wait_event(ev_in[b])
        stream.wait_event(ev_out[k])
        o0, o1 = wl.outs[k]
        call(wl.sets[k], hid_d, cos_d, sin_d, sl_d, o0, o1)
        ev_free[b].record(stream)
        ev_comp[k].record(stream)
        with torch.cuda.stream(s_out):
            s_out.wait_event(ev_comp[k])
            q0_h.copy_(o0, non_blocking=True)
            q1_h.copy_(o1, non_blocking=True)
            ev_out[k].record(s_out)

    def drain():  # the timed region ends when the last read-back has landed
        for e in ev_out:
            stream.wait_event(e)

    for i in range(3):
        e2e_step(i)
    drain()
    torch.cuda.synchronize(dev)
    ranks.barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record(stream)
    s_in.wait_event(f0)   # nothing of the timed steps starts before the start mark
    s_out.wait_event(f0)
    for i in range(steps):
        e2e_step(i)
    drain()
    f1.record(stream)
    torch.cuda.synchronize(dev)
    ranks.barrier()
    ms_local = f0.elapsed_time(f1)
    ms = ranks.max([ms_local])[0]
    h2d = sum(t.numel() * t.element_size() for t in wl.host)
    d2h = sum(t.numel() * t.element_size() for t in (q0_h, q1_h))
    return {"ms": ms, "ms_local": ms_local, "steps": steps, "h2d": h2d, "d2h": d2h, "api": label}


def numa_node_of_this_process():
    try:
        cpus = sorted(os.sched_getaffinity(0))
        nodes = set()
        for node in os.listdir("/sys/devices/system/node"):
            if node.startswith("node"):
                lst = open(f"/sys/devices/system/node/{node}/cpulist").read().strip()
                have = set()
                for part in lst.split(","):
                    a, _, b = part.partition("-")
                    have.update(range(int(a), int(b or a) + 1))
                if have & set(cpus):
                    nodes.add(int(node[4:]))
        return sorted(nodes)
    except Exception:
        return None


# ----------------------------------------------------------------------------------------------- our arm
def hbm_record(wl, ms_step, peak, peak_src):
    op_bytes = algorithmic_bytes(wl.n, wl.h, wl.d, wl.cache_mode)
    ach = op_bytes / (ms_step * 1e-3) / 1e9
    return {"bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4),
            "algorithmic_bytes": op_bytes, "step_us": round(ms_step * 1e3, 2), "frac_of_8TBs_nominal": round(ach / 8000.0, 4),
            "ideal_us_at_peak": round(op_bytes / (peak * 1e9) * 1e6, 2), "peak_source": peak_src}


def run_ours(args):
    import torch
    from sgl_kernel_npu import cabi   # raises if the CUDA library is missing: no fallback

    rank, world, local = rank_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    nvml_index = torch.cuda._get_nvml_device_index(dev) if hasattr(torch.cuda, "_get_nvml_device_index") else local
    try:  # run on the CPUs next to this GPU, so that the pinned host buffers of the e2e leg are allocated on its NUMA node
        import pynvml
        pynvml.nvmlInit()
        pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(nvml_index))
    except Exception:
        pass
    ranks = Ranks("nccl", dev)
    dbg = args.debug_flags
    if dbg:
        cabi.set_debug_flags(dbg)
    # The bench enqueues its calls behind torch copy kernels and behind this library's own kernels only, none of which releases its
    # dependents early, so it may state MLAPO_FLAG_INPUT_COMPLETE (include/mla_preprocess_b200.h); --safe-start times the default.
    flags = 0 if args.safe_start else cabi.FLAG_INPUT_COMPLETE
    peak, peak_src = load_peaks()
    sampler = ClockSampler(nvml_index)
    use_graph = not args.no_graph

    wl = Workload(torch, cabi, dev, rank, world, args.workload, flags=flags, use_graph=use_graph, nz=True)
    block_ms = wl.time_blocks(ranks, args.steps, args.warmup, sampler)
    ms_step, ms_min, ms_max = block_stats(block_ms, args.steps)
    clocks = sampler.summary()

    # ---- e2e through the operator (torch.ops.npu.mla_preprocess, FRACTAL_NZ weights) and through the ctypes mirror
    import sgl_kernel_npu  # noqa: F401  (registers torch.ops.npu.*; raises if the operator library is missing)
    if flags:
        os.environ.setdefault("SGL_KERNEL_NPU_MLAPO_INPUT_COMPLETE", "1")
    op = torch.ops.npu.mla_preprocess

    def call_op(w, hid, c, s, sl, o0, o1):
        op(hid, w["gamma0"], w["beta0"], w["wdqkv_nz"], w["descale0"], w["gamma1"], w["beta1"], w["wuq_nz"], w["descale1"], w["gamma2"],
           c, s, w["wuk"], wl.kv, wl.kvr, sl, w["quant_scale0"], w["quant_offset0"], w["bias0"], w["quant_scale1"], w["quant_offset1"],
           w["bias1"], ctkv_scale=w["ctkv_scale"], q_nope_scale=w["q_nope_scale"], cache_mode=wl.cache_mode, quant_mode=QUANT_MODE,
           q_out0=o0, kv_cache_out0=wl.kv, q_out1=o1, kv_cache_out1=wl.kvr)

    def call_ctypes(w, hid, c, s, sl, o0, o1):
        cabi.mla_preprocess(hid, None, None, w["wdqkv_km"], w["descale0"], w["gamma1"], w["beta1"], w["wuq_km"], w["descale1"],
                            w["gamma2"], c, s, w["wuk"], wl.kv, wl.kvr, sl, w["quant_scale0"], w["quant_offset0"], w["bias0"],
                            w["quant_scale1"], w["quant_offset1"], w["bias1"], ctkv_scale=w["ctkv_scale"],
                            q_nope_scale=w["q_nope_scale"], cache_mode=wl.cache_mode, quant_mode=QUANT_MODE, q_out0=o0,
                            kv_cache_out0=wl.kv, q_out1=o1, kv_cache_out1=wl.kvr, workspace=wl.ws, flags=flags)

    e2e_steps = max(args.steps, min(300, int(math.ceil(MIN_TIMED_MS / max(ms_step * 10, 1e-3)))))
    for k in range(wl.n_sets):   # first sight of each NZ weight repacks it (cached): warm-up, not timed
        call_op(wl.sets[k], wl.hidden, wl.cos, wl.sin, wl.slots, *wl.outs[k])
    torch.cuda.synchronize(dev)
    e_op = time_e2e(torch, wl, ranks, e2e_steps, call_op, "torch.ops.npu.mla_preprocess (FRACTAL_NZ weights, repack cached) -> mla_preprocess_launch (C ABI)")
    e_ct = time_e2e(torch, wl, ranks, e2e_steps, call_ctypes, "sgl_kernel_npu.cabi.mla_preprocess (ctypes, K-major weights, caller workspace)")

    kernels = wl.kernel_times_us(dbg)
    spans = chain_spans_from_instrumented_build(args.workload, dbg, local, flags) if (wl.launches_per_call > 1 and rank == 0) else None

    # ---- sub-records (default invocation only): prefill, the Kimi-K2 rank in every cache mode, TP head shards
    sub = {}
    if args.workload == "c2" and not args.no_subrecords:
        del wl.graph
        torch.cuda.empty_cache()
        since = len(sampler.samples)
        w3 = Workload(torch, cabi, dev, rank, world, "c3", flags=flags, use_graph=False, n_sets=2)
        b3 = w3.time_blocks(ranks, 4, 3, sampler, min_ms=60.0, max_blocks=40)
        ms3, ms3_min, ms3_max = block_stats(b3, 4)
        k3 = w3.kernel_times_us(dbg)
        i8, bf_fl = algorithmic_ops(w3.n, w3.h, w3.d)
        clocks3 = sampler.summary(since)
        hbm3 = hbm_record(w3, ms3, peak, peak_src)
        del w3
        torch.cuda.empty_cache()
        i8_peak = measured_int8_peak() if rank == 0 else None
        t3 = ms3 * 1e-3
        meas = i8_peak["burst"]["tops"] if i8_peak else None
        sub["prefill"] = {
            "config": workload_config("c3", world), "tokens_per_s": round(job_tokens(WORKLOADS["c3"][0], world, "replica") / t3, 1),
            "ms_per_step": round(ms3, 4), "ms_per_step_min": round(ms3_min, 4), "ms_per_step_max": round(ms3_max, 4),
            "steps_per_block": 4, "blocks": len(b3), "launch": "eager (1 call per step, 4 chained kernels)", "kernels_us_serialised": {k: round(v, 1) for k, v in k3.items()},
            "clocks": clocks3,
            "roofline": {"bound": "tensor", "unit": "TOP/s (int8)", "achieved": round(i8 / t3 / 1e12, 1), "int8_ops": i8, "bf16_flops": bf_fl,
                         "peak_nominal": INT8_NOMINAL_TOPS, "frac_nominal": round(i8 / t3 / 1e12 / INT8_NOMINAL_TOPS, 4),
                         "peak_measured": meas, "frac_measured": round(i8 / t3 / 1e12 / meas, 4) if meas else None,
                         "peak_measured_sustained": i8_peak["sustained"]["tops"] if i8_peak else None,
                         "peak_source": "tools/microbench/mlapo_microbench.cu: tcgen05.mma.kind::i8 M128 N256 K32 on resident tiles, every SM"
                                        if i8_peak else "not measured (helper library missing); 2 x MEASURED_PEAKS.json bf16_tflops would be %.0f" % (2 * tensor_peak_bf16()),
                         "frac_int8_plus_bf16_nominal": round((i8 / (INT8_NOMINAL_TOPS * 1e12) + bf_fl / 2.25e15) / t3, 4)},
            "hbm": hbm3,
        }
        c5 = {}
        for cm in ("krope_ctkv", "nzcache", "int8_nzcache"):
            w5 = Workload(torch, cabi, dev, rank, world, "c5", cache_mode=cm, flags=flags, use_graph=use_graph)
            b5 = w5.time_blocks(ranks, args.steps, 3, None)
            m5, m5min, m5max = block_stats(b5, args.steps)
            c5[cm] = {"tokens_per_s": round(job_tokens(w5.n, world, "replica") / (m5 * 1e-3), 1), "ms_per_step": round(m5, 6),
                      "ms_per_step_min": round(m5min, 6), "ms_per_step_max": round(m5max, 6), "launch": w5.launch_mode,
                      "roofline_op": hbm_record(w5, m5, peak, peak_src)}
            del w5
        sub["c5"] = {"config": workload_config("c5", world, cache_mode="all three"), "per_cache_mode": c5}
        if world > 1:
            w4 = Workload(torch, cabi, dev, rank, world, "c4", flags=flags, use_graph=use_graph)
            b4 = w4.time_blocks(ranks, args.steps, 3, None)
            m4, m4min, m4max = block_stats(b4, args.steps)
            sub["tp"] = {"config": workload_config("c4", world), "tp": world, "heads_per_rank": w4.h,
                         "tokens_per_s": round(job_tokens(w4.n, world, "tp") / (m4 * 1e-3), 1), "per_rank_us": round(m4 * 1e3, 2),
                         "per_rank_us_min": round(m4min * 1e3, 2), "per_rank_us_max": round(m4max * 1e3, 2), "scaling": "strong",
                         "launch": w4.launch_mode, "roofline_op_per_rank": hbm_record(w4, m4, peak, peak_src)}
            del w4
        torch.cuda.empty_cache()

    if rank == 0:
        tokens_total = job_tokens(wl.n, world, wl.sharding)
        value = tokens_total / (ms_step * 1e-3)
        dom = max(kernels, key=kernels.get) if kernels else None
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if dom and os.path.exists(tpath):
            try:
                traffic = json.load(open(tpath)).get(args.workload, {}).get(dom)
            except Exception:
                traffic = None
        roofline = None
        if dom:
            kb = kernel_algorithmic_bytes(dom, wl.n, wl.h, wl.d, wl.cache_mode)
            ach = kb / (kernels[dom] * 1e-6) / 1e9
            span = None
            if spans:
                key = "q_path" if dom.startswith("q_") else "mm1" if dom.startswith("mm1") else dom
                if key in spans:
                    span = round(spans[key][1] - spans[key][0], 2)
            roofline = {"bound": "hbm", "kernel": dom, "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                        "frac": round(ach / peak, 4), "traffic": traffic, "algorithmic_bytes": kb,
                        "kernel_us": round(kernels[dom], 2),
                        "kernel_us_is": "serialised pass: CUDA events around each kernel, no programmatic overlap with its neighbours",
                        "kernel_us_in_chain": span,
                        "kernel_us_in_chain_is": "first CTA start to last CTA end of the kernel inside back-to-back chained calls (per-CTA global-timer marks, taken by a child process with the instrumented build of the library, whose kernels are 3-4 % slower than the shipped ones timed everywhere else); it starts under its predecessor's tail",
                        "peak_source": peak_src}
            if wl.n >= 2048 and dom in ("q_fused", "q_pair"):
                ops = 2.0 * wl.n * 1536 * 192 * wl.h + 2.0 * wl.n * 128 * 512 * wl.h
                tach = ops / (kernels[dom] * 1e-6) / 1e12
                roofline.update({"bound": "tensor", "achieved": round(tach, 1), "peak": INT8_NOMINAL_TOPS, "unit": "TOP/s",
                                 "frac": round(tach / INT8_NOMINAL_TOPS, 4), "algorithmic_ops": ops, "peak_source": "nominal dense int8"})
        e2e_value = tokens_total * e_op["steps"] / (e_op["ms"] * 1e-3)
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_step, 6), "ms_per_step_min": round(ms_min, 6), "ms_per_step_max": round(ms_max, 6),
            "blocks": len(block_ms), "timed_ms": round(sum(block_ms), 2),
            "higher_is_better": True, "scaling": "strong" if wl.sharding == "tp" else "weak",
            "vs_baseline": None, "dtype": "s8 x s8 -> s32 (MM1/MM2), bf16 -> f32 (absorb), bf16 activations", "data": "synthetic",
            "config": workload_config(args.workload, world), "launch": wl.launch_mode,
            "start": "MLAPO_FLAG_INPUT_COMPLETE (the calls follow torch copies and this library's own kernels)" if flags else "default (waits for the previous kernel)",
            "e2e": {"value": round(e2e_value, 1), "unit": "tokens/s", "h2d_bytes_per_step": e_op["h2d"], "d2h_bytes_per_step": e_op["d2h"],
                    "steps": e_op["steps"], "api": e_op["api"],
                    "overlap": "copy-in / operator / read-back of consecutive steps on three streams",
                    "pcie_gbs_this_rank": {"h2d_plus_d2h": round((e_op["h2d"] + e_op["d2h"]) * e_op["steps"] / (e_op["ms_local"] * 1e-3) / 1e9, 1),
                                           "d2h": round(e_op["d2h"] * e_op["steps"] / (e_op["ms_local"] * 1e-3) / 1e9, 1)},
                    "host_numa_nodes_of_pinned_buffers": numa_node_of_this_process(),
                    "limiter": "PCIe read-back of q_out0 (%.1f MB per step)" % (e_op["d2h"] / 1e6),
                    "ctypes": {"value": round(tokens_total * e_ct["steps"] / (e_ct["ms"] * 1e-3), 1), "api": e_ct["api"]}},
            "gpu_launches": wl.launches_per_call * args.steps * len(block_ms),
            "clocks": clocks,
            "roofline": roofline,
            "roofline_op": hbm_record(wl, ms_step, peak, peak_src),
            "kernels_us": {k: round(v, 2) for k, v in kernels.items()},
            "kernels_us_is": "serialised pass (their sum exceeds the chained step)",
            "chain_spans_us": spans,
        }
        line.update(sub)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args.workload, budget_s=15.0)
        print(json.dumps(line), flush=True)
    ranks.close()


# ----------------------------------------------------------------------------------------------- CPU legs (oracle port)
def _oracle_case(workload, tokens):
    import torch
    from oracle import mla_oracle as orc     # permitted: cpu_baseline / --impl reference legs only
    n, h, d, cache_mode, _ = WORKLOADS[workload]
    case = orc.gen_case(N=tokens, H=h, D=d, dtype=torch.bfloat16, cache_mode=cache_mode, seed=42,
                        num_blocks=max(16, (tokens * 4 + 127) // 128 + 8))
    return orc, case


def _time_oracle(orc, case, reps):
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        orc.run_case(case, quant_mode=QUANT_MODE, var_mode="torch")
        ts.append(time.perf_counter() - t0)
    return ts


def _sample_tokens(workload, per_step_budget_s):
    """Tokens per CPU step: the full batch when a step of it fits the budget (probed with the full batch itself for decode sizes:
    the cost of a call is far from linear in the tokens), else a bounded sample of it."""
    n = WORKLOADS[workload][0]
    if n <= 512:
        orc, case = _oracle_case(workload, n)
        t = min(_time_oracle(orc, case, 2))
        if t <= per_step_budget_s:
            return n
        return int(max(1, n * per_step_budget_s / t))
    orc, probe = _oracle_case(workload, 64)
    t64 = min(_time_oracle(orc, probe, 2))
    return int(max(1, min(n, 64 * per_step_budget_s / max(t64, 1e-4))))


def cpu_baseline(workload, budget_s=15.0):
    """The same measurement the reference arm makes (full batch per call whenever one call fits ~1.5 s), on all host threads."""
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    n, h = WORKLOADS[workload][0], WORKLOADS[workload][1]
    tokens = _sample_tokens(workload, 1.5)
    orc, case = _oracle_case(workload, tokens)
    t = _time_oracle(orc, case, 2)[-1]                     # warm-up + calibration
    reps = max(3, min(20, int(budget_s / max(t, 1e-3))))
    ts = sorted(_time_oracle(orc, case, reps))
    med = ts[len(ts) // 2]
    return {"value": round(tokens / med, 2), "unit": "tokens/s", "cores": cores, "kind": "port",
            "sample": f"{tokens} of {n} tokens x all {h} heads per call, oracle/mla_oracle.py (torch CPU, {cores} threads), "
                      f"median of {reps} calls, {med * 1e3:.0f} ms/call"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    n, h, d, cache_mode, desc = WORKLOADS[args.workload]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # bounded sample: choose tokens per step so that (steps + warmup) steps end within ~3 minutes
    tokens = _sample_tokens(args.workload, 150.0 / max(1, args.steps + args.warmup))
    orc, case = _oracle_case(args.workload, tokens)
    _time_oracle(orc, case, max(1, min(args.warmup, 5)))
    t0 = time.perf_counter()
    _time_oracle(orc, case, args.steps)
    dt = time.perf_counter() - t0
    value = tokens * args.steps / dt
    sample = (f"{tokens} of {n} tokens x all {h} heads per step, oracle port of golden2_pytorch "
              f"(tests/python/sgl_kernel_npu/test_mla_preprocess.py:407-484), torch CPU {cores} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 2), "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(dt * 1e3 / args.steps, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "s8 x s8 -> s32 (MM1/MM2), bf16 -> f32 (absorb), bf16 activations", "data": "synthetic",
        "config": workload_config(args.workload, world),
        "cpu_baseline": {"value": round(value, 2), "unit": "tokens/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(value, 2), "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "one CPU process on rank 0 whatever --gpus says: only the N=1 ratio against the GPU arm compares like with like",
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of a CUDA graph of one weight rotation")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-subrecords", action="store_true", help="skip the prefill / c5 / tp sub-records of the default line")
    ap.add_argument("--safe-start", action="store_true", help="do not state MLAPO_FLAG_INPUT_COMPLETE (the library's default start)")
    ap.add_argument("--debug-flags", type=int, default=0, help="A/B: extra mlapo_set_debug_flags bits")
    ap.add_argument("--chain-spans-only", action="store_true", help=argparse.SUPPRESS)   # child mode, see chain_spans_from_instrumented_build
    ap.add_argument("--device-index", type=int, default=0, help=argparse.SUPPRESS)
    args = ap.parse_args()
    if args.chain_spans_only:
        run_chain_spans_only(args)
        return
    if args.impl == "reference":
        run_reference(args)
    else:
        try:
            run_ours(args)
        except Exception:
            if args.debug_flags & 2048:   # a kernel died with the stuck-wait log on: say which barrier it was waiting on
                try:
                    from sgl_kernel_npu import cabi
                    buf = (ctypes.c_uint * 1024)()
                    n = cabi.lib.mlapo_debug_read_wait_log(buf, 1024)
                    print(f"[bench] stuck barrier waits logged: {n}", file=sys.stderr)
                    seen = {}
                    for i in range(min(max(n, 0), 255)):   # the log holds the first 255 waiters; group them by warp
                        key = (buf[1 + 4 * i], buf[2 + 4 * i] // 32, buf[3 + 4 * i] & 0xFFFFFF, buf[4 + 4 * i])
                        seen[key] = seen.get(key, 0) + 1
                    for (cta, warp, bar, par), cnt in sorted(seen.items()):
                        print(f"[bench]   cta {cta} warp {warp} ({cnt} threads) barrier smem {bar:#x} parity {par:#x}", file=sys.stderr)
                except Exception as e2:  # noqa: BLE001
                    print(f"[bench] wait log unavailable: {e2}", file=sys.stderr)
            raise


if __name__ == "__main__":
    main()
