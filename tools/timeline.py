"""GPU diagnostic: per-CTA clock marks of mm1_splitk and q_fused (debug flag bit3). Not a test.
Loads the INSTRUMENTED build of the library (the shipped one compiles the marks out): its kernels are 3-4 % slower, so the
periods printed here are for reading the structure of a call, not for quoting."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-npu_b200")]
os.environ.setdefault("MLAPO_LIB", os.path.join(ROOT, "sgl-kernel-npu_b200", "sgl_kernel_npu", "lib", "libmla_preprocess_b200_instr.so"))
import numpy as np
import torch

import bench
from sgl_kernel_npu import cabi

MM1 = ["gtimer", "start", "prologue", "upstream", "first_full", "last_full", "mma_done", "epi_issued", "stores_read", "end"]
QF = ["gtimer", "start", "prologue", "upstream", "first_full", "last_full", "mm2_done", "epi2_done", "unit0", "unit4", "drain_issued",
      "stores_read", "end"]


def main(workload="c2", pdl=False):
    n, h, d, cache_mode, _ = bench.WORKLOADS[workload]
    h = int(os.environ.get("MLAPO_TL_HEADS", h))   # e.g. 16 with c4: one rank of the TP8 split
    dev = torch.device("cuda:0")
    gen = torch.Generator(device=dev).manual_seed(1)
    sets = [bench.make_weight_set(torch, h, d, dev, gen) for _ in range(3)]
    hh, ch, sh, slh = bench.make_host_batch(torch, n, d, 64 * 128, 5)
    hid, cos, sin, sl = (t.to(dev) for t in (hh, ch, sh, slh))
    kv, kvr = bench.cache_tensors(torch, cache_mode, 64, dev)
    q0 = torch.zeros((n, h, 512), dtype=torch.int8 if cache_mode == "int8_nzcache" else torch.bfloat16, device=dev)
    q1 = torch.zeros((n, h, 64), dtype=torch.bfloat16, device=dev)
    cabi.set_debug_flags(8 | (0 if pdl else 1) | int(os.environ.get("MLAPO_EXTRA_FLAGS", "0")))
    ps = []
    for w in sets:
        ps.append(cabi.build_params(hid, w["wdqkv_km"], w["descale0"], w["gamma1"], w["beta1"], w["wuq_km"], w["descale1"], w["gamma2"],
                                    cos, sin, w["wuk"], kv, kvr, sl, w["quant_scale0"], w["quant_offset0"], w["bias0"], w["quant_scale1"],
                                    w["quant_offset1"], w["bias1"], w["ctkv_scale"], w["q_nope_scale"], cache_mode,
                                    "per_tensor_quant_asymm", q0, q1, flags=int(os.environ.get("MLAPO_TL_FLAGS", "1"))))
    lay = cabi.workspace_layout(ps[0])
    ws = torch.zeros(lay.total_bytes, dtype=torch.uint8, device=dev)
    for i in range(6):
        cabi.launch(ps[i % 3], ws)
    torch.cuda.synchronize()
    if pdl:  # steady state as the bench runs it: a CUDA graph of back-to-back calls, PDL edges inside
        gs = torch.cuda.Stream(dev)
        with torch.cuda.stream(gs):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=gs):
                for i in range(6):
                    cabi.launch(ps[i % 3], ws, gs.cuda_stream)
            gr.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(gs)
            for _ in range(20):
                gr.replay()
            e1.record(gs)
        torch.cuda.synchronize()
        print(f"period with the timeline marks on: {e0.elapsed_time(e1) * 1e3 / 120:.2f} us per call")
    tl = ws[lay.timeline:lay.timeline + 4 * 4096 * 16 * 8].view(torch.int64).reshape(4, 4096, 16).cpu().numpy()
    # whole-call picture from the global timer: [kernel] first CTA start .. last CTA end, relative to the quant kernel's start
    spans = {}
    for k, name in [(2, "quant_hidden"), (0, "mm1_splitk"), (3, "norm_rope_cache"), (1, "q_fused")]:
        a = tl[k]
        a = a[a[:, 0] != 0]
        if a.shape[0]:
            spans[name] = (a[:, 0].min(), np.median(a[:, 0]), a[:, 15].max(), a[:, 14].max() if k in (2, 3) else 0, np.median(a[:, 14]) if k in (2, 3) else 0, np.median(a[:, 15]))
    if spans:
        t0 = spans["quant_hidden"][0] if "quant_hidden" in spans else min(v[0] for v in spans.values())
        print("call timeline (us from quant start): kernel first-CTA-start / median-CTA-start / last-CTA-end")
        for name, (a0, am, a1, w, wm, em) in spans.items():
            extra = f"  upstream-wait passed: median {(wm - t0) / 1e3:7.2f} last {(w - t0) / 1e3:7.2f}; median CTA end {(em - t0) / 1e3:7.2f}" if w else ""
            print(f"   {name:16s} {(a0 - t0) / 1e3:7.2f} {(am - t0) / 1e3:7.2f} {(a1 - t0) / 1e3:7.2f}{extra}")
    if os.environ.get("MLAPO_TL_QUANT"):  # when do the quantise kernel's CTAs start, against the previous q kernel's CTA ends?
        qs_ = tl[2][tl[2][:, 0] != 0][:, 0]
        qe_ = tl[1][tl[1][:, 0] != 0][:, 15]
        t0_ = qs_.min()
        print("quant CTA starts (us), percentiles 0/10/25/50/75/90/100: " + " ".join(f"{(np.percentile(qs_, p) - t0_) / 1e3:.2f}" for p in (0, 10, 25, 50, 75, 90, 100)), "of", qs_.size)
        print("q CTA ends of the SAME call (us, same origin), percentiles 0/50/100: " + " ".join(f"{(np.percentile(qe_, p) - t0_) / 1e3:.2f}" for p in (0, 50, 100)))
    QP = ["gtimer", "start", "prologue", "upstream", "first_full"] + [f"u{i}:{w}" for i in range(3) for w in ("mm2_done", "acc_read", "drained")] + ["end"]
    stream = n > 256 and h * ((n + 255) // 256) > 148
    qname, qlabels = ("q_pair", QP) if stream else ("q_fused", QF)
    for k, (name, labels) in enumerate([("mm1_splitk", MM1), (qname, qlabels)]):
        a = tl[k]
        used = a[:, 1] != 0
        a = a[used]
        print(f"{name}: {a.shape[0]} CTAs; cycles since CTA start (median / p10 / p90 / max):")
        for j in range(2, len(labels)):
            dlt = a[:, j] - a[:, 1]
            dlt = dlt[a[:, j] != 0]
            if dlt.size:
                print(f"   {labels[j]:14s} {np.median(dlt):9.0f} {np.percentile(dlt, 10):9.0f} {np.percentile(dlt, 90):9.0f} {dlt.max():9.0f}")
        if name == "q_fused" and a[:, 13].max() > 0:
            order = np.argsort(a[:, 13])
            print("   per SM (smid: last_full, end since CTA start, CTA start offset ns):")
            slow = [i for i in order if a[i, 5] - a[i, 1] > 1.15 * np.median(a[:, 5] - a[:, 1])]
            print("   slow CTAs (smid:cta:last_full): " + " ".join(f"{int(a[i, 13])}:{i}:{int(a[i, 5] - a[i, 1])}" for i in slow))
            absf = (a[:, 0] - a[:, 0].min()) * 1.965 + (a[:, 5] - a[:, 1])
            print("   last_full by CTA id (k-cycles since the first CTA start): " + " ".join(f"{v / 1e3:.0f}" for v in absf))
            abse = (a[:, 0] - a[:, 0].min()) * 1.965 + (a[:, 12] - a[:, 1])
            print("   end by CTA id: " + " ".join(f"{v / 1e3:.0f}" for v in abse))
            print("   idle SMs: " + " ".join(str(x) for x in sorted(set(range(148)) - set(int(v) for v in a[:, 13]))))
        if a[:, 15].max() > 0:
            e = (a[:, 15] - a[:, 0].min()) / 1e3
            st = (a[:, 0] - a[:, 0].min()) / 1e3
            print(f"   absolute (us since the kernel's first CTA start): CTA start p10/p50/p90/max {np.percentile(st, 10):.2f} {np.median(st):.2f} {np.percentile(st, 90):.2f} {st.max():.2f};"
                  f" CTA end p10/p50/p90/max {np.percentile(e, 10):.2f} {np.median(e):.2f} {np.percentile(e, 90):.2f} {e.max():.2f}")
        g = a[:, 0]
        print(f"   CTA start skew (globaltimer ns): p50-min {np.median(g) - g.min():.0f}  max-min {g.max() - g.min():.0f}")
    if qname == "q_pair":
        tr = tl[1][1024:1024 + 8].reshape(-1)
        names = ["mm2_done", "qn_free", "ld0", "deq0", "ld1", "deq1", "ld2", "deq2", "ld3", "deq3", "rope_deq", "fence", "g_bar", "arrive", "rope_st", "fetch"]
        for it in range(2):
            b = it * 32
            print(f"dequant trace unit {it + 2} (cycles since its mm2_done mark): " + " ".join(f"{names[j]}={tr[b + j] - tr[b]}" for j in range(16)))
    cabi.set_debug_flags(0)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "c2", pdl="pdl" in sys.argv[2:])

