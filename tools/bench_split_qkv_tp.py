"""Timing of split_qkv_tp_rmsnorm_rope on a tensor-parallel group (torchrun, one rank per GPU): the one-kernel peer exchange against
local means -> NCCL all-reduce -> apply.  Not part of bench.py's contract; numbers go to profiles/README.md.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/bench_split_qkv_tp.py"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sgl-kernel-npu_b200")]
import torch
import torch.distributed as dist

from sgl_kernel_npu.norm import split_qkv_tp_rmsnorm_rope as m


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    dist.init_process_group("nccl")
    dev = torch.device("cuda", torch.cuda.current_device())
    dt = torch.bfloat16
    for (b, qh, kvh, hd, rd) in [(64, 6144, 1024, 128, 128), (256, 6144, 1024, 128, 128), (2048, 6144, 1024, 128, 128), (8192, 6144, 1024, 128, 128)]:
        sets = max(2, min(32, int(300e6 // (4 * b * (qh + 2 * kvh))) + 1))  # rotate inputs over > L2
        xs = [torch.randn(b, qh + 2 * kvh, device=dev).to(dt) for _ in range(sets)]
        cos, sin = torch.rand(b, rd, device=dev).to(dt), torch.rand(b, rd, device=dev).to(dt)
        qw, kw = torch.randn(qh, device=dev).to(dt), torch.randn(kvh, device=dev).to(dt)
        line = {"shape": [b, qh, kvh, hd, rd], "tp_world": world, "launch": "graph" if "--graph" in sys.argv else "eager"}
        for exchange in (("peer", "allreduce") if world > 1 else ("single",)):
            m.EXCHANGE = exchange
            call = lambda i: m.split_qkv_tp_rmsnorm_rope(xs[i % sets], cos, sin, qh, kvh, hd, 1e-6, qw, kw, rd, world, dist.group.WORLD)
            for i in range(3):
                call(i)
            torch.cuda.synchronize()
            dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if "--graph" in sys.argv:  # device time without the host's share: `sets` calls per graph, outputs kept alive
                graph, stream, outs = torch.cuda.CUDAGraph(), torch.cuda.Stream(), []
                with torch.cuda.stream(stream):
                    with torch.cuda.graph(graph, stream=stream):
                        for i in range(sets):
                            outs.append(call(i))
                    graph.replay()
                    torch.cuda.synchronize()
                    dist.barrier()
                    reps = 20
                    e0.record(stream)
                    for _ in range(reps):
                        graph.replay()
                    e1.record(stream)
                reps *= sets
            else:
                reps = 5 * sets
                e0.record()
                for i in range(reps):
                    call(i)
                e1.record()
            torch.cuda.synchronize()
            us = torch.tensor([e0.elapsed_time(e1) * 1e3 / reps], device=dev)
            dist.all_reduce(us, op=dist.ReduceOp.MAX)
            line[exchange + "_us"] = round(float(us), 2)
            line[exchange + "_GBps_per_gpu"] = round(4 * b * (qh + 2 * kvh) / float(us) / 1e3, 1)
        if rank == 0:
            print(json.dumps(line), flush=True)
    if "--graph" in sys.argv:  # captured graphs hold the communicator: tearing the group down under them stalls for minutes
        sys.stdout.flush()
        os._exit(0)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
