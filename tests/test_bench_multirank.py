"""N>1 plumbing of bench.py on CPU: two gloo ranks rendezvous on 127.0.0.1, barrier, reduce the timing by MAX, and shard
heads the way the TP workload (BASELINE C4) does.  The data path itself has no collective."""
import os
import socket

import pytest
import torch
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import bench
    ranks = bench.Ranks(backend="gloo")
    ranks.barrier()
    got = ranks.max([1.0 + rank, 10.0 - rank])          # per-rank step time, e2e time
    h0, h1 = bench.shard_heads(128, world, rank)
    out.put((rank, got, (h0, h1), bench.job_tokens(512, world, "tp"), bench.job_tokens(256, world, "replica")))
    ranks.barrier()
    ranks.close()


def test_two_gloo_ranks_barrier_max_and_head_sharding():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(out.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, got, heads, tp_tokens, dp_tokens in res:
        assert got == [2.0, 10.0]                         # MAX over ranks, identical on every rank
        assert heads == (rank * 64, rank * 64 + 64)
        assert tp_tokens == 512 and dp_tokens == 512      # TP shares one batch; two replicas process 2 x 256
    assert {r[2] for r in res} == {(0, 64), (64, 128)}


def test_single_rank_needs_no_process_group(monkeypatch):
    import bench
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    r = bench.Ranks(backend="gloo")
    assert r.world == 1 and r.max([3.5]) == [3.5]
    r.barrier()
    with pytest.raises(ValueError):
        bench.shard_heads(128, 3, 0)


def test_algorithmic_bytes_match_survey():
    import bench
    # C2 = 69.66 MB + 256 * 163,204 B (+0.22 MB vectors) = 111.4 MB (+vectors)
    b = bench.algorithmic_bytes(256, 128, 7168, "krope_ctkv")
    assert abs(b - (69_664_768 + 256 * 163_204 + 8 * 2112 + 8 * 192 * 128 + 4 * 1536 + 2 * 512)) == 0
    assert abs(bench.algorithmic_bytes(8, 16, 7168, "krope_ctkv") / 1e6 - 22.2) < 0.2   # C1
