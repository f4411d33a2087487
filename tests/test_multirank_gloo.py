"""World-size-2 run of the multi-rank host logic on CPU (gloo): the contiguous utterance shards, the max-over-ranks
timing reduction bench.py reports, the optional gather of result shards and the rank-0-only reporting path
(SURVEY.md section 8e).  The data path itself has no collective; the GPU kernels are covered by the `-m gpu` tests."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from visual_onoma_to_wave_b200.sharding import gather_shards, max_over_ranks, shard_batch, shard_bounds


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_items, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = shard_bounds(n_items, world, rank)
        job = torch.arange(n_items * 3, dtype=torch.float32).reshape(n_items, 3)       # stands for the utterance batch
        mine = shard_batch(job, world, rank)
        assert mine.shape[0] == hi - lo and torch.equal(mine, job[lo:hi])
        result = mine * 2.0                                                            # the rank's "waveforms"
        # timing: every rank reports its own device time and wall time; the job's time is the slowest rank's
        t_dev, t_wall = 10.0 + rank, 20.0 - rank
        red = max_over_ranks([t_dev, t_wall])
        assert red == [10.0 + world - 1, 20.0], red
        full = gather_shards(result, n_items, world, rank)
        assert torch.equal(full, job * 2.0)
        dist.barrier()
        if rank == 0:                                                                   # rank 0 alone prints the JSON line
            q.put({"value": n_items / (red[0] * 1e-3), "n_gpus": world, "utterances_per_gpu": hi - lo})
    finally:
        dist.destroy_process_group()


def test_two_ranks_shard_reduce_and_report():
    for n_items in (1024, 7):                      # even split (configs[3] at N=2) and a ragged one
        ctx = mp.get_context("spawn")
        q = ctx.SimpleQueue()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, 2, port, n_items, q)) for r in range(2)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0
        line = q.get()
        assert line["n_gpus"] == 2 and line["utterances_per_gpu"] == (n_items + 1) // 2
        assert abs(line["value"] - n_items / 11e-3) < 1e-6
        assert q.empty()                           # exactly one line, from rank 0


def test_max_over_ranks_without_a_process_group():
    assert max_over_ranks([1.5, 2.5]) == [1.5, 2.5]
