"""CPU, world_size 2 on gloo: the chain partitioning and the end-of-run statistics collective."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from montecarlo_b200 import partition


def test_chain_blocks_cover_everything_once():
    for n_total in (1, 7, 1024, 4096, 9216, 9217):
        for world in (1, 2, 3, 8):
            seen = []
            for r in range(world):
                s, c = partition.chain_block(r, world, n_total)
                seen.extend(range(s, s + c))
                assert list(partition.global_seeds(r, world, n_total)) == [i + 1 for i in range(s, s + c)]
            assert seen == list(range(n_total))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    s, c = partition.chain_block(rank, world, n_total)
    g = np.arange(s, s + c, dtype=np.float64)          # a per-chain quantity that encodes the global index
    stats = {k: np.zeros(c) for k in partition.SUM_KEYS}
    stats["sum_e"] = g * 2.0; stats["sum_e2"] = g * g * 4.0; stats["n_stat"] = np.ones(c)
    stats["nsuc"] = np.full(c, 3.0); stats["step"] = np.full(c, 10.0); stats["sitepairs"] = g + 1
    red = partition.reduce_stats(stats)
    allv = partition.gather_per_chain(g * 10.0, n_total)
    if rank == 0:
        q.put((red, allv))
    dist.destroy_process_group()


def test_two_ranks_reduce_and_gather():
    n_total, world = 11, 2
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    [p.start() for p in procs]
    red, allv = q.get()
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    g = np.arange(n_total, dtype=np.float64)
    assert red["sum_e"] == (2 * g).sum() and red["sitepairs"] == (g + 1).sum() and red["step"] == 10 * n_total
    assert abs(red["mean_e"] - (2 * g).mean()) < 1e-12 and abs(red["acceptance"] - 0.3) < 1e-15
    assert abs(red["var_e"] - (2 * g).var()) < 1e-9
    assert np.array_equal(allv, g * 10.0)
