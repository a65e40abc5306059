"""Chains across GPUs: contiguous blocks, nothing exchanged while sampling, one reduction at the end.

The reference's only multi-run parallelism is separate OS processes whose logs are combined afterwards
(PostProcess/PostProcess.F90:229-304).  Here rank r of `world` owns global chains [start, start+count) and
seeds them with (global index + 1); `reduce_stats` / `gather_per_chain` are the single collective step
(NCCL on GPUs, gloo in the CPU tests) that replaces the log scraping.
"""
from __future__ import annotations

import numpy as np

SUM_KEYS = ("sum_e", "sum_e2", "production", "n_stat", "prod_acc", "nsuc", "step", "sitepairs", "n_clamped",
            "uniforms_used")


def chain_block(rank: int, world: int, n_total: int) -> tuple[int, int]:
    """Contiguous block of rank `rank`; the first n_total % world ranks get one chain more."""
    base, rem = divmod(n_total, world)
    count = base + (1 if rank < rem else 0)
    start = rank * base + min(rank, rem)
    return start, count


def global_seeds(rank: int, world: int, n_total: int) -> np.ndarray:
    start, count = chain_block(rank, world, n_total)
    return np.arange(start + 1, start + count + 1, dtype=np.uint64)


def reduce_stats(stats: dict, device=None) -> dict:
    """Sum the additive per-chain counters over all chains of all ranks (one all-reduce of a small fp64 vector)."""
    import torch
    import torch.distributed as dist
    vec = torch.tensor([float(np.sum(stats[k])) for k in SUM_KEYS], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(vec, op=dist.ReduceOp.SUM)
    out = dict(zip(SUM_KEYS, vec.tolist()))
    n = max(out["n_stat"], 1.0)
    out["mean_e"] = out["sum_e"] / n
    out["var_e"] = max(out["sum_e2"] / n - out["mean_e"] ** 2, 0.0)
    out["acceptance"] = out["nsuc"] / max(out["step"], 1.0)
    return out


def gather_per_chain(local: np.ndarray, n_total: int, device=None) -> np.ndarray:
    """All-gather one fp64 value per chain into global chain order (blocks may differ by one chain in length)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return np.asarray(local, dtype=np.float64).copy()
    world = dist.get_world_size()
    width = -(-n_total // world)
    buf = torch.zeros(width, dtype=torch.float64, device=device)
    buf[: len(local)] = torch.as_tensor(np.asarray(local, dtype=np.float64), device=device)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)
    out = np.empty(n_total)
    for r in range(world):
        s, c = chain_block(r, world, n_total)
        out[s:s + c] = parts[r][:c].cpu().numpy()
    return out
