"""Multi-GPU plumbing of the path: sharding by independent frame sequence.

The recurrence runs along time inside one sequence (mainVideoUnshaded.py:417-465); sequences never interact and the
3.6 MB of weights are replicated, so there is NO collective on the data path.  One process per GPU (torchrun);
``torch.distributed`` (NCCL on the GPU box, gloo in the CPU tests) is used only for the barrier around the timed
region, the max-over-ranks of the elapsed time and the gather of small per-rank records / checksums.
"""
import torch
import torch.distributed as dist


def world():
    """(rank, world_size) of the current process; (0, 1) when torch.distributed is not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def sequence_range(n_sequences, rank, world_size):
    """Strong sharding: contiguous, balanced split of ``n_sequences`` over the ranks -> [start, stop).
    The first ``n_sequences % world_size`` ranks own one extra sequence; a rank may own none."""
    if n_sequences < 0 or world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError("bad sharding arguments")
    base, extra = divmod(n_sequences, world_size)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def weak_range(sequences_per_gpu, rank):
    """Weak sharding (BASELINE.json configs[4]): every rank owns ``sequences_per_gpu`` sequences of its own."""
    if sequences_per_gpu < 0 or rank < 0:
        raise ValueError("bad sharding arguments")
    return rank * sequences_per_gpu, (rank + 1) * sequences_per_gpu


def rank_seed(base_seed, rank):
    """Per-rank data seed (data seed + rank); weights use one common seed so every replica is identical."""
    return int(base_seed) + int(rank)


def barrier(device=None):
    if dist.is_available() and dist.is_initialized():
        dist.barrier()
    if device is not None and torch.device(device).type == "cuda":
        torch.cuda.synchronize(device)


def max_over_ranks(value, device="cpu"):
    """Slowest rank's time: throughput of the whole job is total units / max-over-ranks elapsed."""
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value, device="cpu"):
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def gather_records(record):
    """All ranks contribute a small picklable record (units processed, elapsed ms, output checksum); every rank gets
    the list ordered by rank."""
    if dist.is_available() and dist.is_initialized():
        out = [None] * dist.get_world_size()
        dist.all_gather_object(out, record)
        return out
    return [record]


def aggregate_throughput(records, unit_scale=1e6):
    """Whole-job throughput from per-rank records {'units': n, 'ms': t}: sum(units) / max(ms)."""
    ms = max(r["ms"] for r in records)
    units = sum(r["units"] for r in records)
    return units / (ms * 1e-3) / unit_scale if ms > 0 else 0.0


__all__ = ["world", "sequence_range", "weak_range", "rank_seed", "barrier", "max_over_ranks", "sum_over_ranks",
           "gather_records", "aggregate_throughput"]
