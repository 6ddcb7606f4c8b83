"""How independent image sequences are spread over the GPUs of one box (replicas only, no collective on the
data path).  Frames of one sequence are strictly sequential (reference src/tracking.cpp:162,268: trackFrames consumes the
previous frame's points), sequences never interact, so rank r of `world` simply owns its own sequences; torch.distributed is
used for the start/stop barrier and the max-over-ranks time only.  Pure host logic: covered on CPU by gloo tests."""


def owner_of(seq_id, world):
    """Rank that processes global sequence `seq_id` (C5: 64 sequences over 1/2/4/8 GPUs): round robin."""
    return seq_id % world


def sequences_of(rank, world, n_total):
    """Global ids of the sequences owned by `rank` when a fixed set of n_total sequences is sharded (strong scaling)."""
    return [s for s in range(n_total) if owner_of(s, world) == rank]


def stream_seeds(rank, streams_per_gpu, base_seed):
    """Weak scaling (what bench.py measures): every rank runs `streams_per_gpu` sequences of its own; the seeds of
    different ranks never overlap, and rank 0's are those of the single-GPU run."""
    return [base_seed + rank * streams_per_gpu + i for i in range(streams_per_gpu)]


def max_over_ranks(value, world, device=None):
    """Max of a per-rank scalar over all ranks (the job's time is that of its slowest rank)."""
    if world == 1:
        return float(value)
    import torch
    import torch.distributed as dist
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def aggregate_rate(units_per_rank, world, ms_max):
    """Whole-job throughput: units all ranks processed / the slowest rank's time."""
    return world * units_per_rank / (ms_max * 1e-3)
