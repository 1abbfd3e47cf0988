"""Sequence-parallel sharding used by bench.py --gpus N ("replicas only").

One process per GPU, one independent sequence per rank; no collective on the frame path.
The only communication is the barrier around the timed region and the max-over-ranks of the
timed durations (NCCL on GPUs, gloo in the CPU tests)."""
import os


def rank_info():
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0")))


def sequence_seed(rank, base=1234):
    """Scene seed of the sequence a rank processes (BASELINE configs[4]: seeds 1234...1241)."""
    return base + rank


def assign_sequences(n_sequences, world):
    """sequence s -> rank s mod world; returns the list of sequence ids per rank."""
    return [[s for s in range(n_sequences) if s % world == r] for r in range(world)]


def job_throughput(units_per_rank, world, max_seconds):
    """Whole-job throughput: all ranks' units divided by the slowest rank's time."""
    return units_per_rank * world / max_seconds


def reduce_job(local_ms, local_counts, device=None):
    """MAX of the per-rank durations and SUM of the per-rank counters over the default process
    group (identity without one).  local_ms / local_counts: lists of floats."""
    import torch
    import torch.distributed as dist
    ms = torch.tensor(list(local_ms), dtype=torch.float64, device=device)
    cnt = torch.tensor(list(local_counts), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    return [float(v) for v in ms], [float(v) for v in cnt]
