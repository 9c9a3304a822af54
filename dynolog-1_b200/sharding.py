"""Multi-GPU host logic: the path shards by CPU id with no data-path
collective; the one exchange is the merge of counter-group partial sums, an integer
all-reduce followed by the scaling (never the other way round: sum-of-scaled !=
scaled-of-sum, hbt/src/perf_event/PerCpuBase.h:104-119 sums first,
CpuEventsGroup.h:449-459 scales the sum).

Pure torch / torch.distributed plumbing: works on gloo (CPU tests) and nccl."""
from __future__ import annotations

import torch


def shard_cpus(n_cpus: int, world: int, rank: int, mode: str = "block"):
    """CPU ids owned by `rank`.  `block`: contiguous ranges (what bench.py uses,
    cpu_base = first id); `mod`: cpu % world == rank.  A pure function of
    (n_cpus, world, rank) so the oracle can mirror it."""
    if mode == "mod":
        return [c for c in range(n_cpus) if c % world == rank]
    per = (n_cpus + world - 1) // world
    return list(range(min(rank * per, n_cpus), min((rank + 1) * per, n_cpus)))


def allreduce_partial_sums(sums: torch.Tensor, n_events: int, group=None) -> torch.Tensor:
    """sums: [I][3+E] int64 partial group sums of this rank's CPU shard, as written by
    hbt_counts_reduce_dev (word0 = E, or 0 when one of the shard's reads failed).
    In-place SUM all-reduce (u64 wrap == int64 wrap), then word0 is restored: E if
    every shard's word0 was E, else 0 (a failed read anywhere fails the interval,
    PerCpuBase.h:111-113).  Feed the result to hbt_counts_scale_dev."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    ok = sums[:, 0] == world * n_events
    sums[:, 0] = torch.where(ok, torch.full_like(sums[:, 0], n_events), torch.zeros_like(sums[:, 0]))
    return sums


def allreduce_record_counters(counters: torch.Tensor, group=None) -> torch.Tensor:
    """Per-type record counters (num_record_*_, CpuEventsGroup.h:822-832) summed
    over the CPUs of a rank -> totals over all ranks."""
    import torch.distributed as dist
    dist.all_reduce(counters, op=dist.ReduceOp.SUM, group=group)
    return counters
