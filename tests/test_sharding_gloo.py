"""N>1 host logic on CPU: world_size-2 gloo.  Each rank reduces its CPU shard of a
counter-group trace (numpy stands in for hbt_counts_reduce_dev here: plain u64 column
sums), the partial sums are merged by dynolog_1_b200.sharding, and the merged sums
must equal the oracle's sums over all CPUs -- including u64 wrap-around and a failed
read in one shard."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import dynolog_1_b200 as hbt
from dynolog_1_b200 import sharding, tracegen as tg
import _oracle as orc


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, mode, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    I, C, E = 40, 10, 5
    reads = tg.counts_trace(99, I, C, E, multiplexed=True).numpy().view(np.uint64).copy()
    reads[7, :, 3] = np.uint64(2**63 + 5)   # the integer sum wraps mod 2^64
    reads[11, 6, 0] = 0                      # a failed read, lands in exactly one shard
    mine = sharding.shard_cpus(C, world, rank, mode)
    part = reads[:, mine, :]
    sums = part.sum(axis=1, dtype=np.uint64)
    sums[:, 0] = np.where((part[:, :, 0] == E).all(axis=1), E, 0)
    t = torch.from_numpy(sums.view(np.int64).copy())
    sharding.allreduce_partial_sums(t, E)
    counters = torch.tensor([rank + 1] * hbt.NUM_COUNTERS, dtype=torch.int64)
    sharding.allreduce_record_counters(counters)
    if rank == 0:
        np.save(out_path, t.numpy().view(np.uint64))
        assert counters.tolist() == [sum(range(1, world + 1))] * hbt.NUM_COUNTERS
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["block", "mod"])
def test_partial_sums_merge_world2(built, tmp_path, mode):
    out = str(tmp_path / "merged.npy")
    mp.spawn(_worker, args=(2, _free_port(), mode, out), nprocs=2, join=True)
    merged = np.load(out)
    I, C, E = 40, 10, 5
    reads = tg.counts_trace(99, I, C, E, multiplexed=True).numpy().view(np.uint64).copy()
    reads[7, :, 3] = np.uint64(2**63 + 5)
    reads[11, 6, 0] = 0
    want = orc.port_counts(reads)["sums"]
    assert np.array_equal(merged[:, 0], want[:, 0])
    ok = want[:, 0] != 0
    assert np.array_equal(merged[ok], want[ok])
    assert merged[11, 0] == 0 and merged[7, 3] == want[7, 3]


def test_shard_cpus_partitions():
    for n, w in ((256, 8), (10, 3), (5, 8), (512, 4)):
        for mode in ("block", "mod"):
            owned = [sharding.shard_cpus(n, w, r, mode) for r in range(w)]
            flat = sorted(c for o in owned for c in o)
            assert flat == list(range(n))
