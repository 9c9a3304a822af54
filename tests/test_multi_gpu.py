"""The multi-GPU merge of include/hbt_b200.h (hbt_comm_*, hbt_counts_allreduce_dev,
hbt_counts_pipeline_sharded_dev, hbt_fetch_results_all, hbt_hist_gather) over NCCL.

One rank (any GPU box): the communicator has a single member, every collective degenerates to
a local copy, and the results must equal the unsharded entry points and the oracle.
Two ranks (needs >= 2 GPUs, e.g. `gpurun --gpus 2`): each rank owns half of the CPUs; the merged
counter-group outputs must equal the oracle over ALL CPUs (scale after the integer sum), the
gathered per-CPU results and histogram the oracle's per CPU.  The host-side sharding logic is
covered on CPU by test_sharding_gloo.py."""
import os
import socket

import numpy as np
import pytest

import dynolog_1_b200 as hbt
from dynolog_1_b200 import sharding, tracegen as tg
import _oracle as orc

pytestmark = pytest.mark.gpu


def ring_np(stream, ds, tail, fill=77):
    ring, head = tg.place_in_ring(stream, ds, tail, fill_seed=fill)
    return ring.numpy().view(np.uint64), head


def counts_case(I=3000, C=32, E=6):
    reads = tg.counts_trace(4321, I, C, E, multiplexed=True).numpy().view(np.uint64).copy()
    reads[7, :, 3] = np.uint64(2**63 + 5)   # the integer sum wraps mod 2^64
    reads[11, 5, 0] = 0                      # a failed read: lands in exactly one shard
    reads[I - 1, C - 1, 0] = 0
    return reads, [(1, 0), (2, 1)]


def hist_rings(cpus, ds=1 << 18):
    out = []
    for cpu in cpus:
        s = tg.ip_sample_stream(7000 + cpu, (ds - 8192) // 40, cpu=cpu, n_pids=32, n_ips=256, lost_every=501)
        tail = tg.random_tail(cpu + 3, ds)
        ring, head = ring_np(s, ds, tail)
        out.append((ring, tail, head))
    return out


def run_rank(ctx, rank, world, dev):
    """What every rank does; returns nothing, asserts against the oracle."""
    import torch
    reads, ratios = counts_case()
    I, C, W = reads.shape
    E = W - 3
    want = orc.counts(reads, ratios)
    mine = sharding.shard_cpus(C, world, rank)
    part = torch.from_numpy(np.ascontiguousarray(reads[:, mine, :]).view(np.int64)).to(dev)
    sums = torch.zeros((I, W), dtype=torch.int64, device=dev)
    scaled = torch.zeros((I, E), dtype=torch.int64, device=dev)
    rates = torch.zeros((I, E), dtype=torch.float32, device=dev)
    ratio = torch.zeros((I, len(ratios)), dtype=torch.float64, device=dev)
    deltas = torch.zeros((I, E), dtype=torch.int64, device=dev)
    ctx.counts_pipeline_sharded_dev(part, I, len(mine), E, ratios, sums, scaled, rates, ratio, deltas)
    ctx.sync()
    ok = want["sums"][:, 0] != 0
    got_sums = sums.cpu().numpy().view(np.uint64)
    assert np.array_equal(got_sums[:, 0], want["sums"][:, 0])
    assert np.array_equal(got_sums[ok], want["sums"][ok])
    assert np.array_equal(scaled.cpu().numpy().view(np.uint64), want["scaled"])
    assert np.array_equal(deltas.cpu().numpy().view(np.uint64), want["deltas"])
    assert np.array_equal(rates.cpu().numpy(), want["rates"], equal_nan=True)
    fin = np.isfinite(want["ratios"])
    assert np.allclose(ratio.cpu().numpy()[fin], want["ratios"][fin], rtol=1e-12, atol=0.0)
    # the two-step form: reduce, all-reduce, scale
    s2 = torch.zeros_like(sums)
    ctx.counts_reduce_dev(part, I, len(mine), E, s2)
    ctx.counts_allreduce_dev(s2, I, E)
    ctx.sync()
    assert torch.equal(s2[:, 0], sums[:, 0]) and torch.equal(s2[torch.from_numpy(ok).to(dev)], sums[torch.from_numpy(ok).to(dev)])

    # decode: every rank owns n_cpus rings, results of all ranks gathered everywhere
    n_cpus, ds = 3, 1 << 18
    rings = hist_rings(range(rank * n_cpus, (rank + 1) * n_cpus), ds)
    desc = hbt.ip_desc(40)
    ctx.ring_configure(n_cpus, ds)
    ctx.set_cpu_base(rank * n_cpus)
    ctx.reset_state(hbt.INVALID_TS)
    ctx.hist_configure(n_cpus << 15)
    for c, (ring, tail, head) in enumerate(rings):
        ctx.submit_ring_image(c, ring, tail, head)
    ctx.decode(desc)
    res_all = ctx.fetch_results_all()
    assert len(res_all) == world * n_cpus
    want_hist = {}
    all_rings = hist_rings(range(world * n_cpus), ds)
    for g, (ring, tail, head) in enumerate(all_rings):
        w = orc.consume(ring, tail, head, desc, cpu=g)
        assert int(res_all[g]["ret"]) == w["ret"] and int(res_all[g]["n_records"]) == w["n_records"]
        assert int(res_all[g]["new_tail"]) == w["new_tail"]
        assert [int(x) for x in res_all[g]["counters"]] == w["counters"]
        want_hist.update(w["hist"])
    ent, counts = ctx.hist_gather(root=0, cap=len(want_hist) + 16)
    assert int(counts.sum()) == len(want_hist)
    if rank == 0:
        assert orc.hist_to_dict(ent) == want_hist


def test_single_rank_communicator(built):
    import torch
    torch.cuda.set_device(0)
    ctx = hbt.Context(0)
    try:
        ctx.comm_init_single()
        assert ctx.lib.hbt_comm_n_ranks(ctx.h) == 1
        run_rank(ctx, 0, 1, torch.device("cuda", 0))
    finally:
        ctx.close()


def test_collectives_need_a_communicator(built):
    import errno
    import torch
    ctx = hbt.Context(0)
    try:
        s = torch.zeros((4, 4), dtype=torch.int64, device="cuda")
        with pytest.raises(hbt.HbtError) as e:
            ctx.counts_allreduce_dev(s, 4, 1)
        assert e.value.code == -errno.ENOTCONN
    finally:
        ctx.close()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    ctx = hbt.Context(rank)
    try:
        ctx.comm_init(dist)
        run_rank(ctx, rank, world, dev)
    finally:
        ctx.close()
        dist.destroy_process_group()


def test_two_ranks(built):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    mp.spawn(_worker, args=(2, _free_port()), nprocs=2, join=True)
