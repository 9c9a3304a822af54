import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import dynolog_1_b200 as hbt
from dynolog_1_b200 import tracegen as tg
E=8; ring_bytes=16<<20
for ncpu, lost in ((8, 5000), (8, 0), (40, 5000)):
    imgs = torch.empty((ncpu, ring_bytes//8), dtype=torch.int64, device="cuda")
    tails, heads = [], []
    n = (ring_bytes - 65536)//(40+8*E+1)
    for c in range(ncpu):
        seed = tg.stream_seed(3, c)
        s = tg.group_sample_stream(seed, n, E, device="cuda", lost_every=lost)
        tail = tg.random_tail(seed, ring_bytes)
        ring, head = tg.place_in_ring(s, ring_bytes, tail, fill_seed=seed ^ 0xABCDEF)
        imgs[c] = ring; tails.append(tail); heads.append(head)
    torch.cuda.synchronize()
    for segs in ("5", "16"):
        ctx = hbt.Context(0); ctx.ring_configure(ncpu, 0); ctx.set_geometry(int(segs))
        for c in range(ncpu): ctx.submit_ring_dev(c, imgs[c], ring_bytes, tails[c], heads[c])
        ctx.decode(hbt.sampling_desc(E)); res = ctx.fetch_results()
        print("ncpu", ncpu, "lost", lost, "segs", segs, "reruns", ctx.reruns, "slow", ctx.slow_tiles, "rets", set(int(x) for x in res["ret"]))
        ctx.close()
