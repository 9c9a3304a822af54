#!/usr/bin/env python
"""PCIe copy rates on the GPU box: H2D alone, D2H alone, both at once (pinned memory)."""
import time
import torch

dev = torch.device("cuda", 0)
h_in = torch.empty(4 << 30, dtype=torch.uint8).pin_memory()
h_out = torch.empty(1 << 30, dtype=torch.uint8).pin_memory()
d_in = torch.empty(4 << 30, dtype=torch.uint8, device=dev)
d_out = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)


def t(fn, n=3):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n


def h2d(chunks=1):
    with torch.cuda.stream(s1):
        n = h_in.numel() // chunks
        for i in range(chunks):
            d_in[i * n:(i + 1) * n].copy_(h_in[i * n:(i + 1) * n], non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


a = t(h2d); print(f"H2D 4 GiB: {a*1e3:.1f} ms  {4.295/a:.1f} GB/s")
a = t(lambda: h2d(256)); print(f"H2D 4 GiB in 256 copies: {a*1e3:.1f} ms  {4.295/a:.1f} GB/s")
a = t(d2h); print(f"D2H 1 GiB: {a*1e3:.1f} ms  {1.074/a:.1f} GB/s")
a = t(lambda: (h2d(), d2h())); print(f"H2D 4 GiB || D2H 1 GiB: {a*1e3:.1f} ms")


def grouped(G, d2h_stream="own", sync_copy=False):
    ss = [torch.cuda.Stream(dev) for _ in range(G)]
    sd = torch.cuda.Stream(dev)
    n, m = h_in.numel() // G, h_out.numel() // G

    def fn():
        for g in range(G):
            with torch.cuda.stream(ss[g]):
                d_in[g * n:(g + 1) * n].copy_(h_in[g * n:(g + 1) * n], non_blocking=True)
        for g in range(G):
            ss[g].synchronize()
            st = ss[g] if d2h_stream == "own" else sd
            with torch.cuda.stream(st):
                h_out[g * m:(g + 1) * m].copy_(d_out[g * m:(g + 1) * m], non_blocking=not sync_copy)
    return fn


for G in (4, 8):
    for mode in ("own", "shared"):
        a = t(grouped(G, mode)); print(f"G={G} D2H on {mode} stream: {a*1e3:.1f} ms")
    a = t(grouped(G, "own", True)); print(f"G={G} blocking D2H: {a*1e3:.1f} ms")


def chained(G, tiny_d2h=True, copies=64):
    ss = [torch.cuda.Stream(dev) for _ in range(G)]
    small_d = torch.zeros(4, dtype=torch.int32, device=dev)
    small_h = torch.zeros(4, dtype=torch.int32).pin_memory()
    n = h_in.numel() // G
    k = n // copies
    out = []

    def fn():
        ev = None
        for g in range(G):
            with torch.cuda.stream(ss[g]):
                if ev is not None:
                    ss[g].wait_event(ev)
                for i in range(copies):
                    a = g * n + i * k
                    d_in[a:a + k].copy_(h_in[a:a + k], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(ss[g])
                small_d.add_(1)
                if tiny_d2h:
                    small_h.copy_(small_d, non_blocking=True)
        t0 = time.perf_counter()
        tr = []
        for g in range(G):
            ss[g].synchronize()
            tr.append((time.perf_counter() - t0) * 1e3)
        out.append(tr)
    return fn, out


for tiny in (True, False):
    fn, out = chained(4, tiny)
    a = t(fn, 2)
    print(f"chained G=4 tiny_d2h={tiny}: {a*1e3:.1f} ms; group completion:", " ".join(f"{x:.1f}" for x in out[-1]))
