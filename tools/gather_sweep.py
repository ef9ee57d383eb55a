"""Sweeps the launch shape of the frame-gather kernel (per_tune) on a B200 and prints
microseconds per launch and GB/s (algorithmic bytes).  Kernel-only timing: a CUDA
graph of M launches, each with its own index set (so no launch re-reads rows another one left in
L2; the ring is much larger than L2), one CUDA-event pair around the replay.  Every shape's output is
checked byte for byte against torch indexing of a kept copy of the ring.
Usage: python tools/gather_sweep.py [--ring 262144] [--batch 512] [--quick]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ape-x_b200")]

import torch

import __graft_entry__ as entry

entry.build()
import memory
import per_native as pn

ap = argparse.ArgumentParser()
ap.add_argument("--ring", type=int, default=1 << 18)
ap.add_argument("--batch", type=int, default=512)
ap.add_argument("--launches", type=int, default=32, help="gather launches per graph (each with its own indices)")
ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--quick", action="store_true")
args = ap.parse_args()

OBS = (4, 84, 84)
B = args.batch
buf = memory.PrioritizedReplayBuffer(args.ring, 0.6)
gen = torch.Generator(device="cuda").manual_seed(0)
keep_obs, keep_nxt = [], []
for lo in range(0, args.ring, 8192):
    m = min(8192, args.ring - lo)
    obs = torch.randint(0, 256, (m,) + OBS, dtype=torch.uint8, device="cuda", generator=gen)
    nxt = torch.randint(0, 256, (m,) + OBS, dtype=torch.uint8, device="cuda", generator=gen)
    buf.add_batch_device(obs, nxt, torch.arange(lo, lo + m, dtype=torch.int64, device="cuda"),
                         torch.zeros(m, device="cuda"), torch.zeros(m, device="cuda"))
    keep_obs.append(obs)
    keep_nxt.append(nxt)
ring_obs, ring_nxt = torch.cat(keep_obs), torch.cat(keep_nxt)
del keep_obs, keep_nxt
M = args.launches
idxs = [torch.randint(0, args.ring, (B,), dtype=torch.int64, device="cuda", generator=gen).sort().values for _ in range(M)]
o_obs = torch.empty((B,) + OBS, dtype=torch.uint8, device="cuda")
o_nxt = torch.empty_like(o_obs)
o_act = torch.empty(B, dtype=torch.int64, device="cuda")
o_rew = torch.empty(B, device="cuda")
o_done = torch.empty(B, device="cuda")
h = buf._h
bytes_per_launch = 2 * (2 * 28224 + 16) * B
pn.check(pn.lib.per_tune(h, b"pdl", 0))  # launches of one graph must not overlap: kernel-only time


def tune(**kw):
    # shrink first, then grow, so every intermediate shape is valid
    pn.check(pn.lib.per_tune(h, b"gather_mode", 1))
    pn.check(pn.lib.per_tune(h, b"gather_stages", 2))
    pn.check(pn.lib.per_tune(h, b"gather_warps", 1))
    pn.check(pn.lib.per_tune(h, b"gather_piece", 16))
    pn.check(pn.lib.per_tune(h, b"gather_ctas_per_sm", kw.get("gather_ctas_per_sm", 2)))
    pn.check(pn.lib.per_tune(h, b"gather_piece", kw.get("gather_piece", 7056)))
    pn.check(pn.lib.per_tune(h, b"gather_warps", kw.get("gather_warps", 4)))
    pn.check(pn.lib.per_tune(h, b"gather_stages", kw.get("gather_stages", 2)))
    for k in ("gather_depth", "gather_pipe", "gather_vbytes"):
        if k in kw:
            pn.check(pn.lib.per_tune(h, k.encode(), kw[k]))
    pn.check(pn.lib.per_tune(h, b"gather_mode", kw.get("gather_mode", 1)))


def launch(i, sp):
    rc = pn.lib.per_gather(h, pn.ptr(idxs[i % M]), B, pn.ptr(o_obs), pn.ptr(o_nxt), pn.ptr(o_act), pn.ptr(o_rew),
                           pn.ptr(o_done), None, sp)
    if rc:
        pn.check(rc)


def measure():
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        launch(0, pn.stream_ptr(side))
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for i in range(M):
                launch(i, pn.stream_ptr(side))
    torch.cuda.current_stream().wait_stream(side)
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(args.reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / M)
    return best


def correct():
    o_obs.zero_()
    o_nxt.zero_()
    launch(3, pn.stream_ptr())
    torch.cuda.synchronize()
    k = idxs[3]
    return bool(torch.equal(o_obs, ring_obs[k]) and torch.equal(o_nxt, ring_nxt[k]) and torch.equal(o_act, k))


shapes = []  # (label, kwargs)
shapes.append(("ldg16 piece7056 2/sm", dict(gather_mode=1, gather_ctas_per_sm=2)))
shapes.append(("tma 4w2s 4/sm", dict(gather_mode=0, gather_warps=4, gather_stages=2, gather_piece=7056, gather_ctas_per_sm=4)))
for vb, depths in ((16, (4, 8, 12, 16)), (32, (2, 4, 6, 8))):
    for d in depths:
        for pipe in (0, 1):
            for ctas in ((2, 4) if args.quick else (1, 2, 3, 4, 6, 8)):
                shapes.append(("bal vb%d d%d pipe%d %d/sm" % (vb, d, pipe, ctas),
                               dict(gather_mode=2, gather_vbytes=vb, gather_depth=d, gather_pipe=pipe, gather_ctas_per_sm=ctas)))
print("%-28s |   us/launch     GB/s   frac_of_6555.8  ok" % "shape")
best = None
for label, kw in shapes:
    try:
        tune(**kw)
    except Exception as e:
        print(label, "invalid:", str(e)[:60])
        continue
    us = measure()
    ok = correct()
    gbs = bytes_per_launch / us / 1e3
    print("%-28s | %10.2f %9.1f %8.3f   %s" % (label, us, gbs, gbs / 6555.8, ok), flush=True)
    if ok and (best is None or us < best[0]):
        best = (us, label)
print("best:", best)
