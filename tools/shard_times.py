"""Per-rank kernel time of one sharded step WITHOUT cross-rank waiting: `world` shards are
emulated on ONE GPU (tree-only rings, 16-byte observations), every rank's record is in the
mailboxes before any rank samples, and the steps are replayed from a CUDA graph.  The difference
between these numbers and bench.py's multi-GPU step time is what the exchange and the skew
between ranks cost.
Usage: python tools/shard_times.py [--world 2 8] [--cap-local 1048576] [--batch-per-rank 512]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ape-x_b200")]

import torch

import __graft_entry__ as entry

entry.build()
import per_native as pn
import sharded

ap = argparse.ArgumentParser()
ap.add_argument("--world", type=int, nargs="+", default=[2, 8])
ap.add_argument("--cap-local", type=int, default=1 << 20)
ap.add_argument("--batch-per-rank", type=int, default=512)
ap.add_argument("--reps", type=int, default=10)
args = ap.parse_args()
dev = torch.device("cuda")
lib, P = pn.lib, pn.ptr


def graph_time(body, reps):
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        sp = pn.stream_ptr(side)
        body(sp)
        with torch.cuda.graph(g, stream=side):
            for _ in range(reps):
                body(sp)
    torch.cuda.current_stream().wait_stream(side)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (20 * reps)


for world in args.world:
    size = world * args.cap_local
    GB = world * args.batch_per_rank
    shards = [sharded.ShardedReplay(size, 0.6, r, world, obs_shape=(16,), exchange=None) for r in range(world)]
    gen = torch.Generator(device=dev).manual_seed(world)
    for lo in range(0, size, 1 << 18):
        m = min(1 << 18, size - lo)
        o = torch.zeros((m, 16), dtype=torch.uint8, device=dev)
        z = torch.zeros(m, device=dev)
        pr = torch.rand(m, dtype=torch.float64, device=dev, generator=gen) + 1e-3
        for s in shards:
            s.add_batch_device(o, o, torch.zeros(m, dtype=torch.int64, device=dev), z, z, pr)
    sharded.ShardedReplay.enable_local_exchange(shards)
    u = torch.rand(GB, dtype=torch.float64, device=dev, generator=gen)
    idx = [torch.empty(GB, dtype=torch.int64, device=dev) for _ in shards]
    w32 = torch.empty(GB, device=dev)
    q, qn, qt = (torch.randn(GB, 18, device=dev, generator=gen) for _ in range(3))
    act = torch.randint(0, 18, (GB,), device=dev, generator=gen)
    rew = torch.randn(GB, device=dev, generator=gen)
    done = torch.zeros(GB, device=dev)
    o_prio = torch.empty(GB, device=dev)
    o_loss = torch.empty(1, device=dev)

    def publish(sp):
        for s in shards:
            pn.check(lib.per_shard_publish(s._h, size, sp))

    def sample(sp):
        for r, s in enumerate(shards):
            pn.check(lib.per_sample_sharded(s._h, None, P(u), GB, 0.4, size, P(idx[r]), None, P(w32), sp))

    def update(sp):
        for r, s in enumerate(shards):
            pn.check(lib.per_td_priority_update(s._h, P(q), P(qn), P(qt), 18, P(act), P(rew), P(done), P(w32), 0.97, P(idx[r]), GB,
                                                P(o_prio), None, P(o_loss), sp))

    def td_only(sp):
        for r, s in enumerate(shards):
            pn.check(lib.per_td_priority_update(s._h, P(q), P(qn), P(qt), 18, P(act), P(rew), P(done), P(w32), 0.97, None, GB,
                                                P(o_prio), None, P(o_loss), sp))

    t_pub = graph_time(publish, args.reps) / world
    t_ps = graph_time(lambda sp: (publish(sp), sample(sp)), args.reps) / world
    t_psu = graph_time(lambda sp: (publish(sp), sample(sp), update(sp)), args.reps) / world
    t_td = graph_time(td_only, args.reps) / world
    own = int((idx[0] >= 0).sum().item())
    print("world %d, global batch %d (rank 0 owns %d strata), cap/rank 2^%d: per rank and step — publish %.2f us | sample %.2f us | "
          "TD + update %.2f us (TD arithmetic alone %.2f us)" % (world, GB, own, args.cap_local.bit_length() - 1, t_pub, t_ps - t_pub,
                                                                   t_psu - t_ps, t_td))
    for s in shards:
        assert s.status() == 0, s.status()
    del shards
    torch.cuda.empty_cache()
