"""BASELINE configs[4]: sum/min-tree kernel sweep, capacity 2^16..2^26 x batch 256..65536, tree-only ring
(16-byte observations).  Per point: sampling (indices + weights) and priority update of the sampled indices,
GPU time inside CUDA graphs (no launch path), and the algorithmic GB/s of SURVEY section 8(d):
  sample  B * (8 L + 8 + 8 + 4) bytes;   update  2 * [8 B + 24 * sum_l min(B, 2^l)] bytes.
--cpu adds the CPU column of BASELINE.md section 3 loop (ii): the oracle's list-backed port of the reference's tree
(pure Python like memory.py:76-87,106-129, one core) on the same capacity / batch, for capacities up to 2^22.
Usage: python tools/tree_sweep.py [--caps 16 20 24 26] [--batches 256 4096 65536] [--cpu]"""
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
ap.add_argument("--caps", type=int, nargs="+", default=[16, 20, 24, 26])
ap.add_argument("--batches", type=int, nargs="+", default=[256, 512, 1024, 4096, 16384, 65536])
ap.add_argument("--cpu", action="store_true")
ap.add_argument("--dense-lmax", type=int, default=None, help="per_tune dense_update_lmax (experiments)")
args = ap.parse_args()
dev = torch.device("cuda")
lib, P = pn.lib, pn.ptr


def graph_us(body, reps, inner):
    """GPU time per `body()` call: `inner` calls captured as one CUDA graph (no host launch path in the number),
    one event pair around `reps` replays."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        sps = pn.stream_ptr(side)
        body(sps)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for _ in range(inner):
                body(sps)
    torch.cuda.current_stream().wait_stream(side)
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * inner)


def cpu_point(lc, B, budget=1.5):
    """sample (indices + weights) + update of the oracle's scalar tree pair, transitions/s each."""
    import time

    import numpy as np
    from oracle import per_oracle as po
    cap = 1 << lc
    rng = np.random.default_rng(lc)
    leaves = rng.random(cap) + 1e-3
    o = po.ReplayOracle(cap, 0.6, "apex")
    o.sum_tree.v = po.build_tree(leaves, "sum").tolist()
    o.min_tree.v = po.build_tree(leaves, "min").tolist()
    o.rows = [None] * cap
    t_s = t_u = 0.0
    n = 0
    while t_s + t_u < budget:
        u = rng.random(B)
        t0 = time.perf_counter()
        idx = o.sample_indices(B, u)
        o.weights_for(idx, 0.4)
        t1 = time.perf_counter()
        o.update_priorities(idx, [float(p) for p in rng.random(B) + 1e-3])
        t2 = time.perf_counter()
        t_s, t_u, n = t_s + t1 - t0, t_u + t2 - t1, n + 1
    return n * B / t_s, n * B / t_u


print("log2C   batch | sample us   GB/s | update us   GB/s | update path" + ("        | CPU port: sample k/s  update k/s" if args.cpu else ""))
for lc in args.caps:
    cap = 1 << lc
    buf = memory.CustomPrioritizedReplayBuffer(cap, 0.6)
    gen = torch.Generator(device=dev).manual_seed(lc)
    for lo in range(0, cap, 1 << 20):
        m = min(1 << 20, cap - lo)
        o = torch.zeros((m, 16), dtype=torch.uint8, device=dev)
        z = torch.zeros(m, device=dev)
        buf.add_batch_device(o, o, torch.zeros(m, dtype=torch.int64, device=dev), z, z,
                             priorities=torch.rand(m, dtype=torch.float64, device=dev, generator=gen) + 1e-3)
    del o, z
    h, sp = buf._h, pn.stream_ptr()
    if args.dense_lmax is not None:
        pn.check(lib.per_tune(h, b"dense_update_lmax", args.dense_lmax))
    if os.environ.get("UPD_CLUSTER_MAX"):
        pn.check(lib.per_tune(h, b"update_cluster_max", int(os.environ["UPD_CLUSTER_MAX"])))
    for B in args.batches:
        u = torch.rand(B, dtype=torch.float64, device=dev, generator=gen)
        idx = torch.empty(B, dtype=torch.int64, device=dev)
        w32 = torch.empty(B, device=dev)
        pr = torch.rand(B, device=dev, generator=gen) + 1e-3
        # the update is timed as (sample + update of that sample) - (sample): the update of the sample just drawn is the
        # path that matters (the binding to a sample ends with its first update), programmatic chaining off
        pn.check(lib.per_tune(h, b"pdl", 0))

        def sample_only(sps):
            pn.check(lib.per_sample_indices(h, P(u), B, 0.4, P(idx), None, P(w32), sps))

        def sample_update(sps):
            pn.check(lib.per_sample_indices(h, P(u), B, 0.4, P(idx), None, P(w32), sps))
            pn.check(lib.per_update_priorities(h, P(idx), P(pr), pn.PRIO_F32, B, sps))

        reps, inner = (20, 20) if B <= 4096 else (10, 10)
        t_s = graph_us(sample_only, reps, inner)
        t_u = graph_us(sample_update, reps, inner) - t_s
        bytes_s = B * (8 * lc + 8 + 8 + 4)
        bytes_u = 2 * (8 * B + 24 * sum(min(B, 1 << l) for l in range(lc)))
        dense = (12 <= lc <= 26 and B > (1024 if lc <= 20 else 2048) and min(B, cap >> 10) <= 7 * 1184
                 and not os.environ.get('NO_DENSE'))  # dense_update_ok() in csrc/per_b200.cu
        if args.dense_lmax is not None and 12 <= lc <= args.dense_lmax:
            dense = True
        if os.environ.get('NO_DENSE'):
            pn.check(lib.per_tune(h, b'sorted_update', 0))
        path = "dense one-launch" if dense else ("cluster, sorted climb" if B <= 2048 else ("cluster rounds" if B <= 4096 else "multi-launch"))
        cpu = ""
        if args.cpu and lc <= 22:
            cs, cu = cpu_point(lc, B)
            cpu = " | %10.1f %10.1f" % (cs / 1e3, cu / 1e3)
        print("%5d %7d | %9.2f %6.1f | %9.2f %6.1f | %-21s%s" % (lc, B, t_s, bytes_s / t_s / 1e3, t_u, bytes_u / t_u / 1e3, path, cpu), flush=True)
    st = pn.C.c_int64(0)
    lib.per_status(h, pn.C.byref(st), sp)
    assert st.value == 0, st.value
    del buf
    torch.cuda.empty_cache()
