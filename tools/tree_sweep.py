"""BASELINE configs[4]: sum/min-tree kernel sweep, capacity 2^16..2^26 x batch 256..65536, tree-only ring
(16-byte observations).  Per point: sampling (indices + weights) and priority update of the sampled indices,
CUDA-event time over back-to-back launches, and the algorithmic GB/s of SURVEY section 8(d):
  sample  B * (8 L + 8 + 8 + 4) bytes;   update  2 * [8 B + 24 * sum_l min(B, 2^l)] bytes.
Small batches are bound by the launch path of this script (~4-5 us per call), not by the kernels.
Usage: python tools/tree_sweep.py [--caps 16 20 24 26] [--batches 256 4096 65536]"""
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
ap.add_argument("--batches", type=int, nargs="+", default=[256, 512, 4096, 65536])
args = ap.parse_args()
dev = torch.device("cuda")
lib, P = pn.lib, pn.ptr


def timeit(fn, iters):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


print("log2C   batch | sample us   GB/s | update us   GB/s | update path")
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
    if os.environ.get("UPD_CLUSTER_MAX"):
        pn.check(lib.per_tune(h, b"update_cluster_max", int(os.environ["UPD_CLUSTER_MAX"])))
    for B in args.batches:
        u = torch.rand(B, dtype=torch.float64, device=dev, generator=gen)
        idx = torch.empty(B, dtype=torch.int64, device=dev)
        w32 = torch.empty(B, device=dev)
        pr = torch.rand(B, device=dev, generator=gen) + 1e-3
        iters = 200 if B <= 4096 else 50
        t_s = timeit(lambda: lib.per_sample_indices(h, P(u), B, 0.4, P(idx), None, P(w32), sp), iters)
        lib.per_sample_indices(h, P(u), B, 0.4, P(idx), None, P(w32), sp)
        t_u = timeit(lambda: lib.per_update_priorities(h, P(idx), P(pr), pn.PRIO_F32, B, sp), iters)
        bytes_s = B * (8 * lc + 8 + 8 + 4)
        bytes_u = 2 * (8 * B + 24 * sum(min(B, 1 << l) for l in range(lc)))
        path = "cluster, sorted" if B <= 1024 else ("cluster" if B <= 4096 else "multi-launch")
        print("%5d %7d | %9.2f %6.1f | %9.2f %6.1f | %s" % (lc, B, t_s, bytes_s / t_s / 1e3, t_u, bytes_u / t_u / 1e3, path), flush=True)
    st = pn.C.c_int64(0)
    lib.per_status(h, pn.C.byref(st), sp)
    assert st.value == 0, st.value
    del buf
    torch.cuda.empty_cache()
