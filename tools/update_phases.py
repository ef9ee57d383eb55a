"""Fine-grained phase stamps of k_update_cluster (per_update_trace, GPU global timer, thread 0 of CTA 0): sample ->
fused TD + update of that sample, plain launches, median over many launches.
Usage: python tools/update_phases.py [--cap 1048576] [--batch 512 4096]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ape-x_b200")]
import numpy as np
import torch

import __graft_entry__ as entry

entry.build()
import memory
import per_native as pn

ap = argparse.ArgumentParser()
ap.add_argument("--cap", type=int, default=1 << 20)
ap.add_argument("--batch", type=int, nargs="+", default=[512, 1024, 4096])
ap.add_argument("--iters", type=int, default=200)
ap.add_argument("--dense-lmax", type=int, default=None)
args = ap.parse_args()
dev = torch.device("cuda")
buf = memory.CustomPrioritizedReplayBuffer(args.cap, 0.6)
gen = torch.Generator(device=dev).manual_seed(0)
for lo in range(0, args.cap, 1 << 16):
    m = min(1 << 16, args.cap - lo)
    o = torch.zeros((m, 16), dtype=torch.uint8, device=dev)
    buf.add_batch_device(o, o, torch.zeros(m, dtype=torch.int64, device=dev), torch.zeros(m, device=dev),
                         torch.zeros(m, device=dev), priorities=torch.rand(m, dtype=torch.float64, device=dev, generator=gen) + 1e-3)
h, lib, P = buf._h, pn.lib, pn.ptr
sp = pn.stream_ptr()
pn.check(lib.per_tune(h, b"update_trace", 1))
NAMES = ["wait passed", "TD done", "arbitration", "leaves/barrier", "rounds done", "top rebuilt", "TD staged", "CTA max", "TD cl.sync",
         "batch max", "leaves+sib in", "climb done", "top loaded", "top written", "-", "-"]
ORDER = [0, 6, 7, 8, 9, 1, 10, 2, 11, 3, 4, 12, 5, 13]
if args.dense_lmax is not None:
    pn.check(lib.per_tune(h, b"dense_update_lmax", args.dense_lmax))
LOG2 = (args.cap - 1).bit_length()
# the update as TD launch + k_upd_dense_sorted (per_update_trace in include/per_b200.h)
NAMES_DENSE = ["wait passed", "priorities ready (dense)", "-", "fenced (CTA 0)", "-", "top lifted", "TD staged", "CTA max", "TD cl.sync", "batch max",
               "leaves written (CTA 0)", "chunk lifted (CTA 0)", "last CTA elected", "end", "runs found (CTA 0)", "-"]
ORDER_DENSE = [0, 6, 7, 8, 9, 14, 1, 10, 11, 3, 12, 5, 13]


def dense_form(B):  # dense_update_ok() in csrc/per_b200.cu
    if not 12 <= LOG2 <= 26:
        return False
    if args.dense_lmax is not None and LOG2 <= args.dense_lmax:
        return True
    return B > (1024 if LOG2 <= 20 else 2048) and min(B, args.cap >> 10) <= 7 * 1184


for B in args.batch:
    u = [torch.rand(B, dtype=torch.float64, device=dev, generator=gen) for _ in range(8)]
    idx = torch.empty(B, dtype=torch.int64, device=dev)
    w32 = torch.empty(B, device=dev)
    q, qn, qt = (torch.randn(B, 18, device=dev, generator=gen) for _ in range(3))
    act = torch.randint(0, 18, (B,), device=dev, generator=gen)
    rew, done = torch.randn(B, device=dev, generator=gen), torch.zeros(B, device=dev)
    o_prio, o_loss = torch.empty(B, device=dev), torch.empty(1, device=dev)
    rows, clk = [], []
    for it in range(args.iters):
        pn.check(lib.per_sample_indices(h, P(u[it % 8]), B, 0.4, P(idx), None, P(w32), sp))
        pn.check(lib.per_td_priority_update(h, P(q), P(qn), P(qt), 18, P(act), P(rew), P(done), P(w32), 0.97, P(idx), B, P(o_prio),
                                            None, P(o_loss), sp))
        st = (pn.C.c_int64 * 16)()
        pn.check(lib.per_update_trace(h, st, sp))
        rows.append([st[i] - st[0] for i in range(16)])
    med = np.median(np.array(rows[20:]), axis=0) / 1e3
    names, order = (NAMES_DENSE, ORDER_DENSE) if dense_form(B) else (NAMES, ORDER)
    print("batch %d%s: " % (B, " (TD launch + dense form)" if dense_form(B) else "") +
          " | ".join("%s %.2f" % (names[i], med[i]) for i in order if med[i] > 0 or i == 0))
