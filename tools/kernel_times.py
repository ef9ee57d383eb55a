"""Per-kernel steady-state times (CUDA events around back-to-back launches of ONE kernel) at
BASELINE config 2 tree size.  Tree-only ring (16-byte observations) unless --frames.
Usage: python tools/kernel_times.py [--cap 1048576] [--batch 512] [--frames]"""
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
ap.add_argument("--batch", type=int, nargs="+", default=[512])
ap.add_argument("--iters", type=int, default=300)
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


def timeit(fn, iters=args.iters):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def empty_kernel():
    torch.cuda._sleep(0)


print("cap=%d" % args.cap)
for B in args.batch:
    u = torch.rand(B, dtype=torch.float64, device=dev, generator=gen)
    idx = torch.empty(B, dtype=torch.int64, device=dev)
    w32 = torch.empty(B, device=dev)
    w64 = torch.empty(B, dtype=torch.float64, device=dev)
    q, qn, qt = (torch.randn(B, 18, device=dev, generator=gen) for _ in range(3))
    act = torch.randint(0, 18, (B,), device=dev, generator=gen)
    rew = torch.randn(B, device=dev, generator=gen)
    done = torch.zeros(B, device=dev)
    pr32 = torch.rand(B, device=dev, generator=gen) + 1e-3
    pr64 = pr32.double()
    leaf = pr64 ** 0.6
    o_prio = torch.empty(B, device=dev)
    o_loss = torch.empty(1, device=dev)
    t = {}
    t["sample (idx+w32)"] = timeit(lambda: lib.per_sample_indices(h, P(u), B, 0.4, P(idx), None, P(w32), sp))
    t["sample (idx only)"] = timeit(lambda: lib.per_sample_indices(h, P(u), B, 0.4, P(idx), None, None, sp))
    lib.per_sample_indices(h, P(u), B, 0.4, P(idx), None, P(w32), sp)
    t["update_leaves (sorted sampled idx)"] = timeit(lambda: lib.per_update_leaves(h, P(idx), P(leaf), 1.0, B, sp))
    t["update_priorities f32 (device pow)"] = timeit(lambda: lib.per_update_priorities(h, P(idx), P(pr32), pn.PRIO_F32, B, sp))
    t["td only (no tree)"] = timeit(lambda: lib.per_td_priority_update(h, P(q), P(qn), P(qt), 18, P(act), P(rew), P(done), P(w32),
                                                                      0.97, None, B, P(o_prio), None, P(o_loss), sp))
    t["td + update fused"] = timeit(lambda: lib.per_td_priority_update(h, P(q), P(qn), P(qt), 18, P(act), P(rew), P(done), P(w32),
                                                                      0.97, P(idx), B, P(o_prio), None, P(o_loss), sp))
    ridx = torch.randint(0, args.cap, (B,), device=dev, generator=gen)
    t["update_leaves (random idx)"] = timeit(lambda: lib.per_update_leaves(h, P(ridx), P(leaf), 1.0, B, sp))
    t["torch empty-ish kernel (launch floor)"] = timeit(lambda: w32.add_(0.0))
    # phase stamps of the update kernel (GPU global timer), median over 50 launches
    pn.check(lib.per_tune(h, b"update_trace", 1))
    tr = []
    for _ in range(50):
        lib.per_td_priority_update(h, P(q), P(qn), P(qt), 18, P(act), P(rew), P(done), P(w32), 0.97, P(idx), B, P(o_prio), None,
                                   P(o_loss), sp)
        st6 = (pn.C.c_int64 * 16)()
        pn.check(lib.per_update_trace(h, st6, sp))
        tr.append([st6[i + 1] - st6[i] for i in range(5)])
    pn.check(lib.per_tune(h, b"update_trace", 0))
    med = np.median(np.array(tr), axis=0) / 1e3
    print("batch %d" % B)
    print("   td+update phases (us): TD %.2f | stamps %.2f | leaves+max %.2f | sparse rounds %.2f | top %.2f | sum %.2f"
          % (med[0], med[1], med[2], med[3], med[4], med.sum()))
    for k, v in t.items():
        print("   %-40s %8.2f us" % (k, v))
st = pn.C.c_int64(0)
lib.per_status(h, pn.C.byref(st), sp)
print("status", st.value)
