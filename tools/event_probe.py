"""Probe: overhead of CUDA-event brackets around one gather launch — stream-level records between eager
launches vs event-record nodes inside a captured graph (torch external events)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ape-x_b200")]
import numpy as np, torch
import __graft_entry__ as entry
entry.build()
import memory, per_native as pn
dev = torch.device("cuda")
ring, B = 1 << 16, 512
buf = memory.CustomPrioritizedReplayBuffer(ring, 0.6)
gen = torch.Generator(device=dev).manual_seed(0)
for lo in range(0, ring, 8192):
    obs = torch.randint(0, 256, (8192, 4, 84, 84), dtype=torch.uint8, device=dev, generator=gen)
    buf.add_batch_device(obs, obs, torch.zeros(8192, dtype=torch.int64, device=dev), torch.zeros(8192, device=dev), torch.zeros(8192, device=dev),
                         priorities=torch.rand(8192, dtype=torch.float64, device=dev, generator=gen) + 1e-3)
d = buf.sample_device(B, 0.4)
idx = d["idx"]
o_obs, o_nxt = torch.empty_like(d["obs"]), torch.empty_like(d["obs"])
o_act, o_rew, o_done = torch.empty(B, dtype=torch.int64, device=dev), torch.empty(B, device=dev), torch.empty(B, device=dev)
h, P = buf._h, pn.ptr
def gather(sp):
    pn.check(pn.lib.per_gather(h, P(idx), B, P(o_obs), P(o_nxt), P(o_act), P(o_rew), P(o_done), None, sp))
s = torch.cuda.current_stream(); sp = pn.stream_ptr(s)
# 1. back to back (kernel-only incl. launch gap)
for _ in range(10): gather(sp)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(s)
for _ in range(200): gather(sp)
e1.record(s); torch.cuda.synchronize()
print("back-to-back: %.2f us per launch" % (e0.elapsed_time(e1) * 1e3 / 200))
# 2. stream-level events around each launch, other work in between
ts = []
for _ in range(100):
    buf.sample_device(B, 0.4, gather=False)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(s); gather(sp); b.record(s)
    ts.append((a, b))
torch.cuda.synchronize()
print("stream events: %.2f us median" % (np.median([a.elapsed_time(b) for a, b in ts]) * 1e3))
# 3. event-record nodes inside a graph
try:
    side = torch.cuda.Stream(); side.wait_stream(s)
    pairs, graphs = [], []
    with torch.cuda.stream(side):
        sps = pn.stream_ptr(side)
        for i in range(8):
            a, b = torch.cuda.Event(enable_timing=True, external=True), torch.cuda.Event(enable_timing=True, external=True)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                pn.check(pn.lib.per_sample_indices(h, P(torch.rand(B, dtype=torch.float64, device=dev)), B, 0.4, P(idx), None, None, sps))
                a.record(side); gather(sps); b.record(side)
                pn.check(pn.lib.per_update_priorities(h, P(idx), P(torch.rand(B, device=dev) + 1e-3), pn.PRIO_F32, B, sps))
            pairs.append((a, b)); graphs.append(g)
    s.wait_stream(side)
    for g in graphs: g.replay()
    torch.cuda.synchronize()
    for g in graphs: g.replay()
    torch.cuda.synchronize()
    print("graph event nodes: %.2f us median" % (np.median([a.elapsed_time(b) for a, b in pairs]) * 1e3), [round(a.elapsed_time(b) * 1e3, 2) for a, b in pairs])
except Exception as exc:
    print("graph event nodes failed:", str(exc)[:300])
