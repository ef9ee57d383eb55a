"""Where the host-buffer step (memory.sample() + update_priorities()) spends its time: wall-clock of the two calls, of the
native calls inside them, and a plain pinned device-to-host copy of the same bytes for reference.
Usage: python tools/e2e_profile.py [--ring 262144] [--batch 512]"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ape-x_b200")]
import numpy as np
import torch

import __graft_entry__ as entry

entry.build()
import memory
import per_native as pn

ap = argparse.ArgumentParser()
ap.add_argument("--ring", type=int, default=1 << 18)
ap.add_argument("--batch", type=int, default=512)
ap.add_argument("--steps", type=int, default=100)
args = ap.parse_args()
B = args.batch
buf = memory.CustomPrioritizedReplayBuffer(args.ring, 0.6)
gen = torch.Generator(device="cuda").manual_seed(0)
for lo in range(0, args.ring, 8192):
    m = min(8192, args.ring - lo)
    obs = torch.randint(0, 256, (m, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=gen)
    buf.add_batch_device(obs, obs, torch.zeros(m, dtype=torch.int64, device="cuda"), torch.zeros(m, device="cuda"),
                         torch.zeros(m, device="cuda"), priorities=torch.rand(m, dtype=torch.float64, device="cuda", generator=gen) + 1e-3)
prios = np.abs(np.random.default_rng(0).standard_normal(B)) + 1e-6
native = {"per_sample_host": 0.0, "per_update_host": 0.0}
for name in list(native):
    fn = getattr(pn.lib, name)

    def wrapped(*a, _fn=fn, _name=name):
        t = time.perf_counter()
        rc = _fn(*a)
        native[_name] += time.perf_counter() - t
        return rc

    setattr(pn.lib, name, wrapped)
for _ in range(5):
    out = buf.sample(B, 0.4)
    buf.update_priorities(out[-1], prios)
for k in native:
    native[k] = 0.0
t_s = t_u = 0.0
for _ in range(args.steps):
    t0 = time.perf_counter()
    out = buf.sample(B, 0.4)
    t1 = time.perf_counter()
    buf.update_priorities(out[-1], prios)
    t2 = time.perf_counter()
    t_s += t1 - t0
    t_u += t2 - t1
torch.cuda.synchronize()
n = args.steps
print("per step (us): sample() %.1f (native %.1f) | update_priorities() %.1f (native %.1f) | total %.1f -> %.3f M transitions/s"
      % (t_s / n * 1e6, native["per_sample_host"] / n * 1e6, t_u / n * 1e6, native["per_update_host"] / n * 1e6,
         (t_s + t_u) / n * 1e6, B * n / (t_s + t_u) / 1e6))
# reference: the same frame bytes as two plain pinned copies
d = torch.empty(2, B, 4 * 84 * 84, dtype=torch.uint8, device="cuda")
h = torch.empty(2, B, 4 * 84 * 84, dtype=torch.uint8).pin_memory()
for _ in range(3):
    h[0].copy_(d[0], non_blocking=True)
    h[1].copy_(d[1], non_blocking=True)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(50):
    h[0].copy_(d[0], non_blocking=True)
    h[1].copy_(d[1], non_blocking=True)
    torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 50
print("two pinned D2H copies of %.1f MB: %.1f us = %.1f GB/s" % (d.numel() / 2e6, dt * 1e6, d.numel() / dt / 1e9))
