import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ape-x_b200")]
import numpy as np, torch
import __graft_entry__ as entry
entry.build()
import memory
dev = torch.device("cuda")
ring = 1 << 16
buf = memory.CustomPrioritizedReplayBuffer(ring, 0.6)
gen = torch.Generator(device=dev).manual_seed(0)
for lo in range(0, ring, 8192):
    m = 8192
    obs = torch.randint(0, 256, (m, 4, 84, 84), dtype=torch.uint8, device=dev, generator=gen)
    buf.add_batch_device(obs, obs, torch.zeros(m, dtype=torch.int64, device=dev), torch.zeros(m, device=dev), torch.zeros(m, device=dev),
                         priorities=torch.rand(m, dtype=torch.float64, device=dev, generator=gen) + 1e-3)
torch.cuda.synchronize()
pr = np.abs(np.random.default_rng(0).standard_normal(512)).astype(np.float32) + 1e-6
for i in range(5):
    out = buf.sample(512, 0.4); buf.update_priorities(out[-1], pr)
ts = []; tu = []
for i in range(50):
    t0 = time.perf_counter(); out = buf.sample(512, 0.4); t1 = time.perf_counter(); buf.update_priorities(out[-1], pr); torch.cuda.synchronize(); t2 = time.perf_counter()
    ts.append(t1 - t0); tu.append(t2 - t1)
print("sample() %.1f us  update_priorities() %.1f us" % (np.median(ts) * 1e6, np.median(tu) * 1e6))
# raw D2H of the same bytes
d = torch.empty(512 * 2 * 28224, dtype=torch.uint8, device=dev)
for nb in (1, 2, 4):
    hs = [torch.empty(d.numel() // nb, dtype=torch.uint8, pin_memory=True) for _ in range(nb)]
    streams = [torch.cuda.Stream() for _ in range(nb)]
    def go():
        for j in range(nb):
            with torch.cuda.stream(streams[j]):
                hs[j].copy_(d[j * (d.numel() // nb):(j + 1) * (d.numel() // nb)], non_blocking=True)
        torch.cuda.synchronize()
    for _ in range(3): go()
    t0 = time.perf_counter()
    for _ in range(20): go()
    dt = (time.perf_counter() - t0) / 20
    print("D2H %d B in %d concurrent copies: %.1f us = %.1f GB/s" % (d.numel(), nb, dt * 1e6, d.numel() / dt / 1e9))
import cProfile, pstats
prof = cProfile.Profile(); prof.enable()
for i in range(30):
    out = buf.sample(512, 0.4); buf.update_priorities(out[-1], pr)
prof.disable()
pstats.Stats(prof).sort_stats("cumulative").print_stats(18)
