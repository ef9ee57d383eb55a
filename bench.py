#!/usr/bin/env python
"""bench.py — prioritized-replay hot path on B200: sample (prefix-sum descent + importance
weights + frame gather) -> TD-error -> priority -> tree update, per batch.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Prints ONE JSON line (rank 0).  `value` = transitions sampled+updated per second over all
ranks with every input resident in HBM; `e2e` = the same loop through the reference-facing
Python API (memory.py: sample() -> host numpy, update_priorities(host arrays)), host<->device
copies inside the timed region; `roofline` = the frame-gather kernel (the HBM-bound one): one
CUDA-event pair around a graph of 16 of its launches, each on its own freshly sampled index set,
run right after the timed region (an event between two kernels of the step would break their
chaining, so the timed region itself carries none); `roofline_tree`, `aql`, `learner_updates_per_s`:
the other measurement rows of SURVEY.md section 8(d);
`cpu_baseline` = the oracle port of the reference's Python path on this box's host cores.

Workload (BASELINE.json configs[1], SURVEY.md §8d "C2"): ring of 2^20 Atari transitions
(uint8[4,84,84] obs + next_obs, int64 action, f32 reward/done = 59.2 GB in HBM), batch 512,
alpha 0.6, beta 0.4, n-step 3, 18 actions, synthetic data (torch.Generator seed 0).
The ring (59 GB) is far larger than L2 (126 MB), so sampled rows come from HBM; the two
trees (32 MiB) are L2-resident by nature.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "ape-x_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

RING = 1 << 20
BATCH = 512
OBS_SHAPE = (4, 84, 84)
OBS_BYTES = 4 * 84 * 84
N_ACTIONS = 18
ALPHA, BETA, N_STEP, GAMMA = 0.6, 0.4, 3, 0.99
DEFAULT_GATHER_MODE = 2  # per_tune "gather_mode" the library ships with (csrc/gather_kernels.cuh GatherTune)
METRIC = "PER sample+update transitions/s"
UNIT = "transitions/s"
# per sampled transition the gather reads 2*28224 B of frames + 16 B of
# scalars and writes the same.
GATHER_BYTES_PER_ROW = 2 * (2 * OBS_BYTES + 16)


def ncu_gather_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one gather launch at this workload, from the newest
    profiles/r*_gather_traffic.json (written by tools/summarize_ncu.py from an `ncu --set full` capture of
    this very command); None when no capture of the current kernel is on file."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_gather_traffic.json")))
    if not files:
        return None, None
    d = json.load(open(files[-1]))
    return d.get("dram_bytes_per_launch"), d.get("kernel")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index, period=0.02):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(self.period)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------------ CPU baseline
def build_cpu_port(ring, seed=0):
    """The oracle's list-backed port of the reference buffer, pre-filled to `ring` rows.  The
    trees are filled through the vectorised builder (same node values as `ring` sequential
    adds, minutes faster); rows reference a pool of 256 distinct frame pairs, as the
    reference stores rows by reference (memory.py:163-168)."""
    from oracle import per_oracle as po
    rng = np.random.default_rng(seed)
    buf = po.ReplayOracle(ring, ALPHA, "apex")
    buf.lists_like_reference = True  # CustomPrioritizedReplayBuffer._encode_sample (memory.py:348-362): lists of references, no frame copy
    pool = rng.integers(0, 256, size=(256,) + OBS_SHAPE, dtype=np.uint8)
    prio = np.abs(rng.standard_normal(ring)) + 1e-6
    leaves = prio ** ALPHA
    cap = po.next_pow2(ring)
    ls, lm = np.zeros(cap), np.full(cap, np.inf)
    ls[:ring] = lm[:ring] = leaves
    buf.sum_tree.v = po.build_tree(ls, "sum").tolist()
    buf.min_tree.v = po.build_tree(lm, "min").tolist()
    buf.rows = [(pool[i & 255], i % N_ACTIONS, float(i % 3 - 1), pool[(i + 1) & 255], 0.0) for i in range(ring)]
    buf.cursor = 0
    buf.max_priority = float(max(1.0, prio.max()))
    return buf


def cpu_step(buf, rng, batch=None):
    """One reference learner-side replay step: sample(B, beta) then update_priorities."""
    batch = batch or BATCH
    u = rng.random(batch)
    out = buf.sample(batch, BETA, u)
    pr = np.abs(rng.standard_normal(batch)) + 1e-6
    buf.update_priorities(out[-1], [float(p) for p in pr])


def run_cpu_baseline(steps, warmup, budget_s=None, n_gpus=1):
    """The reference's single tree and single process hold the WHOLE replay of the n_gpus configuration:
    n_gpus x 2^20 transitions, global batch n_gpus x 512."""
    buf = build_cpu_port(RING * n_gpus)
    batch = BATCH * n_gpus
    rng = np.random.default_rng(1)
    for _ in range(warmup):
        cpu_step(buf, rng, batch)
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        cpu_step(buf, rng, batch)
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return done * batch / dt, done, dt


def reference_arm(args, rank):
    """--impl reference: the reference's own CPU algorithm (oracle port: pure Python over
    Python lists, like memory.py — it is serial, so 1 core) on the same config and metric."""
    if rank != 0:
        return
    n = max(1, args.gpus)
    steps = min(args.steps, 400)  # ~10-70 ms per step; keeps the run within minutes
    value, done, dt = run_cpu_baseline(steps, min(args.warmup, 5), budget_s=120.0, n_gpus=n)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus,
        "steps": done, "warmup": min(args.warmup, 5), "ms_per_step": 1e3 * dt / max(done, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(n),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "port",
                         "sample": "%d steps of sample(%d)+update_priorities(%d) on one 2^%d-leaf tree pair (the reference "
                                   "keeps the whole replay in one process), oracle/per_oracle.py ReplayOracle with "
                                   "CustomPrioritizedReplayBuffer semantics (pure Python like memory.py; observations returned "
                                   "as lists of references, memory.py:348-362), host has %d cores"
                                   % (done, BATCH * n, BATCH * n, int(np.log2(RING * n)), os.cpu_count())},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_config(n_gpus, ring=None):
    ring = ring or RING
    return {"workload": "Ape-X learner replay step, configs[%d]: ring 2^%d Atari transitions per GPU (%.1f GB HBM), "
                        "batch %d per GPU, alpha 0.6, beta 0.4, n-step 3, 18 actions"
                        % (3 if (n_gpus == 8 and ring == 1 << 21) else 1, int(np.log2(ring)), ring * 56464 / 1e9, BATCH),
            "ring_per_gpu": ring, "batch_per_gpu": BATCH, "obs": "uint8[4,84,84]",
            "sharding": "none" if n_gpus == 1 else "one replay of %d x ring_per_gpu transitions cut into %d contiguous slices; "
                        "per step one exchange of a 32-byte record per rank (peer-memory mailboxes written by the sampling kernel over "
                        "NVLink, or --nccl-exchange), global stratified batch %d" % (n_gpus, n_gpus, BATCH * n_gpus),
            "l2": "inputs (59 GB ring) larger than L2; no flush", "parallelism": "shard%d" % n_gpus}


# ------------------------------------------------------------- extra rows of SURVEY 8(d), N = 1
def _event_us(torch, fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def aql_row(dev):
    """BASELINE configs[2] (C3): AQL candidate-Q + argmax at B=256, K=1000, one continuous action dim —
    the fused tcgen05 kernel (aql.CandidateQ) against the stock PyTorch float32 ops the reference runs for
    the same arithmetic (model.py:294-335: action MLP, repeat + concat, 128->64->1, argmax) on this GPU."""
    import torch
    import torch.nn.functional as F

    import aql
    B, K, A, OBS = 256, 1000, 1, 24
    g = torch.Generator(device=dev).manual_seed(7)
    r = lambda *shape: torch.randn(*shape, device=dev, generator=g)
    wa1, ba1, wa2, ba2 = r(128, A) * 0.5, r(128) * 0.1, r(64, 128) * 0.09, r(64) * 0.1
    wadv1, badv1, wadv2, badv2 = r(64, 128) * 0.09, r(64) * 0.1, r(1, 64) * 0.12, r(1) * 0.1
    f = torch.relu(r(B, 64))
    a_mu = torch.rand(B, K, A, device=dev, generator=g) * 2 - 1
    was_tf32 = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False

    def stock():
        a_out = F.relu(F.linear(F.relu(F.linear(a_mu.reshape(-1, A), wa1, ba1)), wa2, ba2)).reshape(-1, K, 64)
        q_f = f.repeat(1, K).reshape(-1, K, 64)
        x = F.relu(torch.cat([a_out, q_f], dim=2)).reshape(-1, 128)
        q = F.linear(F.relu(F.linear(x, wadv1, badv1)), wadv2, badv2).reshape(B, K)
        return q, q.argmax(1)

    cq = aql.CandidateQ(True, A)
    cq.set_weights(wa1, ba1, wa2, ba2, wadv1, badv1, wadv2, badv2)
    with torch.no_grad():
        q_ref, am_ref = stock()
        q_got, am_got = cq.forward(f, a_mu)
        err = float((q_got - q_ref).abs().max()) / max(float(q_ref.abs().max()), 1.0)
        t_stock = _event_us(torch, stock, 20)
        t_ours = _event_us(torch, lambda: cq.forward(f, a_mu), 50)
    torch.backends.cuda.matmul.allow_tf32 = was_tf32
    flop = 2 * B * K * (A * 128 + 128 * 64 + 64 * 64 + 64) + 2 * B * 64 * 64
    pipe = None
    try:
        import glob
        files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_aql_tensor_pipe.json")))
        if files:
            pipe = json.load(open(files[-1]))
    except Exception:
        pass
    return {"workload": "AQL candidate-Q + argmax, B=256, K=1000, A=1 (BASELINE configs[2])", "fused_us": t_ours,
            "stock_pytorch_fp32_us": t_stock, "speedup": t_stock / t_ours, "gflop_per_forward": flop / 1e9,
            "tflops_fp32_equivalent": flop / t_ours / 1e6, "max_rel_err_vs_stock": err,
            "argmax_agree": float((am_got == am_ref).float().mean()),
            "tensor_pipe_from_ncu": pipe, "timing": "CUDA events around 50 forwards (3 chained launches each), plain launches"}


def learner_row(buf, dev):
    """BASELINE's second metric, Ape-X learner updates/s (ApeX.py:43-82) at configs[1]: the reference's dueling
    Nature-CNN DQN as plain PyTorch fp32 (out of scope), batch 512, n-step 3, double-Q, RMSprop,
    fed by this replay without leaving the GPU: sample_device (gather fused with the uint8 -> float32 widening) ->
    3 forwards + backward + optimiser step -> td_priority_update."""
    import torch
    import torch.nn as nn

    class Dueling(nn.Module):  # model.py:14-66
        def __init__(self):
            super().__init__()
            self.features = nn.Sequential(nn.Conv2d(4, 32, 8, 4), nn.ReLU(), nn.Conv2d(32, 64, 4, 2), nn.ReLU(),
                                          nn.Conv2d(64, 64, 3, 1), nn.ReLU(), nn.Flatten())
            self.advantage = nn.Sequential(nn.Linear(3136, 128), nn.ReLU(), nn.Linear(128, N_ACTIONS))
            self.value = nn.Sequential(nn.Linear(3136, 128), nn.ReLU(), nn.Linear(128, 1))

        def forward(self, x):
            f = self.features(x)
            a = self.advantage(f)
            return self.value(f) + a - a.mean(1, keepdim=True)

    torch.manual_seed(0)
    model, tgt = Dueling().to(dev), Dueling().to(dev)
    tgt.load_state_dict(model.state_dict())
    opt = torch.optim.RMSprop(model.parameters(), 6.25e-5, alpha=0.95, eps=1.5e-7, centered=True)
    B = 512

    def net(d):
        q, qn = model(d["obs"]), model(d["next_obs"])
        with torch.no_grad():
            qt = tgt(d["next_obs"])
        q_a = q.gather(1, d["act"].unsqueeze(1)).squeeze(1)
        nxt = qt.gather(1, qn.max(1)[1].unsqueeze(1)).squeeze(1)
        td = torch.abs((d["rew"] + (GAMMA ** N_STEP) * nxt * (1 - d["done"])).detach() - q_a)
        loss = (torch.where(td < 1, 0.5 * td ** 2, td - 0.5) * d["weights"]).mean()
        opt.zero_grad()
        loss.backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 40)
        opt.step()
        return q.detach(), qn.detach(), qt

    def step():
        d = buf.sample_device(B, BETA, obs_dtype=torch.float32)
        q, qn, qt = net(d)
        buf.td_priority_update(q, qn, qt, d["act"], d["rew"], d["done"], d["weights"], N_STEP, GAMMA, idx=d["idx"], want_loss=False)

    def wall(fn, n):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / n

    t_step = wall(step, 20)
    d0 = buf.sample_device(B, BETA, obs_dtype=torch.float32)
    t_net = wall(lambda: net(d0), 20)
    return {"value": 1.0 / t_step, "unit": "learner updates/s", "ms_per_update": t_step * 1e3,
            "network_alone_ms": t_net * 1e3, "replay_share": max(0.0, 1.0 - t_net / t_step),
            "workload": "ApeX.py:43-82 at configs[1]: dueling Nature-CNN DQN (PyTorch fp32), batch 512, n-step 3, "
                        "double-Q, RMSprop; replay on the device path (sample_device + td_priority_update)",
            "timing": "wall clock over 20 learner steps, synchronised"}


# ------------------------------------------------------------------------------- GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ring", type=int, default=RING, help="rows per GPU (debug: smaller ring)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--parity-size", type=int, default=40000,
                    help="world > 1: transitions of the small replay the start-up parity check runs on (configs[3] as written: 16777216)")
    ap.add_argument("--no-parity-check", action="store_true", help="world > 1: skip the start-up parity check against the oracle")
    ap.add_argument("--no-extras", action="store_true", help="skip the AQL and whole-learner-step rows (N=1)")
    ap.add_argument("--global-batch", type=int, default=0,
                    help="strong scaling: fix the GLOBAL stratified batch (e.g. 4096) and give every GPU global/N strata "
                         "(default: 512 strata per GPU, weak scaling, the contract's configuration)")
    ap.add_argument("--timeline", action="store_true",
                    help="diagnostic: after the timed region replay two graph blocks with the kernels' GPU-timer stamps on "
                         "(per_timeline) and print each rank's median step anatomy to stderr")
    ap.add_argument("--no-early-publish", action="store_true",
                    help="world > 1: the sampling kernel publishes its own record instead of the preceding update kernel")
    ap.add_argument("--eager", action="store_true", help="time plain launches instead of CUDA graphs")
    ap.add_argument("--trace", action="store_true", help="also print per-call GPU time of sample / gather / update (eager, events)")
    ap.add_argument("--nccl-exchange", action="store_true",
                    help="world > 1: exchange the shard records with an NCCL all-gather (eager launches) instead of peer-memory mailboxes")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        reference_arm(args, rank)
        return 0

    import torch
    import torch.distributed as dist

    import __graft_entry__ as entry
    entry.build()
    import memory
    import per_native as pn

    assert torch.cuda.is_available(), "bench.py needs a B200; the product has no CPU path"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    ring = args.ring
    W = max(args.warmup, 3)
    K = args.steps
    global BATCH
    if args.global_batch:
        assert args.global_batch % world == 0
        BATCH = args.global_batch // world
    GB = BATCH * world  # global batch: 512 strata per GPU

    # ---- build + fill this rank's slice ----------------------------------------------------
    # world == 1: the drop-in buffer.  world > 1: slice `rank` of one replay of world*ring
    # transitions (sharded.py); every rank fills the slots it owns.
    if world == 1:
        buf = memory.CustomPrioritizedReplayBuffer(ring, ALPHA)
        shard = None
    else:
        import sharded
        shard = sharded.ShardedReplay(world * ring, ALPHA, rank, world, obs_shape=OBS_SHAPE)
        buf = None
    gen = torch.Generator(device=dev).manual_seed(rank)
    chunk = 8192
    for lo in range(0, ring, chunk):
        m = min(chunk, ring - lo)
        obs = torch.randint(0, 256, (m,) + OBS_SHAPE, dtype=torch.uint8, device=dev, generator=gen)
        nxt = torch.randint(0, 256, (m,) + OBS_SHAPE, dtype=torch.uint8, device=dev, generator=gen)
        act = torch.randint(0, N_ACTIONS, (m,), dtype=torch.int64, device=dev, generator=gen)
        rew = torch.randint(-1, 2, (m,), device=dev, generator=gen).float()
        done = (torch.rand(m, device=dev, generator=gen) < 1e-3).float()
        pr = torch.randn(m, dtype=torch.float64, device=dev, generator=gen).abs() + 1e-6
        if world == 1:
            buf.add_batch_device(obs, nxt, act, rew, done, priorities=pr)
        else:
            pn.check(pn.lib.per_add_batch(shard._h, pn.ptr(obs), pn.ptr(nxt), pn.ptr(act), pn.ptr(rew), pn.ptr(done), None,
                                          pn.ptr(pr), None, 0.0, m, pn.stream_ptr()))
    del obs, nxt
    if shard is not None:
        shard.set_global_len(world * ring, 0)
    if os.environ.get("PER_GATHER_MODE") is not None:  # experiments: 0 TMA rings, 1 one piece per warp, 2 balanced shares
        hh = buf._h if world == 1 else shard._h
        pn.check(pn.lib.per_tune(hh, b"gather_mode", int(os.environ["PER_GATHER_MODE"])))
        for env, key in (("PER_GATHER_CTAS", b"gather_ctas_per_sm"), ("PER_GATHER_PIECE", b"gather_piece"),
                         ("PER_GATHER_DEPTH", b"gather_depth"), ("PER_GATHER_PIPE", b"gather_pipe"),
                         ("PER_GATHER_VBYTES", b"gather_vbytes")):
            if os.environ.get(env):
                pn.check(pn.lib.per_tune(hh, key, int(os.environ[env])))
    if os.environ.get("PER_UPDATE_CTAS") is not None:  # A/B: cap on the update kernel's cluster size (8 = the portable limit)
        pn.check(pn.lib.per_tune(buf._h if world == 1 else shard._h, b"update_ctas", int(os.environ["PER_UPDATE_CTAS"])))
    if os.environ.get("PER_SORTED_UPDATE") is not None:
        pn.check(pn.lib.per_tune(buf._h if world == 1 else shard._h, b"sorted_update", int(os.environ["PER_SORTED_UPDATE"])))
    if os.environ.get("PER_PDL") is not None:  # A/B switch for tools/pdl_ab.sh (default: on)
        pn.check(pn.lib.per_tune(buf._h if world == 1 else shard._h, b"pdl", int(os.environ["PER_PDL"])))
    use_mailbox = shard is not None and not args.nccl_exchange
    if use_mailbox:
        shard.enable_peer_exchange()  # CUDA-IPC mailboxes: the sampling kernel exchanges the records itself
    glen = world * ring
    torch.cuda.synchronize()

    # ---- resident inputs / outputs (same on every rank: common seed) -------------------------
    cgen = torch.Generator(device=dev).manual_seed(1234)
    total_steps = W + K
    n_u = min(total_steps, 1024)  # distinct uniform vectors, cycled
    u_all = torch.rand(n_u, GB, dtype=torch.float64, device=dev, generator=cgen)
    q = torch.randn(GB, N_ACTIONS, device=dev, generator=cgen)
    qn = torch.randn(GB, N_ACTIONS, device=dev, generator=cgen)
    qt = torch.randn(GB, N_ACTIONS, device=dev, generator=cgen)
    idx = torch.empty(GB, dtype=torch.int64, device=dev)
    w32 = torch.empty(GB, dtype=torch.float32, device=dev)
    o_obs = torch.zeros((GB,) + OBS_SHAPE, dtype=torch.uint8, device=dev)
    o_nxt = torch.zeros_like(o_obs)
    o_act = torch.zeros(GB, dtype=torch.int64, device=dev)
    o_rew = torch.zeros(GB, dtype=torch.float32, device=dev)
    o_done = torch.zeros(GB, dtype=torch.float32, device=dev)
    o_prio = torch.empty(GB, dtype=torch.float32, device=dev)
    o_loss = torch.empty(1, dtype=torch.float32, device=dev)
    # world > 1: the TD arithmetic runs over the full batch on every rank (its 0.9*max term is
    # batch-wide); rows owned by other ranks are stood in for by replicated synthetic columns
    t_act = o_act if world == 1 else torch.randint(0, N_ACTIONS, (GB,), dtype=torch.int64, device=dev, generator=cgen)
    t_rew = o_rew if world == 1 else torch.randint(-1, 2, (GB,), device=dev, generator=cgen).float()
    t_done = o_done if world == 1 else torch.zeros(GB, device=dev)
    t_w = w32 if world == 1 else torch.rand(GB, device=dev, generator=cgen)
    shard_rec = torch.zeros(4, dtype=torch.float64, device=dev)
    all_rec = torch.zeros(world * 4, dtype=torch.float64, device=dev)

    lib = pn.lib
    h = buf._h if world == 1 else shard._h
    stream = torch.cuda.current_stream()
    sp = pn.stream_ptr(stream)
    P = pn.ptr
    gamma_n = float(np.float32(GAMMA ** N_STEP))
    a_idx, a_w32 = P(idx), P(w32)
    a_gather = (h, a_idx, GB, P(o_obs), P(o_nxt), P(o_act), P(o_rew), P(o_done), None, sp)
    a_td = (h, P(q), P(qn), P(qt), N_ACTIONS, P(t_act), P(t_rew), P(t_done), P(t_w), gamma_n, a_idx, GB, P(o_prio),
            None, P(o_loss), sp)
    u_ptrs = [pn.C.c_void_p(u_all[i].data_ptr()) for i in range(n_u)]
    a_rec, a_all = P(shard_rec), P(all_rec)
    # kernels per step: k_sample / k_sample_sharded, k_gather_bal, k_update_cluster (TD fused in front; global batches above
    # 1 024 on trees the dense form serves: k_update_cluster as the TD launch + k_upd_dense_sorted); without the mailbox the
    # sharded step adds k_shard_record (+ NCCL's all-gather kernel, not ours)
    dense_lmax = int(os.environ.get("PER_DENSE_LMAX", "0"))
    if "PER_DENSE_LMAX" in os.environ:
        pn.check(pn.lib.per_tune(h, b"dense_update_lmax", dense_lmax))
    tree_log2 = (ring - 1).bit_length()  # this rank's tree
    two_launch_update = os.environ.get("PER_SORTED_UPDATE", "1") != "0" and 12 <= tree_log2 <= 26 and (
        tree_log2 <= dense_lmax or (BATCH > (1024 if tree_log2 <= 20 else 2048) and min(BATCH, (1 << tree_log2) >> 10) <= 7 * 1184))  # dense_update_ok()
    launches_per_step = (3 if (world == 1 or use_mailbox) else 4) + (1 if two_launch_update else 0)

    if world > 1 and not args.no_parity_check:
        # the sharded path against the single-tree oracle on the very GPUs about to be timed (tests/multi_gpu_parity.py;
        # a checker only: a small separate replay, before anything is measured)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import multi_gpu_parity
        multi_gpu_parity.run_check(rank, world, local_rank, size=args.parity_size, steps=4 if args.parity_size < (1 << 20) else 2,
                                   log=sys.stderr.write)
        dist.barrier()
    if use_mailbox:
        # self-check: records through the mailboxes must give the indices the all-gathered records give
        pn.check(lib.per_shard_record(h, glen, a_rec, sp))
        dist.all_gather_into_tensor(all_rec, shard_rec)
        pn.check(lib.per_sample_sharded(h, a_all, u_ptrs[0], GB, BETA, glen, a_idx, None, a_w32, sp))
        idx_collective = idx.clone()
        pn.check(lib.per_sample_sharded(h, None, u_ptrs[0], GB, BETA, glen, a_idx, None, a_w32, sp))
        torch.cuda.synchronize()
        assert torch.equal(idx, idx_collective), "peer-memory exchange and NCCL exchange disagree"
        assert int((idx >= 0).sum().item()) > 0
        if not args.no_early_publish:  # the update kernel's tail publishes the next sample's record
            shard.enable_early_publish()

    def step(i, ev=None):
        if use_mailbox:  # the sampling kernel publishes / awaits the 32-byte records over NVLink itself
            rc = lib.per_sample_sharded(h, None, u_ptrs[i % n_u], GB, BETA, glen, a_idx, None, a_w32, sp)
        elif world > 1:  # the one exchange of the path as an NCCL all-gather
            rc = lib.per_shard_record(h, glen, a_rec, sp)
            dist.all_gather_into_tensor(all_rec, shard_rec)
            rc |= lib.per_sample_sharded(h, a_all, u_ptrs[i % n_u], GB, BETA, glen, a_idx, None, a_w32, sp)
        else:
            rc = lib.per_sample_indices(h, u_ptrs[i % n_u], GB, BETA, a_idx, None, a_w32, sp)
        if ev is not None:
            ev[0].record(stream)
        rc |= lib.per_gather(*a_gather)
        if ev is not None:
            ev[1].record(stream)
        rc |= lib.per_td_priority_update(*a_td)
        if rc:
            pn.check(rc)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for i in range(W):
        step(i)
    barrier()

    if args.trace:  # diagnostic: events between the three calls of a step
        n_tr = 200
        evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(n_tr)]
        for i in range(n_tr):
            evs[i][0].record(stream)
            if use_mailbox:
                lib.per_sample_sharded(h, None, u_ptrs[i % n_u], GB, BETA, glen, a_idx, None, a_w32, sp)
            elif world > 1:
                lib.per_shard_record(h, glen, a_rec, sp)
                dist.all_gather_into_tensor(all_rec, shard_rec)
                lib.per_sample_sharded(h, a_all, u_ptrs[i % n_u], GB, BETA, glen, a_idx, None, a_w32, sp)
            else:
                lib.per_sample_indices(h, u_ptrs[i % n_u], GB, BETA, a_idx, None, a_w32, sp)
            evs[i][1].record(stream)
            lib.per_gather(*a_gather)
            evs[i][2].record(stream)
            lib.per_td_priority_update(*a_td)
            evs[i][3].record(stream)
        barrier()
        tr = [float(np.median([e[k].elapsed_time(e[k + 1]) for e in evs[20:]])) * 1e3 for k in range(3)]
        sys.stderr.write("rank %d trace (us, median): sample %.1f  gather %.1f  td+update %.1f  own rows %d\n"
                         % (rank, tr[0], tr[1], tr[2], int((idx >= 0).sum().item())))

    # ---- CUDA graphs for the launch-bound loop ----------------------------------------------
    # The step is three small launches, so plain Python launches are CPU-bound.  Timed region = the K steps
    # replayed as CUDA graphs of G = min(31, K) whole steps each (plus one graph with the K mod G left-over
    # steps): exactly K steps, 3 K launches, nothing else on the stream.  Falls back to eager launches if capture fails.
    GRAPH_STEPS_MAX, N_GRAPHS = 31, 4
    G = max(1, min(GRAPH_STEPS_MAX, K))
    n_full, rem = K // G, K % G
    graphs, tail_graph, launch_mode = None, None, "eager launches (one Python call per kernel)"

    def capture(body):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            body()
        return g

    def enqueue_head(j, sps):  # TD + tree update of the previous batch, record exchange, sample
        rc = lib.per_td_priority_update(*(a_td[:-1] + (sps,)))
        if world > 1:
            rc |= lib.per_sample_sharded(h, None, u_ptrs[j % n_u], GB, BETA, glen, a_idx, None, a_w32, sps)
        else:
            rc |= lib.per_sample_indices(h, u_ptrs[j % n_u], GB, BETA, a_idx, None, a_w32, sps)
        if rc:
            pn.check(rc)

    def full_block(base, n, sps):
        for t in range(n):
            enqueue_head(base + t, sps)
            rc = lib.per_gather(*(a_gather[:-1] + (sps,)))
            if rc:
                pn.check(rc)

    def draw_batch():  # the graphs run a step as [TD + update of the previous batch, sample, gather]
        if use_mailbox:
            pn.check(lib.per_sample_sharded(h, None, u_ptrs[0], GB, BETA, glen, a_idx, None, a_w32, sp))
        else:
            pn.check(lib.per_sample_indices(h, u_ptrs[0], GB, BETA, a_idx, None, a_w32, sp))
        pn.check(lib.per_gather(*a_gather))

    # (capturing torch's NCCL all-gather inside a graph hung on 2 GPUs in round 1; with the mailbox exchange the
    # step is kernels only, so world > 1 is captured too)
    side = torch.cuda.Stream()
    if not args.eager and (world == 1 or use_mailbox):
        try:
            # the warm-up ended with an update, so draw one batch first (with early publish two updates in a row are refused)
            draw_batch()
            side.wait_stream(stream)
            with torch.cuda.stream(side):
                sp_side = pn.stream_ptr(side)
                built = [capture(lambda j=j: full_block(j * G, G, sp_side)) for j in range(min(N_GRAPHS, n_full))]
                tail = capture(lambda: full_block(N_GRAPHS * G, rem, sp_side)) if rem else None
            stream.wait_stream(side)
            torch.cuda.synchronize()
            graphs, tail_graph = built, tail
            launch_mode = ("cuda graphs: %d replay(s) of a %d-step graph%s; a step = TD+update of the previous batch, sample, "
                           "gather (3 launches)" % (n_full, G, (" + one %d-step graph" % rem) if rem else ""))
        except Exception as exc:
            sys.stderr.write("graph capture failed, timing eager launches: %s\n" % str(exc)[:300])
            graphs = None

    def timed(n_steps, use_graphs):
        """Runs exactly n_steps steps; returns ms (max over ranks)."""
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        # an untimed ~100 us spin in front of the start event: the GPU is busy while the host enqueues the event and the
        # first graph, so the K steps start the moment the event fires (otherwise the host's graph-launch latency, ~15 us,
        # sits inside a timed region that is only 0.6 ms long at K = 20)
        torch.cuda._sleep(200000)
        t0.record(stream)
        if use_graphs:
            assert n_steps == K
            for b_ in range(n_full):
                graphs[b_ % len(graphs)].replay()
            if tail_graph is not None:
                tail_graph.replay()
        else:
            for i in range(n_steps):
                step(W + i)
        t1.record(stream)
        barrier()
        return max_over_ranks(t0.elapsed_time(t1))

    # clocks / throttle reasons are sampled from the graph warm-up to the end of the roofline pass (the timed region
    # itself can be shorter than one NVML sample at small K)
    sampler = ClockSampler(local_rank, period=0.005)
    sampler.start()
    if graphs is not None:
        timed(K, True)  # warm the graphs (untimed)
    # ---- timed region: K steps, CUDA events on the launching stream --------------------------
    if use_mailbox:
        pn.check(lib.per_shard_exchange_stats(h, (pn.C.c_int64 * 3)(), sp))  # clear
    ms = timed(K, graphs is not None)
    exchange_us = None
    if use_mailbox:
        st3 = (pn.C.c_int64 * 3)()
        pn.check(lib.per_shard_exchange_stats(h, st3, sp))
        exchange_us = {"mean": st3[1] / max(st3[0], 1) / 1e3, "max": st3[2] / 1e3, "samples": int(st3[0])}

    # the graphs end with a sample + gather; close the step so that the passes below start from a clean state
    def close_step():
        rc = lib.per_td_priority_update(*a_td)
        if rc:
            pn.check(rc)

    # ---- roofline pass: the gather kernel by itself --------------------------------------------------
    # M launches, each on its own freshly sampled index set (M x 29 MB of distinct rows: nothing is re-read
    # from L2), captured as one graph without programmatic chaining so that the launches run one after the
    # other, one CUDA-event pair around each replay.  average launch duration = elapsed / M.
    M_SETS, REPS = 16, 8
    close_step()
    idx_sets = [torch.empty(GB, dtype=torch.int64, device=dev) for _ in range(M_SETS)]
    for i in range(M_SETS):
        if use_mailbox:
            pn.check(lib.per_sample_sharded(h, None, u_ptrs[(7 * i + 3) % n_u], GB, BETA, glen, P(idx_sets[i]), None, a_w32, sp))
        elif world > 1:
            pn.check(lib.per_shard_record(h, glen, a_rec, sp))
            dist.all_gather_into_tensor(all_rec, shard_rec)
            pn.check(lib.per_sample_sharded(h, a_all, u_ptrs[(7 * i + 3) % n_u], GB, BETA, glen, P(idx_sets[i]), None, a_w32, sp))
        else:
            pn.check(lib.per_sample_indices(h, u_ptrs[(7 * i + 3) % n_u], GB, BETA, P(idx_sets[i]), None, a_w32, sp))
    torch.cuda.synchronize()
    if world > 1:  # this rank's rows of each sample, as a dense index list (strata of other ranks hold -1)
        idx_sets = [t[t >= 0].contiguous() for t in idx_sets]
    rows_per_set = [int(t.numel()) for t in idx_sets]
    pdl_default = 3 if world == 1 else 7
    pn.check(lib.per_tune(h, b"pdl", 0))
    gather_ms = None
    try:
        side.wait_stream(stream)
        with torch.cuda.stream(side):
            sp_side = pn.stream_ptr(side)

            def gather_block():
                for i in range(M_SETS):
                    rc = lib.per_gather(h, P(idx_sets[i]), rows_per_set[i], P(o_obs), P(o_nxt), P(o_act), P(o_rew), P(o_done), None, sp_side)
                    if rc:
                        pn.check(rc)
            gather_block()
            g_gather = capture(gather_block)
        stream.wait_stream(side)
        g_gather.replay()
        torch.cuda.synchronize()
        times = []
        for _ in range(REPS):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            g_gather.replay()
            e1.record(stream)
            torch.cuda.synchronize()
            times.append(e0.elapsed_time(e1) / M_SETS)
        gather_ms = float(np.mean(times))
    finally:
        pn.check(lib.per_tune(h, b"pdl", int(os.environ.get("PER_PDL", pdl_default))))
    clocks = sampler.stop()
    # back to the loop's own buffers: one sample + gather into idx / o_*
    draw_batch()

    # ---- step anatomy: the same graph blocks captured once more with the kernels' GPU-timer stamps on ----
    # (per_timeline: plain stores by CTA 0 of every launch, and for the gather the latest finish over the whole
    # grid).  Outside the timed region; tells how long each kernel takes INSIDE the chained step.
    anatomy = None
    if graphs is not None:
        cap_tl = 8192
        pn.check(lib.per_tune(h, b"timeline", cap_tl))
        if args.timeline:
            pn.check(lib.per_tune(h, b"update_trace", 1))  # + the update kernel's own phase stamps (last launch)
        with torch.cuda.stream(side):
            sp_side = pn.stream_ptr(side)
            probe = [capture(lambda j=j: full_block(j * GRAPH_STEPS_MAX, GRAPH_STEPS_MAX, sp_side)) for j in range(3)]
        stream.wait_stream(side)
        arr = (pn.C.c_int64 * cap_tl)()
        cnt = pn.C.c_int64(0)
        for g in probe:  # warm
            g.replay()
        barrier()
        pn.check(lib.per_timeline(h, arr, cap_tl, pn.C.byref(cnt), sp))  # clear
        barrier()
        for g in probe:
            g.replay()
        barrier()
        pn.check(lib.per_timeline(h, arr, cap_tl, pn.C.byref(cnt), sp))
        pn.check(lib.per_tune(h, b"timeline", 0))
        if args.timeline:
            st16 = (pn.C.c_int64 * 16)()
            pn.check(lib.per_update_trace(h, st16, sp))
            pn.check(lib.per_tune(h, b"update_trace", 0))
            ph = ["wait passed", "TD done", "arbitration", "leaves/barrier", "rounds done", "top rebuilt", "TD staged", "CTA max",
                  "TD cl.sync", "batch max", "leaves+sib in", "climb done", "top loaded", "top written"]
            order = [6, 7, 8, 9, 1, 10, 2, 11, 3, 4, 12, 5, 13]
            sys.stderr.write("rank %d update phases (us): " % rank + " | ".join("%s %.2f" % (ph[i], (st16[i] - st16[0]) / 1e3) for i in order if st16[i] > st16[0]) + "\n")
        ent = sorted(((int(arr[i]) & ((1 << 56) - 1), (int(arr[i]) >> 56) & 0xff) for i in range(cnt.value)))
        names = {7: "update entered", 8: "update past wait", 12: "TD done", 9: "update CTA0 done", 1: "sample entered",
                 10: "records in", 2: "sample past wait", 11: "strata known", 3: "sample CTA0 done", 4: "gather entered",
                 5: "gather past wait", 6: "gather CTA0 done", 13: "gather grid done", 20: "sample shared part in", 21: "descent done"}
        steps_tl, cur = [], None
        for t, tag in ent:  # a step of the graph = [update, sample, gather]; anchor = update past its wait
            if tag == 8:
                if cur:
                    steps_tl.append(cur)
                cur = {"t0": t}
            elif cur is not None and tag not in cur:
                cur[tag] = t - cur["t0"]
        med = lambda xs: float(np.median(xs)) / 1e3 if len(xs) else None
        body = steps_tl[2:]
        periods = np.diff([c["t0"] for c in steps_tl]) if len(steps_tl) > 2 else np.array([])
        span = lambda a_, b_: med([c[b_] - (c[a_] if a_ != 8 else 0) for c in body if b_ in c and (a_ == 8 or a_ in c)])
        anatomy = {"steps": len(body), "period": med(periods), "td_arithmetic": span(8, 12), "update": span(8, 9),
                   "sample": span(2, 3), "gather_cta0": span(5, 6), "gather_grid": span(5, 13),
                   "how": "GPU global-timer stamps written by the kernels (per_timeline) in 3 replayed 31-step graph blocks after "
                          "the timed region: past the programmatic-launch wait -> done"}
        if args.timeline:
            msg = "rank %d timeline (us after 'update past wait', median of %d steps; period %.2f): " % (rank, len(body), anatomy["period"] or 0.0)
            for tag in (12, 9, 1, 10, 2, 20, 11, 21, 3, 4, 5, 6, 13, 7):
                vals = [c[tag] for c in body if tag in c]
                if vals:
                    msg += "%s %.2f | " % (names[tag], med(vals))
            sys.stderr.write(msg + "\n")
    value = GB * K / (ms * 1e-3)
    rows_this_rank = int((idx >= 0).sum().item())
    status = pn.C.c_int64(0)
    pn.check(lib.per_status(h, pn.C.byref(status), sp))
    assert status.value == 0, "device status word set during the bench: %d" % status.value
    # the same steps launched one call at a time (shows what the graphs remove)
    eager_value = None
    if graphs is not None:
        k2 = min(K, 300)
        ms2 = timed(k2, False)
        eager_value = GB * k2 / (ms2 * 1e-3)

    # ---- e2e: the public API with host buffers ---------------------------------------------
    e2e_steps = max(3, min(K, 100))
    if world == 1:
        # reference-facing API: sample() -> host numpy (frames cross PCIe), update_priorities(host arrays)
        prios_host = [np.abs(np.random.default_rng(i).standard_normal(BATCH)).astype(np.float32) + 1e-6 for i in range(8)]
        import random as _random
        _random.seed(0)
        for i in range(3):
            out = buf.sample(BATCH, BETA)
            buf.update_priorities(out[-1], prios_host[i % 8])
        barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            out = buf.sample(BATCH, BETA)                      # uniforms H2D; frames, scalars, weights, idx D2H
            buf.update_priorities(out[-1], prios_host[i % 8])  # idx + leaves H2D
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        d2h = sum(int(np.asarray(x).nbytes) for x in out[:-1]) + 8 * BATCH
        h2d = 8 * BATCH + 8 * BATCH + 8 * BATCH
        api = "memory.CustomPrioritizedReplayBuffer.sample()/update_priorities() with host numpy in/out"
    else:
        # sharded replay through host buffers, the reference's calling convention on every rank: the global batch's
        # uniforms go up, this rank's share of the batch (frames included) comes back to host memory, and the global
        # batch's (indices, priorities) go up again for the update
        rng_e = np.random.default_rng(5)
        u_host = [rng_e.random(GB) for _ in range(8)]
        prios_host = [np.abs(rng_e.standard_normal(GB)) + 1e-6 for _ in range(8)]
        own_rows = torch.zeros(1, dtype=torch.float64, device=dev)

        def e2e_round(i):
            out = shard.sample_host(GB, BETA, u_host[i % 8])
            idx_all = torch.full((GB,), -1, dtype=torch.int64, device=dev)
            idx_all[out["lo"]:out["hi"]] = torch.as_tensor(out["idx"]).to(dev)
            dist.all_reduce(idx_all, op=dist.ReduceOp.MAX)  # every rank learns the whole batch's indices (32 KB)
            shard.update_priorities_host(idx_all.cpu().numpy(), prios_host[i % 8])
            return out

        for i in range(3):
            out = e2e_round(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            out = e2e_round(i)
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        rows = out["hi"] - out["lo"]
        own_rows[0] = rows
        dist.all_reduce(own_rows)
        # bytes over PCIe per step, summed over the ranks: frames + scalars + weights + indices down; uniforms, indices, leaves up
        d2h = int(own_rows.item()) * (2 * OBS_BYTES + 8 + 4 + 4 + 8) + world * 8 * GB * 2
        h2d = world * (8 * GB * 3)
        api = ("sharded.ShardedReplay.sample_host()/update_priorities_host() on every rank: host uniforms in, the rank's share of "
               "the batch (frames included) out to pinned host memory, global (indices, priorities) back in")
    e2e_s = max_over_ranks(e2e_s)
    e2e = {"value": GB * e2e_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "steps": e2e_steps, "api": api}

    # ---- learner-ward delivery (SURVEY 8e data movement): every rank's rows into rank 0's memory over NVLink ----------
    p2p = None
    if world > 1 and use_mailbox and not args.no_extras:
        lb = shard.open_learner_batch(GB, learner_rank=0)
        if rank == 0:
            lb["prio"].copy_(torch.rand(GB, device=dev) + 1e-3)
        lb.fence()

        def p2p_step(i):
            d = shard.sample_to_learner(lb, BETA, u_all[i % n_u])   # sample + gather into the learner's buffers + idx / weights
            shard.update_from_learner(lb, d["idx"])                 # owned priorities read back over NVLink + tree update
            return d

        for i in range(5):
            d_last = p2p_step(i)
        lb.fence()
        n_p2p = 50
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(n_p2p):
            d_last = p2p_step(5 + i)
        e1.record(stream)
        lb.fence()
        p2p_us = max_over_ranks(e0.elapsed_time(e1) * 1e3 / n_p2p)
        own0 = torch.tensor([int((d_last["idx"] >= 0).sum().item()) if rank == 0 else 0], device=dev)
        dist.all_reduce(own0)
        remote_rows = GB - int(own0.item())
        nv_peak = 770.0  # measured peer copy per direction on this pool (B200_PROFILING.md)
        p2p = {"us_per_step": p2p_us, "rows_per_s_into_learner": GB / (p2p_us * 1e-6),
               "rows_crossing_nvlink_per_step": remote_rows,
               "nvlink_ingest_GBps": remote_rows * (2 * OBS_BYTES + 16) / (p2p_us * 1e-6) / 1e9,
               "nvlink_peak_GBps": nv_peak, "frac": remote_rows * (2 * OBS_BYTES + 16) / (p2p_us * 1e-6) / 1e9 / nv_peak,
               "api": "sharded.ShardedReplay.sample_to_learner() / update_from_learner(): rows, indices and weights stored into "
                      "rank 0's batch buffers by the owners' kernels (CUDA IPC mappings, plain stores over NVLink), priorities read "
                      "back from rank 0's memory; plain launches, CUDA events, max over ranks",
               "note": "one learner GPU takes at most ~0.77 TB/s: the sample+update rate of the headline is with every GPU training "
                       "on the strata it owns"}
        del lb

    # ---- the other measurement rows of SURVEY section 8(d), N = 1 only -------------------------------
    aql_block = learner = None
    if rank == 0 and world == 1 and not args.no_extras:
        try:
            aql_block = aql_row(dev)
        except Exception as exc:  # an extra row must never cost the headline line
            aql_block = {"error": str(exc)[:200]}
        try:
            learner = learner_row(buf, dev)
        except Exception as exc:
            learner = {"error": str(exc)[:200]}

    # ---- CPU baseline on this box (rank 0, N=1 only) -----------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, done_steps, dt = run_cpu_baseline(10 ** 6, 3, budget_s=15.0)
        cpu = {"value": v, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": "%d steps (%.1f s) of sample(512)+update_priorities(512) on a 2^20-leaf tree with the "
                         "oracle's pure-Python port of memory.py's CustomPrioritizedReplayBuffer (the class ApeX.py uses: "
                         "observations come back as lists of references, memory.py:348-362, no frame copy); host has %d "
                         "cores, the reference path is serial Python" % (done_steps, dt, os.cpu_count())}

    if rank == 0:
        peak, peak_src = peaks()
        mean_rows = float(np.mean(rows_per_set))
        launch_bytes = GATHER_BYTES_PER_ROW * mean_rows  # rows one gather launch of the roofline pass moved on this rank
        achieved = launch_bytes / (gather_ms * 1e-3) / 1e9
        traffic, traffic_kernel = ncu_gather_traffic()
        mode = int(os.environ.get("PER_GATHER_MODE", DEFAULT_GATHER_MODE))
        kernel_name = {0: "k_gather_tma", 1: "k_gather_ldg<16>", 2: "k_gather_bal"}[mode]
        L_tree = int(np.log2(ring))
        tree_bytes_sample = BATCH * (8 * L_tree + 20)
        tree_bytes_update = 2 * (8 * BATCH + 24 * sum(min(BATCH, 2 ** l) for l in range(L_tree)))
        roofline_tree = None
        if anatomy and anatomy.get("sample") and anatomy.get("update"):
            upd_us = anatomy["update"] - (anatomy.get("td_arithmetic") or 0.0)
            roofline_tree = {
                "bound": "latency (L2-resident trees; SURVEY 8(d) names HBM)", "peak": peak, "unit": "GB/s",
                "sample": {"kernel": "k_sample" if world == 1 else "k_sample_sharded", "bytes_per_launch": tree_bytes_sample,
                           "in_step_us": anatomy["sample"], "achieved": tree_bytes_sample / anatomy["sample"] / 1e3,
                           "frac": tree_bytes_sample / anatomy["sample"] / 1e3 / peak},
                "update": {"kernel": "k_update_cluster (tree part, TD prologue excluded)", "bytes_per_launch": tree_bytes_update,
                           "in_step_us": upd_us, "achieved": tree_bytes_update / upd_us / 1e3,
                           "frac": tree_bytes_update / upd_us / 1e3 / peak},
                "how": "SURVEY 8(d) algorithmic bytes / the kernel's in-step duration (step_anatomy_us); both kernels are chains of "
                       "dependent L2 round trips at batch 512 (0.4 MB per step), see profiles/ for the large-batch sweep where the "
                       "same kernels are bandwidth-bound",
            }
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong" if args.global_batch else "weak", "vs_baseline": None,
            "dtype": "f64 trees / u8 frames / f32 TD", "data": "synthetic", "config": workload_config(world, ring),
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches_per_step * K,
            "roofline": {"bound": "hbm", "kernel": kernel_name, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic if (world == 1 and traffic_kernel == kernel_name) else None,
                         "peak_source": peak_src, "bytes_per_launch": launch_bytes, "avg_launch_us": gather_ms * 1e3,
                         "how": "one CUDA-event pair around a graph of %d back-to-back launches (no programmatic chaining), each on "
                                "its own sampled index set; mean of %d replays, right after the timed region" % (M_SETS, REPS),
                         "in_step": None if not (anatomy and anatomy.get("gather_grid")) else {
                             "avg_launch_us": anatomy["gather_grid"],
                             "achieved": GATHER_BYTES_PER_ROW * rows_this_rank / (anatomy["gather_grid"] * 1e-6) / 1e9,
                             "frac": GATHER_BYTES_PER_ROW * rows_this_rank / (anatomy["gather_grid"] * 1e-6) / 1e9 / peak,
                             "how": "supplementary: the launch as it runs inside the chained step, timed by the kernel's own "
                                    "GPU-timer stamps (wait passed -> last lane of the grid done)"}},
            "roofline_tree": roofline_tree,
            "cpu_baseline": cpu, "launch": launch_mode, "value_eager_launches": eager_value,
            "learner_step_us": {"total": 1e3 * ms / K, "gather_alone": gather_ms * 1e3},
            "record_exchange_us": exchange_us, "step_anatomy_us": anatomy,
            "aql": aql_block, "learner_updates_per_s": learner, "p2p": p2p,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
