#!/usr/bin/env python
"""bench.py — prioritized-replay hot path on B200: sample (prefix-sum descent + importance
weights + frame gather) -> TD-error -> priority -> tree update, per batch.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Prints ONE JSON line (rank 0).  `value` = transitions sampled+updated per second over all
ranks with every input resident in HBM; `e2e` = the same loop through the reference-facing
Python API (memory.py: sample() -> host numpy, update_priorities(host arrays)), host<->device
copies inside the timed region; `roofline` = the frame-gather kernel (the HBM-bound one)
timed with CUDA events around each of its launches inside the timed region;
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
METRIC = "PER sample+update transitions/s"
UNIT = "transitions/s"
# per sampled transition the gather reads 2*28224 B of frames + 16 B of
# scalars and writes the same.
GATHER_BYTES_PER_ROW = 2 * (2 * OBS_BYTES + 16)
# dram__bytes_read.sum + dram__bytes_write.sum of one k_gather_ldg<16> launch at this workload, from the
# ncu --set full capture summarised in profiles/r01e_ncu_summary.csv (the 28.9 MB written stay in the 126 MB L2)
NCU_GATHER_TRAFFIC = 29224192 + 315392


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


def cpu_step(buf, rng):
    """One reference learner-side replay step: sample(B, beta) then update_priorities."""
    u = rng.random(BATCH)
    out = buf.sample(BATCH, BETA, u)
    pr = np.abs(rng.standard_normal(BATCH)) + 1e-6
    buf.update_priorities(out[-1], [float(p) for p in pr])


def run_cpu_baseline(steps, warmup, budget_s=None):
    buf = build_cpu_port(RING)
    rng = np.random.default_rng(1)
    for _ in range(warmup):
        cpu_step(buf, rng)
    t0 = time.perf_counter()
    done = 0
    for _ in range(steps):
        cpu_step(buf, rng)
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    return done * BATCH / dt, done, dt


def reference_arm(args, rank):
    """--impl reference: the reference's own CPU algorithm (oracle port: pure Python over
    Python lists, like memory.py — it is serial, so 1 core) on the same config and metric."""
    if rank != 0:
        return
    steps = min(args.steps, 400)  # ~25 ms per step; keeps the run within minutes
    value, done, dt = run_cpu_baseline(steps, min(args.warmup, 5))
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus,
        "steps": done, "warmup": min(args.warmup, 5), "ms_per_step": 1e3 * dt / max(done, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(1),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "port",
                         "sample": "%d steps of sample(512)+update_priorities(512) on a 2^20-leaf tree, "
                                   "oracle/per_oracle.py ReplayOracle (pure Python like memory.py), "
                                   "host has %d cores" % (done, os.cpu_count())},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_config(n_gpus):
    return {"workload": "Ape-X learner replay step, configs[1]: ring 2^20 Atari transitions per GPU (59.2 GB HBM), "
                        "batch %d per GPU, alpha 0.6, beta 0.4, n-step 3, 18 actions" % BATCH,
            "ring_per_gpu": RING, "batch_per_gpu": BATCH, "obs": "uint8[4,84,84]",
            "sharding": "none" if n_gpus == 1 else "one replay of %d x 2^20 transitions cut into %d contiguous slices; "
                        "per step one exchange of a 32-byte record per rank (peer-memory mailboxes written by the sampling kernel over "
                        "NVLink, or --nccl-exchange), global stratified batch %d" % (n_gpus, n_gpus, BATCH * n_gpus),
            "l2": "inputs (59 GB ring) larger than L2; no flush", "parallelism": "shard%d" % n_gpus}


# ------------------------------------------------------------------------------- GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--ring", type=int, default=RING, help="rows per GPU (debug: smaller ring)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    if os.environ.get("PER_GATHER_MODE") is not None:  # experiment: 1 = register-staged gather, no shared memory
        hh = buf._h if world == 1 else shard._h
        pn.check(pn.lib.per_tune(hh, b"gather_mode", int(os.environ["PER_GATHER_MODE"])))
        pn.check(pn.lib.per_tune(hh, b"gather_ctas_per_sm", int(os.environ.get("PER_GATHER_CTAS", "2"))))
        if os.environ.get("PER_GATHER_PIECE"):
            pn.check(pn.lib.per_tune(hh, b"gather_piece", int(os.environ["PER_GATHER_PIECE"])))
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
    # kernels per step: k_sample, k_gather_ldg, k_update_cluster (TD fused in front); sharded adds
    # k_shard_record (+ NCCL's all-gather kernel, not ours) and B > 2048 uses the multi-launch update
    launches_per_step = 3 if (world == 1 or use_mailbox) else 4

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
    # The step is four small launches, so plain Python launches are CPU-bound.  Timed region =
    # blocks of GRAPH_STEPS whole steps replayed as ONE CUDA graph, each block followed by one
    # instrumented step whose gather is launched by hand between two timing events (the
    # roofline's per-launch duration comes from those).  Falls back to eager launches if capture fails.
    GRAPH_STEPS, N_GRAPHS = 31, 4  # one instrumented step in 32: 62 gather timings in the default 2000 steps
    graphs, head_graphs, launch_mode = None, None, "eager launches"

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

    # (capturing torch's NCCL all-gather inside a graph hung on 2 GPUs in round 1; with the mailbox exchange the
    # step is kernels only, so world > 1 is captured too)
    if not args.eager and (world == 1 or use_mailbox):
        try:
            # the graphs run a step as [TD + update of the previous batch, sample, gather]; the warm-up ended with
            # an update, so draw one batch first (with early publish two updates in a row are refused)
            if use_mailbox:
                pn.check(lib.per_sample_sharded(h, None, u_ptrs[0], GB, BETA, glen, a_idx, None, a_w32, sp))
            else:
                pn.check(lib.per_sample_indices(h, u_ptrs[0], GB, BETA, a_idx, None, a_w32, sp))
            pn.check(lib.per_gather(*a_gather))
            side = torch.cuda.Stream()
            side.wait_stream(stream)
            with torch.cuda.stream(side):
                sp_side = pn.stream_ptr(side)

                def full_block(base):
                    for t in range(GRAPH_STEPS):
                        enqueue_head(base + t, sp_side)
                        rc = lib.per_gather(*(a_gather[:-1] + (sp_side,)))
                        if rc:
                            pn.check(rc)

                built = [capture(lambda j=j: full_block(j * GRAPH_STEPS)) for j in range(N_GRAPHS)]
                heads = [capture(lambda j=j: enqueue_head(N_GRAPHS * GRAPH_STEPS + j, sp_side)) for j in range(N_GRAPHS)]
            stream.wait_stream(side)
            torch.cuda.synchronize()
            graphs, head_graphs = built, heads
            launch_mode = ("cuda graphs: blocks of %d whole steps per graph + 1 instrumented step "
                           "(gather launched between timing events)" % GRAPH_STEPS)
        except Exception as exc:
            sys.stderr.write("graph capture failed, timing eager launches: %s\n" % str(exc)[:300])
            graphs = None

    def timed(n_steps, use_graphs):
        """Runs exactly n_steps steps; returns (ms over all ranks, mean ms of the instrumented gather launches)."""
        evs = []
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        t0.record(stream)
        if use_graphs:
            done, blk = 0, 0
            while done < n_steps:
                if n_steps - done > GRAPH_STEPS:
                    graphs[blk % N_GRAPHS].replay()
                    done += GRAPH_STEPS
                ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                head_graphs[blk % N_GRAPHS].replay()
                ev[0].record(stream)
                rc = lib.per_gather(*a_gather)
                ev[1].record(stream)
                if rc:
                    pn.check(rc)
                evs.append(ev)
                done += 1
                blk += 1
        else:
            for i in range(n_steps):
                ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                step(W + i, ev)
                evs.append(ev)
        t1.record(stream)
        barrier()
        return max_over_ranks(t0.elapsed_time(t1)), float(np.mean([x.elapsed_time(y) for x, y in evs]))

    if graphs is not None:
        timed(min(K, 64), True)  # warm the graphs
    # ---- timed region: K steps, CUDA events on the launching stream --------------------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    if use_mailbox:
        pn.check(lib.per_shard_exchange_stats(h, (pn.C.c_int64 * 3)(), sp))  # clear
    ms, gather_ms = timed(K, graphs is not None)
    clocks = sampler.stop()
    exchange_us = None
    if use_mailbox:
        st3 = (pn.C.c_int64 * 3)()
        pn.check(lib.per_shard_exchange_stats(h, st3, sp))
        exchange_us = {"mean": st3[1] / max(st3[0], 1) / 1e3, "max": st3[2] / 1e3, "samples": int(st3[0])}
    # ---- step anatomy: the same graph blocks captured once more with the kernels' GPU-timer stamps on ----
    # (per_timeline: plain stores by CTA 0 of every launch, and for the gather the latest finish over the whole
    # grid).  Outside the timed region; tells how long each kernel takes INSIDE the chained step, which the
    # event-bracketed gather above cannot: an event between two kernels breaks their programmatic chaining.
    anatomy = None
    if graphs is not None:
        cap_tl = 8192
        pn.check(lib.per_tune(h, b"timeline", cap_tl))
        with torch.cuda.stream(side):
            probe = [capture(lambda j=j: full_block(j * GRAPH_STEPS)) for j in range(3)]
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
        ent = sorted(((int(arr[i]) & ((1 << 56) - 1), (int(arr[i]) >> 56) & 0xff) for i in range(cnt.value)))
        names = {7: "update entered", 8: "update past wait", 12: "TD done", 9: "update CTA0 done", 1: "sample entered",
                 10: "records in", 2: "sample past wait", 11: "strata known", 3: "sample CTA0 done", 4: "gather entered",
                 5: "gather past wait", 6: "gather CTA0 done", 13: "gather grid done"}
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
                   "how": "GPU global-timer stamps written by the kernels (per_timeline) in 3 replayed graph blocks after "
                          "the timed region: past the programmatic-launch wait -> done"}
        if args.timeline:
            msg = "rank %d timeline (us after 'update past wait', median of %d steps; period %.2f): " % (rank, len(body), anatomy["period"] or 0.0)
            for tag in (12, 9, 1, 10, 2, 11, 3, 4, 5, 6, 13, 7):
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
        ms2, _ = timed(k2, False)
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
        # sharded API: uniforms arrive from pinned host memory, frames stay on the GPUs that train on
        # them, the step's result (loss + the batch's new priorities) is read back to the host
        u_host = [torch.rand(GB, dtype=torch.float64, generator=torch.Generator().manual_seed(i)).pin_memory() for i in range(8)]
        res_host = torch.empty(GB + 1, dtype=torch.float32).pin_memory()
        outbuf = {"obs": o_obs, "next_obs": o_nxt, "act": o_act, "rew": o_rew, "done": o_done,
                  "idx": idx, "weights": w32, "weights64": torch.empty(GB, dtype=torch.float64, device=dev)}
        u_dev = torch.empty(GB, dtype=torch.float64, device=dev)
        o_res = torch.empty(GB + 1, dtype=torch.float32, device=dev)  # priorities | loss: one read-back per step

        def e2e_step(i):
            u_dev.copy_(u_host[i % 8], non_blocking=True)
            d = shard.sample_device(GB, BETA, u_dev, gather=True, out=outbuf)
            loss, prio = shard.td_priority_update(q, qn, qt, t_act, t_rew, t_done, t_w, N_STEP, GAMMA, d["idx"],
                                                  out_prio=o_res[:GB], out_loss=o_res[GB:])
            res_host.copy_(o_res, non_blocking=True)
            torch.cuda.current_stream().synchronize()

        for i in range(3):
            e2e_step(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            e2e_step(i)
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        h2d, d2h = 8 * GB, 4 * (GB + 1)
        api = "sharded.ShardedReplay.sample_device()/td_priority_update(): uniforms from pinned host memory, loss+priorities to host"
    e2e_s = max_over_ranks(e2e_s)
    e2e = {"value": GB * e2e_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "steps": e2e_steps, "api": api}

    # ---- CPU baseline on this box (rank 0, N=1 only) -----------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, done_steps, dt = run_cpu_baseline(10 ** 6, 3, budget_s=15.0)
        cpu = {"value": v, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": "%d steps (%.1f s) of sample(512)+update_priorities(512) on a 2^20-leaf tree with the "
                         "oracle's pure-Python port of memory.py; host has %d cores, the reference path is serial"
                         % (done_steps, dt, os.cpu_count())}

    if rank == 0:
        peak, peak_src = peaks()
        launch_bytes = GATHER_BYTES_PER_ROW * rows_this_rank  # rows this rank's gather actually moved (last step)
        achieved = launch_bytes / (gather_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong" if args.global_batch else "weak", "vs_baseline": None,
            "dtype": "f64 trees / u8 frames / f32 TD", "data": "synthetic", "config": workload_config(world),
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches_per_step * K,
            "roofline": {"bound": "hbm", "kernel": "k_gather_ldg<16>" if os.environ.get("PER_GATHER_MODE", "1") == "1" else "k_gather_tma", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": NCU_GATHER_TRAFFIC if world == 1 else None,
                         "peak_source": peak_src, "bytes_per_launch": launch_bytes, "avg_launch_us": gather_ms * 1e3,
                         "in_step": None if not (anatomy and anatomy.get("gather_grid")) else {
                             "avg_launch_us": anatomy["gather_grid"],
                             "achieved": launch_bytes / (anatomy["gather_grid"] * 1e-6) / 1e9,
                             "frac": launch_bytes / (anatomy["gather_grid"] * 1e-6) / 1e9 / peak,
                             "how": "supplementary: the same launch timed by the kernel's own GPU-timer stamps inside the chained "
                                    "step (wait passed -> last lane of the grid done); the event-bracketed figure above runs the "
                                    "kernel unchained from its neighbours"}},
            "cpu_baseline": cpu, "launch": launch_mode, "value_eager_launches": eager_value,
            "learner_step_us": {"total": 1e3 * ms / K, "gather": gather_ms * 1e3},
            "record_exchange_us": exchange_us, "step_anatomy_us": anatomy,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
