"""BASELINE configs[4] at 1/2/4/8 GPUs: the sum/min-tree kernels of a replay of C transitions cut into world slices
(tree-only rows: 16-byte observations), global stratified batch B.  One step = priority update of the sample just drawn
(every rank applies the strata it owns) + the next sample (record exchange through the peer mailboxes inside the kernel,
early publish from the update's tail); K steps are replayed as CUDA graphs, timed on the device, max over ranks.
Reported: microseconds per step, M transitions/s (B per step, whole job) and the algorithmic GB/s of SURVEY 8(d) for the
pair (sample B (8 L + 20) + update 2 [8 B + 24 sum_l min(B, 2^l)] with L = log2 C of the WHOLE replay's tree).
world == 1 runs the unsharded buffer through the same loop.
Usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/tree_sweep_multi.py
       [--caps 16 20 24 26] [--batches 512 4096 65536]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ape-x_b200")]
import torch
import torch.distributed as dist

import __graft_entry__ as entry

ap = argparse.ArgumentParser()
ap.add_argument("--caps", type=int, nargs="+", default=[16, 20, 24, 26])
ap.add_argument("--batches", type=int, nargs="+", default=[512, 4096, 65536])
ap.add_argument("--steps", type=int, default=20, help="steps per graph")
ap.add_argument("--reps", type=int, default=10)
args = ap.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
if rank == 0:
    entry.build()
if world > 1:
    dist.barrier()
import memory
import per_native as pn
import sharded

lib, P = pn.lib, pn.ptr
ALPHA, BETA = 0.6, 0.4


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


keep = []
if rank == 0:
    print("N=%d  log2C   batch | step us  M trans/s  GB/s (8d)" % world, flush=True)
for lc in args.caps:
    cap = 1 << lc
    if cap % world or cap // world < 64:
        continue
    ring = cap // world
    gen = torch.Generator(device=dev).manual_seed(100 + rank)
    if world == 1:
        buf = memory.CustomPrioritizedReplayBuffer(cap, ALPHA)
    else:
        buf = sharded.ShardedReplay(cap, ALPHA, rank, world, obs_shape=(16,))
        h = buf._h
    chunk = 1 << 16
    for lo in range(0, ring, chunk):
        m = min(chunk, ring - lo)
        o = torch.zeros((m, 16), dtype=torch.uint8, device=dev)
        z = torch.zeros(m, device=dev)
        pr = torch.rand(m, dtype=torch.float64, device=dev, generator=gen) + 1e-3
        if world == 1:
            buf.add_batch_device(o, o, torch.zeros(m, dtype=torch.int64, device=dev), z, z, priorities=pr)
        else:
            pn.check(lib.per_add_batch(h, P(o), P(o), P(torch.zeros(m, dtype=torch.int64, device=dev)), P(z), P(z), None, P(pr),
                                       None, 0.0, m, pn.stream_ptr()))
    if world == 1:
        h = buf._h
    else:
        buf.set_global_len(cap, 0)
        buf.enable_peer_exchange()
        buf.enable_early_publish()
    torch.cuda.synchronize()
    cgen = torch.Generator(device=dev).manual_seed(7)  # the same on every rank
    for B in args.batches:
        if B > cap:
            continue
        n_u = 8
        u = torch.rand(n_u, B, dtype=torch.float64, device=dev, generator=cgen)
        newp = torch.rand(B, device=dev, generator=cgen) + 1e-3
        idx = torch.empty(B, dtype=torch.int64, device=dev)
        w32 = torch.empty(B, device=dev)
        u_ptrs = [pn.C.c_void_p(u[i].data_ptr()) for i in range(n_u)]

        def sample(i, sps):
            if world == 1:
                pn.check(lib.per_sample_indices(h, u_ptrs[i % n_u], B, BETA, P(idx), None, P(w32), sps))
            else:
                pn.check(lib.per_sample_sharded(h, None, u_ptrs[i % n_u], B, BETA, cap, P(idx), None, P(w32), sps))

        def step(i, sps):  # update of the batch in hand, then the next sample
            pn.check(lib.per_update_priorities(h, P(idx), P(newp), pn.PRIO_F32, B, sps))
            sample(i, sps)

        sp = pn.stream_ptr()
        sample(0, sp)  # (no update in front: this sample publishes its record itself; afterwards the updates do)
        for i in range(3):
            step(i, sp)
        barrier()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            sps = pn.stream_ptr(side)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                for i in range(args.steps):
                    step(i, sps)
        torch.cuda.current_stream().wait_stream(side)
        g.replay()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(200000)
        e0.record()
        for _ in range(args.reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (args.reps * args.steps)
        if world > 1:
            t = torch.tensor([us], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            us = float(t.item())
        status = pn.C.c_int64(0)
        pn.check(lib.per_status(h, pn.C.byref(status), sp))
        assert status.value == 0, "status %d at 2^%d x %d" % (status.value, lc, B)
        bytes_pair = B * (8 * lc + 20) + 2 * (8 * B + 24 * sum(min(B, 1 << l) for l in range(lc)))
        if rank == 0:
            print("N=%d  %5d %7d | %7.2f %10.1f %9.1f" % (world, lc, B, us, B / us, bytes_pair / us / 1e3), flush=True)
        barrier()
    keep.append(buf)  # (peers hold CUDA-IPC mappings of its mailbox: released together at exit)
    barrier()
if world > 1:
    dist.destroy_process_group()
