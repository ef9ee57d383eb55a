"""SURVEY section 8(e) data movement on real GPUs: every rank's gather kernel writes the rows it owns STRAIGHT into
the learner GPU's batch buffers (rank 0's memory, mapped into the other processes with CUDA IPC; plain stores over
NVLink, the kernels only see addresses), and rank 0 applies nothing but reads: after a barrier its buffers must hold,
for every stratum, exactly the row the owning rank holds for that index.

Checks (bit-exact): each rank also gathers its rows into a private local buffer; per-row checksums of rank 0's
peer-written buffers must equal the owners' local ones.  Then times the step [record exchange + sample + gather] with
local and with peer destinations.
Usage: torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/p2p_gather.py [--ring 16384] [--batch 512]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ape-x_b200")]
import numpy as np
import torch
import torch.distributed as dist

import __graft_entry__ as entry

ap = argparse.ArgumentParser()
ap.add_argument("--ring", type=int, default=1 << 14, help="transitions per GPU")
ap.add_argument("--batch", type=int, default=512, help="strata per GPU")
ap.add_argument("--steps", type=int, default=200)
args = ap.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
if rank == 0:
    entry.build()
dist.barrier()
import per_native as pn
import sharded

OBS = (4, 84, 84)
OB = int(np.prod(OBS))
GB = args.batch * world
size = world * args.ring
shard = sharded.ShardedReplay(size, 0.6, rank, world, obs_shape=OBS)
gen = torch.Generator(device=dev).manual_seed(100 + rank)
for lo in range(0, args.ring, 4096):
    m = min(4096, args.ring - lo)
    obs = torch.randint(0, 256, (m,) + OBS, dtype=torch.uint8, device=dev, generator=gen)
    nxt = torch.randint(0, 256, (m,) + OBS, dtype=torch.uint8, device=dev, generator=gen)
    pn.check(pn.lib.per_add_batch(shard._h, pn.ptr(obs), pn.ptr(nxt), pn.ptr(torch.randint(0, 18, (m,), device=dev, generator=gen)),
                                  pn.ptr(torch.randn(m, device=dev, generator=gen)), pn.ptr(torch.zeros(m, device=dev)), None,
                                  pn.ptr(torch.rand(m, dtype=torch.float64, device=dev, generator=gen) + 1e-3), None, 0.0, m,
                                  pn.stream_ptr()))
shard.set_global_len(size, 0)
shard.enable_peer_exchange()

# the learner's batch buffers live on rank 0; everybody else maps them
nbytes = {"obs": GB * OB, "next_obs": GB * OB, "act": GB * 8, "rew": GB * 4, "done": GB * 4}
if rank == 0:
    own, handles = {}, {}
    for k, nb in nbytes.items():
        own[k], handles[k] = sharded.PeerBuffer.alloc(nb, local)
    box = [handles]
else:
    box = [None]
dist.broadcast_object_list(box, src=0)
learner = own if rank == 0 else {k: sharded.PeerBuffer.open(box[0][k], nbytes[k], local) for k in nbytes}
private = {"obs": torch.zeros((GB,) + OBS, dtype=torch.uint8, device=dev), "next_obs": torch.zeros((GB,) + OBS, dtype=torch.uint8, device=dev),
           "act": torch.zeros(GB, dtype=torch.int64, device=dev), "rew": torch.zeros(GB, device=dev), "done": torch.zeros(GB, device=dev)}
cgen = torch.Generator(device=dev).manual_seed(7)  # same uniforms on every rank
u_all = torch.rand(64, GB, dtype=torch.float64, device=dev, generator=cgen)


def step(i, dest):
    d = shard.sample_device(GB, 0.4, u_all[i % 64], gather=False)
    shard.gather_device(d["idx"], dest)
    return d["idx"]


# ---- correctness: one step into both destinations, checksums per row ----
if rank == 0:
    for k in ("obs", "next_obs"):
        learner[k].view(torch.uint8, (GB * OB,)).zero_()
torch.cuda.synchronize()
dist.barrier()
idx = step(0, learner)
torch.cuda.synchronize()
dist.barrier()  # every rank's rows have landed in rank 0's memory
step_idx = idx.clone()
# the same draw again into the private buffers (nothing changed the trees in between)
idx2 = step(0, private)
assert torch.equal(step_idx, idx2)
mine = idx2 >= 0
sums = torch.zeros(GB, 2, dtype=torch.int64, device=dev)
sums[mine, 0] = private["obs"][mine].view(int(mine.sum()), -1).sum(1, dtype=torch.int64)
sums[mine, 1] = private["next_obs"][mine].view(int(mine.sum()), -1).sum(1, dtype=torch.int64)
own_mask = mine.to(torch.int64)
dist.all_reduce(sums)      # each stratum has exactly one owner
dist.all_reduce(own_mask)
assert int(own_mask.min()) == 1 and int(own_mask.max()) == 1, "every stratum must have exactly one owner"
if rank == 0:
    got = torch.stack([learner["obs"].view(torch.uint8, (GB, OB)).sum(1, dtype=torch.int64),
                       learner["next_obs"].view(torch.uint8, (GB, OB)).sum(1, dtype=torch.int64)], 1)
    assert torch.equal(got, sums), "rows written over NVLink differ from the owners' rows"
    # and byte for byte for rank 0's own strata
    assert torch.equal(learner["obs"].view(torch.uint8, (GB,) + OBS)[mine], private["obs"][mine])
    print("P2P gather: %d strata from %d ranks landed in rank 0's buffers, checksums of all rows match their owners" % (GB, world))


# ---- timing ----
def timed(dest):
    for i in range(10):
        step(i, dest)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(i, dest)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) * 1e3 / args.steps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


t_local = timed(private)
t_peer = timed(learner)
if rank == 0:
    rows_remote = GB - int(mine.sum())
    print("sample + gather per step (plain launches, max over ranks): local buffers %.1f us | learner's buffers on rank 0 %.1f us "
          "(%d of %d rows cross NVLink, %.1f MB)" % (t_local, t_peer, rows_remote, GB, rows_remote * 2 * OB / 1e6))
assert shard.status() == 0
dist.barrier()
dist.destroy_process_group()
