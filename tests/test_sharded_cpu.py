"""Sharded-replay protocol on the CPU: geometry / routing host logic, the claim that slices +
records reproduce the single tree bit for bit (oracle only), and the record exchange over a
real 2-process `gloo` group."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT
from oracle import per_oracle as po
from oracle import shard_oracle as so


def geometry():
    import sharded
    return sharded.ShardGeometry


def test_geometry_and_routing():
    G = geometry()
    g = G(100000, 8)
    assert g.cap == 131072 and g.cap_local == 16384
    assert [g.ring_local(r) for r in range(8)] == [16384] * 6 + [1696, 0]
    assert g.owner(16383) == 0 and g.owner(16384) == 1 and g.owner(99999) == 6
    # a batch that crosses two slice borders and wraps the ring
    segs = g.route(99990, 30)
    assert segs == [(6, 0, 10, 99990 - 6 * 16384), (0, 10, 30, 0)]
    segs = g.route(16380, 16390)
    assert [(r, hi - lo) for r, lo, hi, _ in segs] == [(0, 4), (1, 16384), (2, 2)]
    # every slot is routed exactly once over a full lap
    seen = np.zeros(100000, dtype=int)
    cur = 0
    for n in (7, 50, 4096, 33333, 62514):
        for r, lo, hi, local in g.route(cur, n):
            seen[r * g.cap_local + local: r * g.cap_local + local + (hi - lo)] += 1
        cur = (cur + n) % g.size
    assert cur == 0 and np.all(seen == 1)
    with pytest.raises(AssertionError):
        G(100, 3)


@pytest.mark.parametrize("size,world,fill", [(4096, 2, 4096), (4096, 8, 4096), (5000, 4, 5000), (5000, 8, 3001),
                                             (100000, 8, 100000), (1 << 14, 16, 9999)])
def test_slices_reproduce_single_tree(size, world, fill):
    """Oracle-only proof of the protocol: for every rank, the sharded sampler restated in
    oracle/shard_oracle.py yields exactly the single-tree indices on that rank's strata."""
    rng = np.random.default_rng(size + world)
    cap = po.next_pow2(size)
    leaves = np.zeros(cap)
    leaves[:fill] = np.abs(rng.standard_normal(fill)) + 1e-6
    lm = np.where(leaves > 0, leaves, np.inf)
    sv, mv = po.build_tree(leaves, "sum"), po.build_tree(lm, "min")
    batch = 512
    u = rng.random(batch)
    want = po.descend_many(sv, po.stratified_masses(po.prefix_total(sv, fill - 1), batch, u))
    slices = [(so.slice_of(sv, r, world), so.slice_of(mv, r, world)) for r in range(world)]
    recs = np.stack([so.shard_record(s, m, r, fill) for r, (s, m) in enumerate(slices)])
    assert so.top_tree(recs)[1].view(np.uint64) == sv[1].view(np.uint64)
    assert np.float64(so.global_mass_total(recs, cap // world, fill)).view(np.uint64) == \
        np.float64(po.prefix_total(sv, fill - 1)).view(np.uint64)
    got = np.full(batch, -1, dtype=np.int64)
    for r, (s, _m) in enumerate(slices):
        part = so.sharded_sample(recs, s, r, u, batch, fill)
        assert np.all(got[part >= 0] == -1)  # every stratum has exactly one owner
        got[part >= 0] = part[part >= 0]
    assert np.array_equal(got, want)


WORKER = r"""
import os, sys
sys.path[:0] = [{root!r}, os.path.join({root!r}, "ape-x_b200")]
import numpy as np, torch, torch.distributed as dist
from oracle import per_oracle as po, shard_oracle as so
import sharded
dist.init_process_group("gloo", rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, world = dist.get_rank(), dist.get_world_size()
size, fill, batch = 6000, 5123, 256
rng = np.random.default_rng(0)                       # same stream on both ranks
cap = po.next_pow2(size)
leaves = np.zeros(cap); leaves[:fill] = np.abs(rng.standard_normal(fill)) + 1e-6
sv = po.build_tree(leaves, "sum"); mv = po.build_tree(np.where(leaves > 0, leaves, np.inf), "min")
u = rng.random(batch)
mine_s, mine_m = so.slice_of(sv, rank, world), so.slice_of(mv, rank, world)   # this rank only sees its slice
rec = torch.from_numpy(so.shard_record(mine_s, mine_m, rank, fill))
recs = sharded.exchange_records(rec).numpy().reshape(world, 4)                 # the one collective, over gloo
part = so.sharded_sample(recs, mine_s, rank, u, batch, fill)
allp = [torch.empty(batch, dtype=torch.int64) for _ in range(world)]
dist.all_gather(allp, torch.from_numpy(part))
merged = torch.stack(allp).max(0).values.numpy()
want = po.descend_many(sv, po.stratified_masses(po.prefix_total(sv, fill - 1), batch, u))
assert np.array_equal(merged, want), "sharded indices differ from the single tree"
assert sum(int((p >= 0).sum()) for p in allp) == batch
dist.destroy_process_group()
print("rank", rank, "ok")
"""


def test_record_exchange_gloo_world2(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(root=ROOT))
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29531", WORLD_SIZE="2")
    procs = [subprocess.Popen([sys.executable, str(script)], env=dict(env, RANK=str(r)), stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, "rank %d failed:\n%s" % (r, o[-2000:])
        assert "ok" in o
