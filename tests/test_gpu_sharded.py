"""Sharded replay on the GPU.  `world` ranks are emulated in ONE process on ONE device: each
rank is a ShardedReplay (its own native handle, trees and ring), the all-gather is a
concatenation, and the ranks run one after the other (they never wait on one another inside a
kernel, so sequential execution is equivalent).  Checked against a single unsharded buffer
holding the same data (which test_gpu_parity.py ties to the reference) and against the
oracle: indices bit-exact, every shard tree equal to the matching subtree of the single tree."""
import numpy as np
import pytest
import torch

from oracle import per_oracle as po
from oracle import shard_oracle as so

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import __graft_entry__ as entry
    entry.build()
    import memory
    import per_native
    import sharded
    return memory, sharded, per_native


def bits(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64)).view(np.uint64)


def make_data(n, obs_shape, seed):
    gen = torch.Generator(device="cuda").manual_seed(seed)
    obs = torch.randint(0, 256, (n,) + obs_shape, dtype=torch.uint8, device="cuda", generator=gen)
    nxt = torch.randint(0, 256, (n,) + obs_shape, dtype=torch.uint8, device="cuda", generator=gen)
    act = torch.randint(0, 18, (n,), dtype=torch.int64, device="cuda", generator=gen)
    rew = torch.randn(n, device="cuda", generator=gen)
    done = (torch.rand(n, device="cuda", generator=gen) < 0.05).float()
    prio = torch.rand(n, dtype=torch.float64, device="cuda", generator=gen) + 1e-3
    return obs, nxt, act, rew, done, prio


@pytest.mark.parametrize("size,world,n_add,batch", [
    (4096, 2, 4096, 256), (5000, 4, 5000, 512), (5000, 8, 3001, 512), (100000, 8, 100000, 4096),
    (5000, 8, 7300, 512),      # wraps the global ring: early slots overwritten
    (1 << 16, 16, 60000, 512),
    (1 << 16, 2, 1 << 16, 8192), (70000, 4, 70000, 20000),  # thousands of own strata: one thread per stratum (k_sample_sharded<true>)
])
def test_sharded_equals_single_tree(mods, size, world, n_add, batch):
    memory, sharded, pn = mods
    obs_shape = (2, 8)
    data = make_data(n_add, obs_shape, size + world)
    single = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    store = {}
    shards = [sharded.ShardedReplay(size, 0.6, r, world, obs_shape=obs_shape,
                                    exchange=lambda rec, r=r: store["all"]) for r in range(world)]
    for lo in range(0, n_add, 1777):  # uneven chunks crossing slice borders
        hi = min(n_add, lo + 1777)
        part = [t[lo:hi].contiguous() for t in data]
        single.add_batch_device(part[0], part[1], part[2], part[3], part[4], priorities=part[5])
        for s in shards:
            s.add_batch_device(*part)
    assert all(len(s) == len(single) for s in shards)
    cap_l = shards[0].tree_capacity_local
    rng = np.random.default_rng(7)

    def exchange_all():
        recs = []
        for s in shards:
            left_by_update = s._rec.clone() if s._rec_fresh else None
            pn.check(pn.lib.per_shard_record(s._h, len(s), pn.ptr(s._rec), pn.stream_ptr()))
            if left_by_update is not None:  # the update kernel's tail must have left the same record
                assert np.array_equal(bits(left_by_update.cpu().numpy()), bits(s._rec.cpu().numpy()))
            recs.append(s._rec.clone())
        store["all"] = torch.cat(recs)

    for step in range(3):
        sv, mv = single._it_sum._value, single._it_min._value
        for r, s in enumerate(shards):  # every shard tree is the matching subtree of the single tree
            assert np.array_equal(bits(s.tree_values(pn.TREE_SUM))[1:], bits(so.slice_of(sv, r, world))[1:])
            assert np.array_equal(bits(s.tree_values(pn.TREE_MIN))[1:], bits(so.slice_of(mv, r, world))[1:])
        u = rng.random(batch)
        want = single.sample_device(batch, 0.4, uniforms=u)
        exchange_all()
        assert np.array_equal(bits(store["all"].cpu().numpy().reshape(world, 4)[:, :3]),
                              bits(np.stack([so.shard_record(so.slice_of(sv, r, world), so.slice_of(mv, r, world), r,
                                                             len(single)) for r in range(world)])[:, :3]))
        merged = torch.full((batch,), -1, dtype=torch.int64, device="cuda")
        w_merged = torch.zeros(batch, dtype=torch.float64, device="cuda")
        out = None
        for s in shards:
            d = s.sample_device(batch, 0.4, u, gather=False)
            mine = d["idx"] >= 0
            assert not bool((merged[mine] >= 0).any())  # one owner per stratum
            merged[mine] = d["idx"][mine]
            w_merged[mine] = d["weights64"][mine]
            # all ranks gather into the SAME full-batch buffers, as they would into a learner's peer memory
            out = s.gather_device(d["idx"], out)
            assert s.status() == 0
        assert torch.equal(merged, want["idx"])
        assert np.array_equal(merged.cpu().numpy(), po.descend_many(sv, po.stratified_masses(
            po.prefix_total(sv, len(single) - 1), batch, u)))
        np.testing.assert_allclose(w_merged.cpu().numpy(), want["weights64"].cpu().numpy(), rtol=1e-12)
        for k in ("obs", "next_obs", "act", "rew", "done"):
            assert torch.equal(out[k], want[k]), k
        # priorities come back for the full batch; each rank applies its own strata
        newp = torch.rand(batch, device="cuda") + 1e-3
        single.update_priorities_device(want["idx"], newp)
        for s in shards:
            d_idx = torch.where((merged // cap_l) == s.rank, merged, torch.full_like(merged, -1))
            s.update_priorities_device(d_idx, newp)
            assert s.status() == 0


@pytest.mark.parametrize("size,world,batch", [(1 << 14, 4, 512), (1 << 15, 2, 4096), (1 << 16, 2, 3000)])
def test_sharded_fused_td_update(mods, size, world, batch):
    """The fused TD update of the strata a shard has just drawn: k_update_cluster alone, or (more than 1 024 own strata)
    as the TD launch in front of k_upd_dense_sorted, which then visits the owned range only."""
    memory, sharded, pn = mods
    data = make_data(size, (2, 8), 5)
    single = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    store = {}
    shards = [sharded.ShardedReplay(size, 0.6, r, world, obs_shape=(2, 8), exchange=lambda rec: store["all"])
              for r in range(world)]
    single.add_batch_device(*data[:5], priorities=data[5])
    for s in shards:
        s.add_batch_device(*data)
    u = np.random.default_rng(3).random(batch)
    want = single.sample_device(batch, 0.4, uniforms=u)
    recs = []
    for s in shards:
        pn.check(pn.lib.per_shard_record(s._h, len(s), pn.ptr(s._rec), pn.stream_ptr()))
        recs.append(s._rec.clone())
    store["all"] = torch.cat(recs)
    q, qn, qt = (torch.randn(batch, 18, device="cuda") for _ in range(3))
    loss0, prio0, _ = single.td_priority_update(q, qn, qt, want["act"], want["rew"], want["done"], want["weights"], 3, 0.99,
                                                idx=want["idx"])
    for s in shards:
        d = s.sample_device(batch, 0.4, u, gather=False)
        loss, prio = s.td_priority_update(q, qn, qt, want["act"], want["rew"], want["done"], want["weights"], 3, 0.99, d["idx"])
        assert torch.equal(prio, prio0) and float(loss) == float(loss0)
    sv, mv = single._it_sum._value, single._it_min._value
    for r, s in enumerate(shards):
        assert s.status() == 0
        assert np.array_equal(bits(s.tree_values(pn.TREE_SUM))[1:], bits(so.slice_of(sv, r, world))[1:])
        assert np.array_equal(bits(s.tree_values(pn.TREE_MIN))[1:], bits(so.slice_of(mv, r, world))[1:])


@pytest.mark.parametrize("size,world,batch", [(6000, 2, 512), (1 << 15, 2, 4096), (1 << 13, 2, 3000)])
def test_early_publish_from_bound_updates(mods, size, world, batch):
    """Early publish with every update applied to the sample's own index buffer (the learner loop): the record of the
    next sample comes from the tail of k_update_cluster (the sorted climb) or, above 1 024 own strata, of
    k_upd_dense_sorted."""
    memory, sharded, pn = mods
    data = make_data(size, (2, 8), 23 + world)
    single = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    single.add_batch_device(*data[:5], priorities=data[5])
    shards = [sharded.ShardedReplay(size, 0.6, r, world, obs_shape=(2, 8), exchange=None) for r in range(world)]
    for s in shards:
        s.add_batch_device(*data)
    sharded.ShardedReplay.enable_local_exchange(shards)
    for s in shards:
        s.enable_early_publish()
        s.publish_record()
    rng = np.random.default_rng(9)
    for step in range(4):
        u = rng.random(batch)
        want = single.sample_device(batch, 0.4, uniforms=u, gather=False)
        drawn = [s.sample_device(batch, 0.4, u, gather=False) for s in shards]
        merged = torch.full((batch,), -1, dtype=torch.int64, device="cuda")
        for d in drawn:
            mine = d["idx"] >= 0
            assert not bool((merged[mine] >= 0).any())
            merged[mine] = d["idx"][mine]
        assert torch.equal(merged, want["idx"]), step
        newp = torch.rand(batch, device="cuda") + 1e-3
        single.update_priorities_device(want["idx"], newp)
        for s, d in zip(shards, drawn):
            s.update_priorities_device(d["idx"], newp)  # bound: the owned range of the sample's buffer
            assert s.status() == 0
        sv, mv = single._it_sum._value, single._it_min._value
        for r, s in enumerate(shards):
            assert np.array_equal(bits(s.tree_values(pn.TREE_SUM))[1:], bits(so.slice_of(sv, r, world))[1:])
            assert np.array_equal(bits(s.tree_values(pn.TREE_MIN))[1:], bits(so.slice_of(mv, r, world))[1:])


@pytest.mark.parametrize("size,world,batch", [(6000, 2, 512), (1 << 15, 2, 4096)])
def test_early_publish_from_fused_td_updates(mods, size, world, batch):
    """The learner loop on shards with early publish: sample, fused TD update of the sample's own buffer.  Above 1 024 own
    strata the update is k_update_cluster as the TD launch + k_upd_dense_sorted, whose tail publishes the next record."""
    memory, sharded, pn = mods
    data = make_data(size, (2, 8), 31 + world)
    single = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    single.add_batch_device(*data[:5], priorities=data[5])
    shards = [sharded.ShardedReplay(size, 0.6, r, world, obs_shape=(2, 8), exchange=None) for r in range(world)]
    for s in shards:
        s.add_batch_device(*data)
    sharded.ShardedReplay.enable_local_exchange(shards)
    for s in shards:
        s.enable_early_publish()
        s.publish_record()
    rng = np.random.default_rng(13)
    gen = torch.Generator(device="cuda").manual_seed(2)
    for step in range(3):
        u = rng.random(batch)
        want = single.sample_device(batch, 0.4, uniforms=u)
        drawn = [s.sample_device(batch, 0.4, u, gather=False) for s in shards]
        merged = torch.full((batch,), -1, dtype=torch.int64, device="cuda")
        for d in drawn:
            mine = d["idx"] >= 0
            merged[mine] = d["idx"][mine]
        assert torch.equal(merged, want["idx"]), step
        q, qn, qt = (torch.randn(batch, 18, device="cuda", generator=gen) for _ in range(3))
        loss0, prio0, _ = single.td_priority_update(q, qn, qt, want["act"], want["rew"], want["done"], want["weights"], 3, 0.99,
                                                    idx=want["idx"])
        for s, d in zip(shards, drawn):
            loss, prio = s.td_priority_update(q, qn, qt, want["act"], want["rew"], want["done"], want["weights"], 3, 0.99, d["idx"])
            assert torch.equal(prio, prio0) and float(loss) == float(loss0)
            assert s.status() == 0
        sv, mv = single._it_sum._value, single._it_min._value
        for r, s in enumerate(shards):
            assert np.array_equal(bits(s.tree_values(pn.TREE_SUM))[1:], bits(so.slice_of(sv, r, world))[1:])
            assert np.array_equal(bits(s.tree_values(pn.TREE_MIN))[1:], bits(so.slice_of(mv, r, world))[1:])


@pytest.mark.parametrize("world,early", [(2, False), (8, False), (2, True), (8, True)])
def test_peer_mailbox_exchange_matches_collective(mods, world, early):
    """The record exchange through mailboxes (per_shard_attach_exchange) must give the same
    indices as the all-gathered records.  Ranks share one GPU here, so every rank publishes with
    per_shard_publish before any of them samples (on separate GPUs the sampling kernel does both).
    early: the update kernels publish the next sample's record themselves (per_tune early_publish)."""
    memory, sharded, pn = mods
    size, batch = 6000, 512
    data = make_data(size, (2, 8), 11 + world)
    single = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    single.add_batch_device(*data[:5], priorities=data[5])
    shards = [sharded.ShardedReplay(size, 0.6, r, world, obs_shape=(2, 8), exchange=None) for r in range(world)]
    for s in shards:
        s.add_batch_device(*data)
    sharded.ShardedReplay.enable_local_exchange(shards)
    if early:
        for s in shards:
            s.enable_early_publish()
    rng = np.random.default_rng(5)
    for step in range(4):  # several epochs: both mailbox parities, flags from earlier epochs left behind
        u = rng.random(batch)
        want = single.sample_device(batch, 0.4, uniforms=u, gather=False)
        if not early or step == 0:
            for s in shards:
                s.publish_record()
        merged = torch.full((batch,), -1, dtype=torch.int64, device="cuda")
        for s in shards:
            d = s.sample_device(batch, 0.4, u, gather=False)
            mine = d["idx"] >= 0
            merged[mine] = d["idx"][mine]
            assert s.status() == 0  # in particular no exchange timeout
        assert torch.equal(merged, want["idx"])
        newp = torch.rand(batch, device="cuda") + 1e-3
        single.update_priorities_device(want["idx"], newp)
        for s in shards:
            own = torch.where((merged // s.tree_capacity_local) == s.rank, merged, torch.full_like(merged, -1))
            s.update_priorities_device(own, newp)
    if early:  # the contract is enforced: nothing may touch the tree between an update and the next sample
        s = shards[0]
        with pytest.raises(Exception, match="early publish"):
            s.update_priorities_device(own, newp)
        with pytest.raises(Exception, match="early publish"):
            s.add_batch_device(*[t[:8].contiguous() for t in data])
        with pytest.raises(Exception, match="early publish"):
            s.enable_early_publish(False)
