"""Parity of the sharded replay on REAL peers: one process per GPU, records exchanged through the
peer-memory mailboxes inside the sampling kernel (CUDA IPC over NVLink) and published early by the
update kernel's tail, checked against the single-tree oracle (oracle/per_oracle.py, shard_oracle.py).

Used two ways:
  * tests/test_gpu_multi.py launches it as `python tests/multi_gpu_parity.py` once per GPU (RANK /
    WORLD_SIZE / LOCAL_RANK / MASTER_* in the environment, NCCL for the test's own all-gathers);
  * bench.py --gpus N calls run_check() at start-up, so every scaling run proves parity on the GPUs
    it is about to time (a line on stderr says so).

What is checked, every step, on every rank:
  indices   the union over ranks of per_sample_sharded's output == the reference's
            _sample_proportional on ONE tree holding all leaves (bit-exact), one owner per stratum;
  weights   == memory.py:289-295 on that tree (1e-9 relative; the bar is 1e-5);
  rows      gathered bytes of the owned strata == the stored rows;
  trees     every rank's local tree == the matching subtree of the single tree rebuilt by the oracle
            from the leaves (bit-exact), after fused TD updates and after plain priority updates;
  records   travel only through the mailboxes: a stale or torn record would change the next
            sample's indices, which are compared again.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "ape-x_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)


def bits(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64)).view(np.uint64)


def run_check(rank, world, device, size=40000, batch=None, steps=5, obs_shape=(2, 8), alpha=0.6, beta=0.4, seed=0,
              early_publish=True, log=None):
    """Builds a sharded replay of `size` rows over `world` ranks, runs `steps` sample -> gather -> update
    rounds through the product path and compares each with the oracle.  Collectives of the CHECK itself
    (gathering every rank's answer) go through torch.distributed; the product path uses none."""
    import torch
    import torch.distributed as dist

    import per_native as pn
    import sharded
    from oracle import per_oracle as po
    from oracle import shard_oracle as so

    batch = batch or 512 * world
    dev = torch.device("cuda", device)
    rng = np.random.default_rng(seed)  # the same stream on every rank
    n_add = size - size // 7           # partially filled: the prefix boundary falls inside a shard
    obs = rng.integers(0, 256, size=(n_add,) + obs_shape, dtype=np.uint8)
    nxt = rng.integers(0, 256, size=(n_add,) + obs_shape, dtype=np.uint8)
    act = rng.integers(0, 18, size=n_add).astype(np.int64)
    rew = rng.standard_normal(n_add).astype(np.float32)
    done = (rng.random(n_add) < 0.05).astype(np.float32)
    prio = rng.random(n_add) + 1e-3
    t = lambda a: torch.as_tensor(a).to(dev)
    shard = sharded.ShardedReplay(size, alpha, rank, world, obs_shape=obs_shape, device=dev)
    chunk = max(3777, n_add // 97)   # uneven chunks that cross slice borders
    for lo in range(0, n_add, chunk):
        hi = min(n_add, lo + chunk)
        shard.add_batch_device(t(obs[lo:hi]), t(nxt[lo:hi]), t(act[lo:hi]), t(rew[lo:hi]), t(done[lo:hi]), t(prio[lo:hi]))
    shard.enable_peer_exchange()
    cap = po.next_pow2(size)
    cap_l = cap // world

    def gather_leaves():
        """All ranks' local leaves -> the single tree's leaf vector (device leaves are CUDA pow: <= 2 ulp from libm)."""
        mine = torch.as_tensor(shard.tree_values(pn.TREE_SUM)[cap_l:]).to(dev)
        mine_min = torch.as_tensor(shard.tree_values(pn.TREE_MIN)[cap_l:]).to(dev)
        out = [torch.empty_like(mine) for _ in range(world)]
        out_m = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(out, mine)
        dist.all_gather(out_m, mine_min)
        return torch.cat(out).cpu().numpy(), torch.cat(out_m).cpu().numpy()

    def check_trees(tag):
        ls, lm = gather_leaves()
        sv, mv = po.build_tree(ls, "sum"), po.build_tree(lm, "min")
        assert np.array_equal(bits(shard.tree_values(pn.TREE_SUM))[1:], bits(so.slice_of(sv, rank, world))[1:]), tag + ": sum tree"
        assert np.array_equal(bits(shard.tree_values(pn.TREE_MIN))[1:], bits(so.slice_of(mv, rank, world))[1:]), tag + ": min tree"
        return sv, mv

    sv, mv = check_trees("after the inserts")
    # the leaves the inserts produced are priority ** alpha within 2 ulp of libm's
    np.testing.assert_allclose(sv[cap:cap + n_add], prio ** alpha, rtol=1e-15 * 4)
    if early_publish:
        shard.enable_early_publish()
    q, qn, qt = (torch.as_tensor(rng.standard_normal((batch, 18)).astype(np.float32)).to(dev) for _ in range(3))
    for step in range(steps):
        u = rng.random(batch)
        d = shard.sample_device(batch, beta, torch.as_tensor(u).to(dev), gather=True)
        idx = d["idx"]
        all_idx = [torch.empty_like(idx) for _ in range(world)]
        dist.all_gather(all_idx, idx)
        stack = torch.stack(all_idx)
        assert bool(((stack >= 0).sum(0) == 1).all()), "every stratum must have exactly one owner"
        merged = stack.max(0).values
        want = po.descend_many(sv, po.stratified_masses(po.prefix_total(sv, n_add - 1), batch, u))
        assert np.array_equal(merged.cpu().numpy(), want), "step %d: sharded indices differ from the single tree" % step
        w = d["weights64"].clone()
        dist.all_reduce(w)  # non-owners hold 0
        np.testing.assert_allclose(w.cpu().numpy(), po.is_weights(sv, mv, want, n_add, beta), rtol=1e-9)
        mine = (idx >= 0).cpu().numpy()
        own = want[mine]
        assert np.all(own // cap_l == rank)
        assert np.array_equal(d["obs"].cpu().numpy()[mine], obs[own]) and np.array_equal(d["next_obs"].cpu().numpy()[mine], nxt[own])
        assert np.array_equal(d["act"].cpu().numpy()[mine], act[own]) and np.array_equal(d["rew"].cpu().numpy()[mine], rew[own])
        assert np.array_equal(d["done"].cpu().numpy()[mine], done[own])
        if step % 2 == 0:  # fused TD -> priority -> update over the full batch (utils.py:71-80 + memory.py:300-320)
            w32 = w.float()
            loss, pr = shard.td_priority_update(q, qn, qt, t(act[want]), t(rew[want]), t(done[want]), w32, 3, 0.99, idx)
            ol, op, _ = po.td_priorities(q.cpu().numpy(), qn.cpu().numpy(), qt.cpu().numpy(), act[want], rew[want], done[want],
                                         w32.cpu().numpy(), 3, 0.99)
            np.testing.assert_allclose(pr.cpu().numpy(), op, rtol=1e-5)
            np.testing.assert_allclose(float(loss), ol, rtol=1e-5)
            newp = pr.cpu().numpy().astype(np.float64)
        else:              # plain priorities, duplicates included (the sample itself has them)
            newp = rng.random(batch) + 1e-3
            shard.update_priorities_device(idx, t(newp))
        assert shard.status() == 0, "device status word set (exchange timeout / bad index)"
        exp_leaves = po.apply_updates(sv[cap:], want, newp ** alpha)
        sv, mv = check_trees("step %d" % step)
        np.testing.assert_allclose(sv[cap:], exp_leaves, rtol=1e-15 * 4)
    # the host-buffer forms (numpy in / numpy out on every rank, the reference's calling convention)
    u = rng.random(batch)
    h = shard.sample_host(batch, beta, u)
    want = po.descend_many(sv, po.stratified_masses(po.prefix_total(sv, n_add - 1), batch, u))
    lo, hi = h["lo"], h["hi"]
    assert np.array_equal(h["idx"], want[lo:hi]) and np.all(want[lo:hi] // cap_l == rank)
    assert hi - lo == int(np.sum(want // cap_l == rank))  # all of this rank's strata, nothing else (none at all is possible)
    if hi > lo:
        assert (lo == 0 or want[lo - 1] // cap_l < rank) and (hi == batch or want[hi] // cap_l > rank)
    np.testing.assert_allclose(h["weights"], po.is_weights(sv, mv, want[lo:hi], n_add, beta), rtol=1e-9)
    assert np.array_equal(h["obs"], obs[want[lo:hi]]) and np.array_equal(h["next_obs"], nxt[want[lo:hi]])
    assert np.array_equal(h["act"], act[want[lo:hi]]) and np.array_equal(h["rew"], rew[want[lo:hi]]) and np.array_equal(h["done"], done[want[lo:hi]])
    newp = rng.random(batch) + 1e-3
    shard.update_priorities_host(want, newp)  # every rank gets the whole batch's pairs and keeps its own
    exp_leaves = po.apply_updates(sv[cap:], want, po.pow_alpha(newp, alpha))
    sv, mv = check_trees("host update")
    assert np.array_equal(bits(sv[cap:]), bits(exp_leaves)), "host-path leaves must be bit-identical (libm pow)"
    assert shard.status() == 0
    if log:
        log("multi-GPU parity ok on rank %d of %d: %d steps of %d strata, indices bit-exact vs the single-tree oracle, "
            "shard trees == subtrees, records through the mailboxes%s\n" % (rank, world, steps, batch, " (early publish)" if early_publish else ""))
    del shard
    return True


def run_p2p_check(rank, world, device, rows_per_gpu=4096, batch=None, steps=3, obs_shape=(4, 84, 84), alpha=0.6, beta=0.4,
                  seed=21, log=None):
    """Learner-ward delivery on real peers (SURVEY.md section 8e data movement): every rank's gather stores the rows it
    owns into rank 0's batch buffers over NVLink, its strata's indices / weights likewise; rank 0 (the learner) checks
    the complete batch against the single-tree oracle and the stored rows byte for byte, writes new priorities into its
    own memory, and every rank applies its share by reading them over NVLink; the trees are compared with the oracle's."""
    import torch
    import torch.distributed as dist

    import per_native as pn
    import sharded
    from oracle import per_oracle as po
    from oracle import shard_oracle as so

    batch = batch or 256 * world
    dev = torch.device("cuda", device)
    size = rows_per_gpu * world
    rng = np.random.default_rng(seed)
    obs = rng.integers(0, 256, size=(size,) + obs_shape, dtype=np.uint8)
    nxt = rng.integers(0, 256, size=(size,) + obs_shape, dtype=np.uint8)
    act = rng.integers(0, 18, size=size).astype(np.int64)
    rew = rng.standard_normal(size).astype(np.float32)
    done = (rng.random(size) < 0.05).astype(np.float32)
    prio = rng.random(size) + 1e-3
    t = lambda a: torch.as_tensor(a).to(dev)
    shard = sharded.ShardedReplay(size, alpha, rank, world, obs_shape=obs_shape, device=dev)
    for lo in range(0, size, 1024):
        hi = min(size, lo + 1024)
        shard.add_batch_device(t(obs[lo:hi]), t(nxt[lo:hi]), t(act[lo:hi]), t(rew[lo:hi]), t(done[lo:hi]), t(prio[lo:hi]))
    shard.enable_peer_exchange()
    shard.enable_early_publish()
    lb = shard.open_learner_batch(batch, learner_rank=0)
    cap = po.next_pow2(size)
    cap_l = cap // world

    def single_tree():
        mine = torch.as_tensor(shard.tree_values(pn.TREE_SUM)[cap_l:]).to(dev)
        mine_m = torch.as_tensor(shard.tree_values(pn.TREE_MIN)[cap_l:]).to(dev)
        a, b = [torch.empty_like(mine) for _ in range(world)], [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(a, mine)
        dist.all_gather(b, mine_m)
        return po.build_tree(torch.cat(a).cpu().numpy(), "sum"), po.build_tree(torch.cat(b).cpu().numpy(), "min")

    sv, mv = single_tree()
    for step in range(steps):
        u = rng.random(batch)
        if lb.local:
            for k in ("obs", "next_obs"):
                lb[k].zero_()
            lb["idx"].fill_(-7)
        lb.fence()
        d = shard.sample_to_learner(lb, beta, torch.as_tensor(u).to(dev))
        lb.fence()  # every rank's rows, indices and weights are in the learner's memory
        want = po.descend_many(sv, po.stratified_masses(po.prefix_total(sv, size - 1), batch, u))
        if lb.local:
            assert np.array_equal(lb["idx"].cpu().numpy(), want), "indices delivered to the learner differ from the single tree's"
            np.testing.assert_allclose(lb["weights"].cpu().numpy(), po.is_weights(sv, mv, want, size, beta).astype(np.float32), rtol=1e-5)
            assert np.array_equal(lb["obs"].cpu().numpy(), obs[want]) and np.array_equal(lb["next_obs"].cpu().numpy(), nxt[want])
            assert np.array_equal(lb["act"].cpu().numpy(), act[want]) and np.array_equal(lb["rew"].cpu().numpy(), rew[want])
            assert np.array_equal(lb["done"].cpu().numpy(), done[want])
            newp = (rng.random(batch) + 1e-3).astype(np.float32)
            lb["prio"].copy_(torch.as_tensor(newp))
        else:
            newp = (rng.random(batch) + 1e-3).astype(np.float32)  # same stream on every rank: the oracle's copy
        lb.fence()  # the learner's priorities are in place
        shard.update_from_learner(lb, d["idx"])
        assert shard.status() == 0
        exp = po.apply_updates(sv[cap:], want, newp.astype(np.float64) ** alpha)
        sv, mv = single_tree()
        np.testing.assert_allclose(sv[cap:], exp, rtol=4e-15)
        assert np.array_equal(bits(shard.tree_values(pn.TREE_SUM))[1:], bits(so.slice_of(sv, rank, world))[1:])
        assert np.array_equal(bits(shard.tree_values(pn.TREE_MIN))[1:], bits(so.slice_of(mv, rank, world))[1:])
    lb.fence()
    if log:
        log("P2P delivery ok on rank %d of %d: %d steps of %d strata gathered into rank 0's buffers over NVLink (rows byte-exact, "
            "indices = single-tree oracle), priorities read back from the learner's memory, shard trees == subtrees\n" % (rank, world, steps, batch))
    del shard
    return True


if __name__ == "__main__":
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import __graft_entry__ as entry
    entry.build()
    for early in (False, True):
        run_check(rank, world, local, early_publish=early, seed=3 + int(early), log=sys.stderr.write)
    run_check(rank, world, local, size=(1 << 16) + 77, batch=4096, steps=3, seed=9, log=sys.stderr.write)
    run_p2p_check(rank, world, local, log=sys.stderr.write)
    dist.barrier()
    dist.destroy_process_group()
    print("rank", rank, "ok")
