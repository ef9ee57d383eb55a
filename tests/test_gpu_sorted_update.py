"""The sorted fast path of the one-launch priority update (update_kernels.cuh: stage 1 per CTA,
one cluster barrier, stage 2 along the borders) against the level-synchronous path and the oracle.

The native library takes the fast path when it is handed the very index buffer the last sample
wrote (ascending indices); a clone of that buffer takes the generic path.  Twin buffers with the
same contents must end up with bit-identical trees, whatever the priority distribution does to
the sample: spread out, concentrated in a few leaves (duplicates, CTAs sharing low ancestors),
everything in one CTA's range, tiny trees, batches that do not fill the cluster."""
import numpy as np
import pytest
import torch

from oracle import per_oracle as po

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import __graft_entry__ as entry
    entry.build()
    import memory
    import per_native
    return memory, per_native


def bits(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64)).view(np.uint64)


def priorities(kind, n, gen):
    if kind == "uniform":
        return torch.rand(n, dtype=torch.float64, device="cuda", generator=gen) + 1e-3
    if kind == "few_hot":  # almost all mass in 5 leaves: long runs of duplicates
        p = torch.full((n,), 1e-9, dtype=torch.float64, device="cuda")
        hot = torch.randint(0, n, (5,), device="cuda", generator=gen)
        p[hot] = 1.0 + torch.rand(5, dtype=torch.float64, device="cuda", generator=gen)
        return p
    if kind == "one_block":  # all mass inside one aligned block of 64 leaves: every CTA shares low ancestors
        p = torch.full((n,), 1e-12, dtype=torch.float64, device="cuda")
        lo = (n // 3) // 64 * 64
        w = min(64, n - lo)
        p[lo:lo + w] = torch.rand(w, dtype=torch.float64, device="cuda", generator=gen) + 0.5
        return p
    if kind == "two_ends":  # mass at both ends of the ring: borders between far-apart leaves
        p = torch.full((n,), 1e-12, dtype=torch.float64, device="cuda")
        p[:min(40, n // 2)] = 1.0
        p[-min(40, n // 2):] = 2.0
        return p
    if kind == "steps":  # geometric decay: samples crowd the front
        return torch.exp(-torch.arange(n, dtype=torch.float64, device="cuda") * (20.0 / n)) + 1e-9
    raise ValueError(kind)


@pytest.mark.parametrize("size,batch", [(1 << 16, 512), (1 << 16, 64), (1 << 16, 1000), (5000, 512), (3000, 777), (1 << 11, 256),
                                        (700, 128), (40, 16), (1 << 20, 512), (1 << 14, 1), (1 << 20, 2048), (1 << 16, 1300), (1 << 22, 1025),
                                        (1 << 12, 1100), (1 << 13, 3000), (1 << 18, 5000), (1 << 21, 9000), (1 << 22, 17000), (1 << 22, 2500), (3000000, 4096)])
@pytest.mark.parametrize("kind", ["uniform", "few_hot", "one_block", "two_ends", "steps"])
def test_sorted_path_equals_generic_path(mods, size, batch, kind):
    memory, pn = mods
    gen = torch.Generator(device="cuda").manual_seed(size * 7 + batch)
    obs = torch.zeros((size, 8), dtype=torch.uint8, device="cuda")
    z = torch.zeros(size, device="cuda")
    pr = priorities(kind, size, gen)
    fast = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    slow = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    for b in (fast, slow):
        b.add_batch_device(obs, obs, torch.zeros(size, dtype=torch.int64, device="cuda"), z, z, priorities=pr)
    rng = np.random.default_rng(batch)
    for step in range(4):
        u = rng.random(batch)
        df = fast.sample_device(batch, 0.4, uniforms=u, gather=False)
        ds = slow.sample_device(batch, 0.4, uniforms=u, gather=False)
        assert torch.equal(df["idx"], ds["idx"])
        idx_host = df["idx"].cpu().numpy()
        assert np.all(np.diff(idx_host) >= 0)  # what the fast path relies on
        newp = torch.rand(batch, dtype=torch.float64, device="cuda", generator=gen) * (10.0 if step % 2 else 1e-3) + 1e-6
        fast.update_priorities_device(df["idx"], newp)          # the sample's own buffer: sorted fast path
        slow.update_priorities_device(ds["idx"].clone(), newp)  # a copy: generic path
        for which in (pn.TREE_SUM, pn.TREE_MIN):
            tf, ts = fast.tree_values(which), slow.tree_values(which)
            assert np.array_equal(bits(tf)[1:], bits(ts)[1:]), (kind, size, batch, step, which)
        fast._raise_status()
        slow._raise_status()
    # and against the oracle's tree built from the final leaves (the invariant node = op(children))
    cap = fast.tree_values(pn.TREE_SUM).shape[0] // 2
    leaves = fast.tree_values(pn.TREE_SUM)[cap:]
    want = np.zeros(2 * cap)
    want[cap:] = leaves
    for i in range(cap - 1, 0, -1):
        want[i] = want[2 * i] + want[2 * i + 1]
    assert np.array_equal(bits(fast.tree_values(pn.TREE_SUM))[1:], bits(want)[1:])


@pytest.mark.parametrize("size,batch", [(1 << 15, 512), (1 << 15, 2048), (1 << 20, 4096), (1 << 12, 1500)])
def test_sorted_path_fused_td(mods, size, batch):
    """The fused TD prologue in front of the sorted path (the bench's call); above 1 024 items the TD launch of
    k_update_cluster followed by k_upd_dense_sorted."""
    memory, pn = mods
    gen = torch.Generator(device="cuda").manual_seed(3)
    obs = torch.zeros((size, 8), dtype=torch.uint8, device="cuda")
    z = torch.zeros(size, device="cuda")
    pr = priorities("steps", size, gen)
    fast = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    slow = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    for b in (fast, slow):
        b.add_batch_device(obs, obs, torch.zeros(size, dtype=torch.int64, device="cuda"), z, z, priorities=pr)
    rng = np.random.default_rng(1)
    for step in range(3):
        u = rng.random(batch)
        df = fast.sample_device(batch, 0.4, uniforms=u, gather=False)
        ds = slow.sample_device(batch, 0.4, uniforms=u, gather=False)
        q, qn, qt = (torch.randn(batch, 18, device="cuda", generator=gen) for _ in range(3))
        act = torch.randint(0, 18, (batch,), device="cuda", generator=gen)
        rew = torch.randn(batch, device="cuda", generator=gen)
        done = torch.zeros(batch, device="cuda")
        lf, pf, _ = fast.td_priority_update(q, qn, qt, act, rew, done, df["weights"], 3, 0.99, idx=df["idx"])
        ls, ps, _ = slow.td_priority_update(q, qn, qt, act, rew, done, ds["weights"], 3, 0.99, idx=ds["idx"].clone())
        assert torch.equal(pf, ps) and float(lf) == float(ls)
        for which in (pn.TREE_SUM, pn.TREE_MIN):
            assert np.array_equal(bits(fast.tree_values(which))[1:], bits(slow.tree_values(which))[1:])


def test_scribbled_index_buffer_is_reported(mods):
    """The fast path is taken on pointer identity with the last sample's index buffer; if the caller has
    reordered that buffer in place the kernel cannot recover, but it must say so (BAD_INDEX status)."""
    memory, pn = mods
    size, batch = 1 << 14, 512
    gen = torch.Generator(device="cuda").manual_seed(9)
    obs = torch.zeros((size, 8), dtype=torch.uint8, device="cuda")
    z = torch.zeros(size, device="cuda")
    buf = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    buf.add_batch_device(obs, obs, torch.zeros(size, dtype=torch.int64, device="cuda"), z, z, priorities=priorities("uniform", size, gen))
    d = buf.sample_device(batch, 0.4, uniforms=np.random.default_rng(0).random(batch), gather=False)
    d["idx"].copy_(d["idx"].flip(0).clone())
    buf.update_priorities_device(d["idx"], torch.rand(batch, dtype=torch.float64, device="cuda", generator=gen) + 1e-3)
    with pytest.raises(AssertionError):
        buf._raise_status()
