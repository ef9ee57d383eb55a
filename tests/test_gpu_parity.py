"""GPU parity tests: the CUDA path (called through the C ABI by the drop-in classes of
ape-x_b200/memory.py) against the oracle (oracle/per_oracle.py) and against the golden
fixtures generated from the real reference (tests/golden/*.npz).

Bars: tree nodes and sampled indices bit-exact; importance weights, returns, TD errors and
priorities within 1e-5 relative (BASELINE.json north_star); gathered bytes identical.
"""
import hashlib
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import per_oracle as po

pytestmark = pytest.mark.gpu

RTOL = 1e-5


@pytest.fixture(scope="module")
def memory():
    import __graft_entry__ as entry
    entry.build()
    import memory as m
    return m


def load(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def bits(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64)).view(np.uint64)


def sha(values):
    return np.frombuffer(hashlib.sha256(np.asarray(values, dtype=np.float64).tobytes()).digest(), dtype=np.uint8)


class Draws:
    """Replays a uniform vector through memory.random.random (the reference draws from the
    global `random` module, memory.py:247)."""

    def __init__(self, memory_mod, monkeypatch):
        self.q = []
        monkeypatch.setattr(memory_mod.random, "random", lambda: self.q.pop())

    def load(self, u):
        self.q = list(u)[::-1]


# ------------------------------------------------------------------------------- segment trees
def test_segment_trees_golden(memory):
    g = load("tree_ops.npz")
    cap = int(g["cap"])
    st, mt = memory.SumSegmentTree(cap), memory.MinSegmentTree(cap)
    # all 200 point sets as ONE batch: duplicates must resolve like sequential execution
    st.set_batch(g["set_idx"], g["set_val"])
    mt.set_batch(g["set_idx"], g["set_val"])
    assert np.array_equal(bits(st._value), bits(g["sum_nodes"]))
    assert np.array_equal(bits(mt._value), bits(g["min_nodes"]))
    for (a, b), s, m in zip(g["ranges"][:120], g["range_sum"], g["range_min"]):
        assert bits(st.sum(int(a), int(b))) == bits(s)
        assert bits(mt.min(int(a), int(b))) == bits(m)
    for (a, b), s in zip(g["neg_ranges"], g["neg_sum"]):
        assert bits(st.sum(int(a), int(b))) == bits(s)
    assert bits(st.sum()) == bits(g["full_sum"]) and bits(mt.min()) == bits(g["full_min"])
    assert list(st.find_prefixsum_idx_batch(g["masses"])) == list(g["found"])
    assert list(st.find_prefixsum_idx_batch(g["tie_masses"])) == list(g["tie_found"])  # strict > (Q4)
    assert st.find_prefixsum_idx(float(g["masses"][0])) == int(g["found"][0])
    assert st[int(g["set_idx"][-1])] == float(g["set_val"][-1])


def test_segment_tree_scalar_setitem_and_asserts(memory):
    st = memory.SumSegmentTree(16)
    ref = po.HeapTree(16, "sum")
    for k, v in [(3, 1.5), (3, 0.25), (15, 2.0), (0, 1e-9)]:
        st[k] = v
        ref.set(k, v)
    assert np.array_equal(bits(st._value)[1:], bits(ref.v)[1:])
    with pytest.raises(AssertionError):
        memory.SumSegmentTree(24)
    with pytest.raises(AssertionError):
        st[16]
    with pytest.raises(AssertionError):
        st.find_prefixsum_idx(1e9)  # memory.py:121
    with pytest.raises(AssertionError):
        st.set_batch([99], [1.0])


@pytest.mark.parametrize("cap,n", [(1 << 12, 700), (1 << 16, 4096), (1 << 16, 5000), (1 << 20, 70000), (1 << 22, 300000)])
def test_tree_batched_update_vs_oracle(memory, cap, n):
    """Random batches with duplicates through both update forms (one-CTA for n <= 4096, the
    multi-launch form above), applied twice, against the oracle's sequential scatter + rebuild."""
    rng = np.random.default_rng(cap + n)
    st, mt = memory.SumSegmentTree(cap), memory.MinSegmentTree(cap)
    leaves_s, leaves_m = np.zeros(cap), np.full(cap, np.inf)
    for rnd in range(2):
        idx = rng.integers(0, cap, size=n)
        idx[: n // 10] = idx[n // 2: n // 2 + n // 10]  # forced duplicates
        val = np.abs(rng.standard_normal(n)) + 1e-9
        st.set_batch(idx, val)
        mt.set_batch(idx, val)
        leaves_s = po.apply_updates(leaves_s, idx, val)
        leaves_m = po.apply_updates(leaves_m, idx, val)
    assert np.array_equal(bits(st._value)[1:], bits(po.build_tree(leaves_s, "sum"))[1:])
    assert np.array_equal(bits(mt._value)[1:], bits(po.build_tree(leaves_m, "min"))[1:])
    sv = st._value
    masses = rng.random(3000) * sv[1]
    assert np.array_equal(st.find_prefixsum_idx_batch(masses), po.descend_many(sv, masses))
    for n_leaves in (1, 2, 3, cap // 2, cap // 2 + 1, cap - 1, cap, int(rng.integers(1, cap))):
        assert bits(st.sum(0, n_leaves)) == bits(po.prefix_total(sv, n_leaves))


# ------------------------------------------------------------------- replay buffers vs golden
def replay_rows(g, tag, variant):
    for i in range(g[tag + "_obs"].shape[0]):
        row = [g[tag + "_obs"][i], int(g[tag + "_act"][i]), float(g[tag + "_rew"][i]), g[tag + "_next"][i],
               float(g[tag + "_done"][i])]
        if variant == "apex":
            row.append(float(g[tag + "_extra"][i]))
        elif variant == "aql":
            row.append(g[tag + "_extra"][i])
        yield row


@pytest.mark.parametrize("variant,fname", [("per", "per_small.npz"), ("apex", "per_apex.npz"), ("aql", "per_aql.npz")])
def test_replay_golden_sequences(memory, monkeypatch, variant, fname):
    """add / sample / update_priorities interleavings recorded from the reference, incl. ring
    wrap-around, unsorted update indices with duplicates, and the AQL a_mu column."""
    g = load(fname)
    cls = {"per": memory.PrioritizedReplayBuffer, "apex": memory.CustomPrioritizedReplayBuffer,
           "aql": memory.CustomPrioritizedReplayBuffer_AQL}[variant]
    buf = cls(int(g["size"]), float(g["alpha"]))
    draws = Draws(memory, monkeypatch)
    beta, batch = float(g["beta"]), int(g["batch"])
    for row in replay_rows(g, "fill", variant):
        buf.add(*row)
    assert np.array_equal(bits(buf._it_sum._value), bits(g["after_fill_sum"]))
    assert np.array_equal(bits(buf._it_min._value), bits(g["after_fill_min"]))
    for r in range(int(g["rounds"])):
        draws.load(g[f"r{r}_u"])
        res = buf.sample(batch, beta)
        assert list(res[-1]) == list(g[f"r{r}_idx"])
        np.testing.assert_allclose(res[-2], g[f"r{r}_w"], rtol=RTOL)
        assert res[-2].dtype == np.float64
        assert np.array_equal(res[0], g[f"r{r}_obs"]) and np.array_equal(res[3], g[f"r{r}_next"])
        assert np.array_equal(res[1], g[f"r{r}_act"])
        assert np.array_equal(res[2], g[f"r{r}_rew"]) and np.array_equal(res[4], g[f"r{r}_done"])
        if variant == "aql":
            assert np.array_equal(res[5], g[f"r{r}_amu"])
        # the callers' next step (ApeX.py:50-55) must accept what sample returned
        torch.FloatTensor(res[0]), torch.LongTensor(res[1]), torch.FloatTensor(res[2]), torch.FloatTensor(res[-2])
        buf.update_priorities([int(k) for k in g[f"r{r}_upd_idx"]], g[f"r{r}_upd_prio"])
        assert np.array_equal(bits(buf._it_sum._value), bits(g[f"r{r}_sum"]))
        assert np.array_equal(bits(buf._it_min._value), bits(g[f"r{r}_min"]))
        assert buf._max_priority == float(g[f"r{r}_maxp"])
        for row in replay_rows(g, f"r{r}_add", variant):
            buf.add(*row)
        assert int(buf._next_idx) == int(g[f"r{r}_cursor"]) and len(buf) == int(g[f"r{r}_len"])
        assert len(buf._storage) == len(buf)
    assert np.array_equal(bits(buf._it_sum._value), bits(g["final_sum"]))
    assert np.array_equal(bits(buf._it_min._value), bits(g["final_min"]))


def test_config1_golden(memory, monkeypatch):
    """BASELINE config 1: size 100000 -> 2^17 leaves, batch 512, alpha .6, beta .4."""
    g = load("per_c1.npz")
    size, batch = int(g["size"]), int(g["batch"])
    rng = np.random.default_rng(int(g["seed"]))
    prio0 = np.abs(rng.standard_normal(size)) + 1e-6
    buf = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    dummy = np.zeros(16, dtype=np.uint8)
    for i in range(size):
        buf.add(dummy, 0, 0.0, dummy, 0.0, float(prio0[i]))
    assert np.array_equal(sha(buf._it_sum._value), g["fill_sum_sha"])
    assert np.array_equal(sha(buf._it_min._value), g["fill_min_sha"])
    draws = Draws(memory, monkeypatch)
    for r in range(int(g["rounds"])):
        draws.load(rng.random(batch))
        res = buf.sample(batch, 0.4)
        assert np.array_equal(np.asarray(res[-1]), g[f"r{r}_idx"])
        np.testing.assert_allclose(res[-2], g[f"r{r}_w"], rtol=RTOL)
        buf.update_priorities(res[-1], np.abs(rng.standard_normal(batch)) + 1e-6)
        assert np.array_equal(sha(buf._it_sum._value), g[f"r{r}_sum_sha"])
        assert np.array_equal(sha(buf._it_min._value), g[f"r{r}_min_sha"])


# --------------------------------------------------------------------- full-size, device API
def fill_device(buf, n, obs_shape, rng_seed, chunk=8192, with_prio=True):
    """Pre-fills through add_batch_device with rows generated on the GPU; returns host copies
    of the priorities and a checksum column (first 8 bytes of every observation)."""
    gen = torch.Generator(device="cuda").manual_seed(rng_seed)
    prios, tags = [], []
    done_rows = 0
    while done_rows < n:
        m = min(chunk, n - done_rows)
        obs = torch.randint(0, 256, (m,) + obs_shape, dtype=torch.uint8, device="cuda", generator=gen)
        nxt = torch.randint(0, 256, (m,) + obs_shape, dtype=torch.uint8, device="cuda", generator=gen)
        act = torch.randint(0, 18, (m,), dtype=torch.int64, device="cuda", generator=gen)
        rew = torch.randint(-1, 2, (m,), device="cuda", generator=gen).float()
        done = (torch.rand(m, device="cuda", generator=gen) < 0.001).float()
        pr = (torch.randn(m, dtype=torch.float64, device="cuda", generator=gen).abs() + 1e-6) if with_prio else None
        buf.add_batch_device(obs, nxt, act, rew, done, priorities=pr)
        if pr is not None:
            prios.append(pr.cpu().numpy())
        tags.append(obs.reshape(m, -1)[:, :8].cpu().numpy().copy())
        done_rows += m
    return (np.concatenate(prios) if prios else None), np.concatenate(tags)


@pytest.mark.parametrize("size,batch", [(1 << 17, 512), (1 << 20, 512), (1 << 20, 2048), (1 << 22, 1500), (1 << 20, 4096), (1 << 20, 8191), (1 << 20, 8192), (300000, 20000), (1 << 20, 65536), (1 << 22, 65536), (1 << 16, 4096), (1 << 13, 2048), (1 << 12, 1500), (1 << 17, 2048), (1 << 14, 2500), (1 << 21, 3000), (1 << 23, 5000)])
def test_sample_update_full_size_properties(memory, size, batch):
    """BASELINE config 2 tree sizes, tree-only rows (16-byte observations so the test stays
    light).  Device pow: leaves within 2 ulp of libm's; every internal node must then be the
    exact op of its children (the invariant that makes the tree a function of its leaves),
    indices must equal the oracle's descent on the read-back tree, weights within 1e-5."""
    rng = np.random.default_rng(size + batch)
    buf = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    prio, _ = fill_device(buf, size, (16,), rng_seed=3, chunk=1 << 16)
    cap = po.next_pow2(size)
    sv, mv = buf._it_sum._value, buf._it_min._value
    want_leaves = po.pow_alpha(prio[:70000], 0.6)
    ulp = np.abs(bits(sv[cap:cap + 70000]).astype(np.int64) - bits(want_leaves).astype(np.int64))
    assert ulp.max() <= 2
    assert np.array_equal(bits(sv)[1:], bits(po.build_tree(sv[cap:], "sum"))[1:])
    assert np.array_equal(bits(mv)[1:], bits(po.build_tree(mv[cap:], "min"))[1:])
    for step in range(3):
        u = rng.random(batch)
        d = buf.sample_device(batch, 0.4, uniforms=u, gather=False)
        total = po.prefix_total(sv, size - 1)
        want_idx = po.descend_many(sv, po.stratified_masses(total, batch, u))
        idx = d["idx"].cpu().numpy()
        assert np.array_equal(idx, want_idx)
        assert np.all(np.diff(idx) >= 0) and idx.max() <= size - 2  # stratified => sorted; newest slot excluded (Q1)
        np.testing.assert_allclose(d["weights64"].cpu().numpy(), po.is_weights(sv, mv, idx, size, 0.4), rtol=RTOL)
        np.testing.assert_allclose(d["weights"].cpu().numpy(), po.is_weights(sv, mv, idx, size, 0.4), rtol=RTOL)
        newp = (torch.randn(batch, device="cuda").abs() + 1e-6)
        buf.update_priorities_device(d["idx"], newp)
        sv2, mv2 = buf._it_sum._value, buf._it_min._value
        # leaves: last write wins among duplicates, value = double(p)**alpha within 2 ulp
        lv = po.pow_alpha(newp.double().cpu().numpy(), 0.6)
        want_s = po.apply_updates(sv[cap:], idx, lv)
        d_ulp = np.abs(bits(sv2[cap:]).astype(np.int64) - bits(want_s).astype(np.int64))
        assert d_ulp.max() <= 2
        assert np.array_equal(bits(sv2)[1:], bits(po.build_tree(sv2[cap:], "sum"))[1:])
        assert np.array_equal(bits(mv2)[1:], bits(po.build_tree(mv2[cap:], "min"))[1:])
        assert np.array_equal(bits(sv2[cap:cap + size]), bits(mv2[cap:cap + size]))  # slots >= size keep the neutral leaves (Q9)
        sv, mv = sv2, mv2
    assert buf._max_priority >= 1.0


def test_exact_leaf_path_full_size(memory, monkeypatch):
    """Host-facing update_priorities at 2^20 leaves: leaves come from libm pow, so the whole
    node array must equal the oracle's bit for bit."""
    size, batch = 1 << 20, 512
    rng = np.random.default_rng(77)
    buf = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    fill_device(buf, size, (16,), rng_seed=4, chunk=1 << 16)
    cap = size
    leaves_s = buf._it_sum._value[cap:].copy()
    leaves_m = buf._it_min._value[cap:].copy()
    draws = Draws(memory, monkeypatch)
    for step in range(3):
        sv = po.build_tree(leaves_s, "sum")
        u = rng.random(batch)
        draws.load(u)
        res = buf.sample(batch, 0.4)
        want_idx = po.descend_many(sv, po.stratified_masses(po.prefix_total(sv, size - 1), batch, u))
        assert np.array_equal(np.asarray(res[-1]), want_idx)
        newp = np.abs(rng.standard_normal(batch)) + 1e-6
        buf.update_priorities(res[-1], newp.astype(np.float32))  # float32 in, like utils.py:77
        lv = po.pow_alpha(newp.astype(np.float32).astype(np.float64), 0.6)
        leaves_s = po.apply_updates(leaves_s, want_idx, lv)
        leaves_m = po.apply_updates(leaves_m, want_idx, lv)
        assert np.array_equal(bits(buf._it_sum._value)[1:], bits(po.build_tree(leaves_s, "sum"))[1:])
        assert np.array_equal(bits(buf._it_min._value)[1:], bits(po.build_tree(leaves_m, "min"))[1:])


# ------------------------------------------------------------------------------------ gather
@pytest.mark.parametrize("obs_shape,size,batch", [((4, 84, 84), 6000, 512), ((4, 84, 84), 300, 37), ((3, 5), 1000, 64),
                                                  ((17,), 500, 33), ((48,), 700, 129), ((16,), 900, 257), ((4, 84, 84), 5000, 3000)])
@pytest.mark.parametrize("mover", ["tma", "ldg", "bal16", "bal16_nopipe", "bal32", "bal32_d2"])
def test_gather_matches_stored_rows(memory, obs_shape, size, batch, mover):
    """Every mover of the gather (per_tune "gather_mode": 0 = TMA bulk copies through shared-memory rings,
    1 = one 7 056-byte piece per warp through registers, 2 = balanced shares, 16- or 32-byte vectors, with and
    without the software pipeline) returns the stored bytes; rows whose size is not a multiple of 16 bytes take
    the byte mover whatever the mode.  Includes ring wrap-around (more adds than slots)."""
    import per_native as pn
    gen = torch.Generator(device="cuda").manual_seed(9)
    buf = memory.PrioritizedReplayBuffer(size, 0.6)
    tune = {"tma": dict(gather_mode=0), "ldg": dict(gather_mode=1), "bal16": dict(gather_mode=2, gather_vbytes=16, gather_depth=8, gather_pipe=1),
            "bal16_nopipe": dict(gather_mode=2, gather_vbytes=16, gather_depth=16, gather_pipe=0),
            "bal32": dict(gather_mode=2, gather_vbytes=32, gather_depth=4, gather_pipe=1),
            "bal32_d2": dict(gather_mode=2, gather_vbytes=32, gather_depth=2, gather_pipe=0, gather_ctas_per_sm=8)}[mover]
    total = size + size // 3
    obs = torch.randint(0, 256, (total,) + obs_shape, dtype=torch.uint8, device="cuda", generator=gen)
    nxt = torch.randint(0, 256, (total,) + obs_shape, dtype=torch.uint8, device="cuda", generator=gen)
    act = torch.randint(0, 18, (total,), dtype=torch.int64, device="cuda", generator=gen)
    rew = torch.randn(total, device="cuda", generator=gen)
    done = (torch.rand(total, device="cuda", generator=gen) < 0.1).float()
    for lo in range(0, total, 777):
        hi = min(total, lo + 777)
        buf.add_batch_device(obs[lo:hi].contiguous(), nxt[lo:hi].contiguous(), act[lo:hi].contiguous(),
                             rew[lo:hi].contiguous(), done[lo:hi].contiguous())
    assert len(buf) == size and buf._next_idx == total % size
    for k, v in tune.items():
        pn.check(pn.lib.per_tune(buf._h, k.encode(), v))
    slot_src = np.arange(size)
    slot_src[: total - size] = np.arange(size, total)  # slots overwritten after the wrap
    d = buf.sample_device(batch, 0.5, uniforms=np.random.default_rng(1).random(batch))
    idx = d["idx"].cpu().numpy()
    src = torch.as_tensor(slot_src[idx], device="cuda")
    assert torch.equal(d["obs"], obs[src]) and torch.equal(d["next_obs"], nxt[src])
    assert torch.equal(d["act"], act[src]) and torch.equal(d["rew"], rew[src]) and torch.equal(d["done"], done[src])
    # host-facing uniform sampler (memory.py:184-205) and explicit _encode_sample
    enc = buf._encode_sample([0, size - 1, 5, 5])
    assert np.array_equal(enc[0], obs[torch.as_tensor(slot_src[[0, size - 1, 5, 5]], device="cuda")].cpu().numpy())
    assert len(memory.ReplayBuffer.sample(buf, 8)) == 5


# --------------------------------------------------------------------------------- TD kernel
@pytest.mark.parametrize("case", [0, 1, 2])
def test_td_priority_golden(memory, case):
    g = load("td.npz")
    c = f"c{case}_"
    B = g[c + "q"].shape[0]
    buf = memory.CustomPrioritizedReplayBuffer(4096, 0.6)
    fill_device(buf, 4096, (16,), rng_seed=1)
    dev = lambda a: torch.as_tensor(a).cuda()
    loss, prio, td = buf.td_priority_update(dev(g[c + "q"]), dev(g[c + "qn"]), dev(g[c + "qt"]), dev(g[c + "a"]),
                                            dev(g[c + "r"]), dev(g[c + "d"]), dev(g[c + "w"]), int(g[c + "n"]), 0.99)
    np.testing.assert_allclose(prio.cpu().numpy(), g[c + "prios"], rtol=RTOL)
    np.testing.assert_allclose(float(loss), float(g[c + "loss"]), rtol=RTOL)
    ol, op, otd = po.td_priorities(g[c + "q"], g[c + "qn"], g[c + "qt"], g[c + "a"], g[c + "r"], g[c + "d"], g[c + "w"],
                                   int(g[c + "n"]), 0.99)
    np.testing.assert_allclose(td.cpu().numpy(), otd, rtol=RTOL, atol=1e-7)


@pytest.mark.parametrize("batch", [512, 4096, 20000])
def test_td_fused_update_feeds_the_trees(memory, batch):
    """K5 -> K1 without leaving the device (ApeX.py:58-63): priorities match the oracle's
    float32 arithmetic, leaves = double(prio)**alpha within 2 ulp, nodes exact given leaves.
    batch 20000 exercises the multi-launch TD + update form."""
    size = 1 << 16
    rng = np.random.default_rng(batch)
    buf = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    fill_device(buf, size, (16,), rng_seed=2, chunk=1 << 15)
    cap = size
    d = buf.sample_device(batch, 0.4, uniforms=rng.random(batch))
    q, qn, qt = (torch.randn(batch, 18, device="cuda") for _ in range(3))
    before = buf._it_sum._value[cap:].copy()
    loss, prio, td = buf.td_priority_update(q, qn, qt, d["act"], d["rew"], d["done"], d["weights"], 3, 0.99, idx=d["idx"])
    ol, op, otd = po.td_priorities(q.cpu().numpy(), qn.cpu().numpy(), qt.cpu().numpy(), d["act"].cpu().numpy(),
                                   d["rew"].cpu().numpy(), d["done"].cpu().numpy(), d["weights"].cpu().numpy(), 3, 0.99)
    np.testing.assert_allclose(prio.cpu().numpy(), op, rtol=RTOL)
    np.testing.assert_allclose(td.cpu().numpy(), otd, rtol=RTOL, atol=1e-7)
    np.testing.assert_allclose(float(loss), float(ol), rtol=1e-4)
    sv, mv = buf._it_sum._value, buf._it_min._value
    want = po.apply_updates(before, d["idx"].cpu().numpy(), po.pow_alpha(prio.double().cpu().numpy(), 0.6))
    assert np.abs(bits(sv[cap:]).astype(np.int64) - bits(want).astype(np.int64)).max() <= 2
    assert np.array_equal(bits(sv)[1:], bits(po.build_tree(sv[cap:], "sum"))[1:])
    assert np.array_equal(bits(mv)[1:], bits(po.build_tree(mv[cap:], "min"))[1:])
    buf._sync_max_priority()
    assert buf._max_priority >= float(prio.max())


# ------------------------------------------------------------------------------------ quirks
def test_quirks_and_asserts(memory, monkeypatch):
    buf = memory.PrioritizedReplayBuffer(100, 0.6)  # Q9: 100 slots, 128 leaves
    o = np.zeros((2, 2), dtype=np.uint8)
    with pytest.raises(AssertionError):
        memory.PrioritizedReplayBuffer(10, -1.0)  # memory.py:224
    for i in range(100):
        buf.add(o, i, 0.0, o, 0.0)
    sv = buf._it_sum._value
    assert np.all(sv[128 + 100:] == 0.0) and np.all(buf._it_min._value[128 + 100:] == np.inf)  # Q9
    with pytest.raises(AssertionError):
        buf.sample(8, 0.0)  # memory.py:284
    with pytest.raises(AssertionError):
        buf.update_priorities([1, 2], [1.0])  # memory.py:313
    with pytest.raises(AssertionError):
        buf.update_priorities([1], [0.0])  # memory.py:315
    with pytest.raises(AssertionError):
        buf.update_priorities([100], [1.0])  # memory.py:316
    # Q1: the newest slot (len-1) carries all the mass but can never be drawn
    buf.update_priorities([99], [1e6])
    draws = Draws(memory, monkeypatch)
    draws.load(np.full(16, 0.999999))
    assert max(buf.sample(16, 0.4)[-1]) <= 98
    # Q6: duplicates -> last wins; running max never decays
    buf.update_priorities([5, 5, 5], [3.0, 2.0, 0.5])
    assert buf._it_sum[5] == 0.5 ** 0.6 and buf._max_priority == 1e6
    buf.add(o, 0, 0.0, o, 0.0)  # wraps to slot 0 with max_priority ** alpha (memory.py:239)
    assert buf._next_idx == 1 and len(buf) == 100
    assert buf._it_sum[0] == 1e6 ** 0.6
    # device-side validation: bad entries are skipped and reported
    idx = torch.tensor([1, 2, 3], device="cuda")
    buf.update_priorities_device(idx, torch.tensor([1.0, -1.0, 2.0], device="cuda"))
    with pytest.raises(AssertionError):
        buf._raise_status()
    assert buf._it_sum[1] == 1.0 and buf._it_sum[3] == pytest.approx(2.0 ** 0.6, rel=1e-15)
    # Q8: a float size (ApeX.py:16 passes 1e6) must not break the ring
    fb = memory.CustomPrioritizedReplayBuffer(8.0, 0.6)
    for i in range(20):
        fb.add(o, 0, 0.0, o, 0.0, 1.0 + i)
    assert len(fb) == 8 and fb._next_idx == 4 and fb._max_priority == 20.0


def test_add_chunk_equals_row_by_row_add(memory):
    """The batchrecorder ingest path (batchrecorder.py:133-137): one add_chunk per actor chunk must
    leave the buffer exactly as one add() per transition does, including ring wrap-around."""
    rng = np.random.default_rng(21)
    size, n = 300, 470
    obs = rng.integers(0, 256, size=(n, 4, 84, 84), dtype=np.uint8)
    nxt = rng.integers(0, 256, size=(n, 4, 84, 84), dtype=np.uint8)
    act = rng.integers(0, 18, size=n)
    rew = rng.standard_normal(n)
    done = (rng.random(n) < 0.1).astype(np.float32)
    prio = np.abs(rng.standard_normal(n)) + 1e-6
    a = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    b = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    for i in range(n):
        a.add(obs[i], int(act[i]), float(rew[i]), nxt[i], float(done[i]), float(prio[i]))
    for lo in range(0, n, 50):
        hi = min(n, lo + 50)
        b.add_chunk(list(obs[lo:hi]), list(act[lo:hi]), list(rew[lo:hi]), list(nxt[lo:hi]), list(done[lo:hi]), prio[lo:hi])
    assert len(a) == len(b) == size and a._next_idx == b._next_idx == n % size and a._max_priority == b._max_priority
    assert np.array_equal(bits(a._it_sum._value), bits(b._it_sum._value))
    assert np.array_equal(bits(a._it_min._value), bits(b._it_min._value))
    idx = list(range(0, size, 7))
    ea, eb = a._encode_sample(idx), b._encode_sample(idx)
    for x, y in zip(ea, eb):
        assert np.array_equal(x, y)


@pytest.mark.parametrize("dtype,scale", [(torch.float32, 1.0), (torch.float32, 1.0 / 255.0), (torch.bfloat16, 1.0), (torch.float16, 1.0 / 255.0)])
def test_gather_cast_equals_torch_widening(memory, dtype, scale):
    """Gather fused with the learner's uint8 -> float conversion (ApeX.py:50,53): must equal
    gathering uint8 and converting with torch (float32 exact; 16-bit types = torch's rounding)."""
    gen = torch.Generator(device="cuda").manual_seed(3)
    size, batch = 3000, 257
    buf = memory.PrioritizedReplayBuffer(size, 0.6)
    obs = torch.randint(0, 256, (size, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=gen)
    nxt = torch.randint(0, 256, (size, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=gen)
    z = torch.zeros(size, device="cuda")
    buf.add_batch_device(obs, nxt, torch.arange(size, device="cuda"), z, z)
    u = np.random.default_rng(2).random(batch)
    plain = buf.sample_device(batch, 0.4, uniforms=u)
    cast = buf.sample_device(batch, 0.4, uniforms=u, obs_dtype=dtype, obs_scale=scale)
    assert torch.equal(plain["idx"], cast["idx"]) and torch.equal(plain["act"], cast["act"])
    for k in ("obs", "next_obs"):
        want = (plain[k].float() * np.float32(scale)).to(dtype)
        assert cast[k].dtype == dtype and torch.equal(cast[k], want), k


# ------------------------------------------------------- TD kernel: AQL's [B, K = 1000] shape, large batches
def _golden_module():
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(GOLDEN, "make_golden.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


@pytest.mark.parametrize("case", [0, 1])
def test_td_priority_golden_compute_loss_aql(memory, case):
    """utils.compute_loss_AQL (utils.py:44-61) at its real shape, K = 1000 candidate actions per row: the Q rows of
    one CTA do not fit the staging buffer, so the prologue of k_update_cluster reads them in place (the branch
    `staged == false`).  Golden priorities / loss from the real reference (tests/golden/td_aql.npz; the inputs are
    regenerated from the fixture's seed)."""
    q, qn, qt, a, r, d, w, n_steps = _golden_module().td_aql_inputs(case)
    g = load("td_aql.npz")
    B = q.shape[0]
    buf = memory.CustomPrioritizedReplayBuffer(4096, 0.6)
    fill_device(buf, 4096, (16,), rng_seed=1)
    dev = lambda x: torch.as_tensor(x).cuda()
    smp = buf.sample_device(B, 0.4, uniforms=np.random.default_rng(case).random(B), gather=False)
    before = buf._it_sum._value[4096:].copy()
    loss, prio, td = buf.td_priority_update(dev(q), dev(qn), dev(qt), dev(a), dev(r), dev(d), dev(w), n_steps, 0.99, idx=smp["idx"])
    np.testing.assert_allclose(prio.cpu().numpy(), g[f"c{case}_prios"], rtol=RTOL)
    np.testing.assert_allclose(float(loss), float(g[f"c{case}_loss"]), rtol=RTOL)
    ol, op, otd = po.td_priorities(q, qn, qt, a, r, d, w, n_steps, 0.99)
    np.testing.assert_allclose(td.cpu().numpy(), otd, rtol=RTOL, atol=1e-7)
    sv = buf._it_sum._value
    want = po.apply_updates(before, smp["idx"].cpu().numpy(), po.pow_alpha(prio.double().cpu().numpy(), 0.6))
    assert np.abs(bits(sv[4096:]).astype(np.int64) - bits(want).astype(np.int64)).max() <= 2
    assert np.array_equal(bits(sv)[1:], bits(po.build_tree(sv[4096:], "sum"))[1:])


@pytest.mark.parametrize("batch,nA", [(20000, 4), (20000, 1000), (5000, 1000), (4096, 37)])
def test_td_priority_large_batches_and_action_counts(memory, batch, nA):
    """The multi-launch TD form (k_td_pass1 / k_td_pass2, batches above 4096) and the one-launch form with action
    counts other than Atari's 18, against the oracle's float32 restatement of utils.py:71-80; then the leaves."""
    size = 1 << 15
    rng = np.random.default_rng(batch + nA)
    buf = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    fill_device(buf, size, (16,), rng_seed=3, chunk=1 << 15)
    idx = torch.as_tensor(rng.integers(0, size, size=batch)).cuda()
    f32 = lambda *s: torch.as_tensor(rng.standard_normal(s).astype(np.float32)).cuda()
    q, qn, qt = f32(batch, nA), f32(batch, nA), f32(batch, nA)
    a = torch.as_tensor(rng.integers(0, nA, size=batch)).cuda()
    r, d, w = f32(batch), torch.as_tensor((rng.random(batch) < 0.1).astype(np.float32)).cuda(), f32(batch).abs()
    before = buf._it_sum._value[size:].copy()
    loss, prio, td = buf.td_priority_update(q, qn, qt, a, r, d, w, 3, 0.99, idx=idx)
    ol, op, otd = po.td_priorities(q.cpu().numpy(), qn.cpu().numpy(), qt.cpu().numpy(), a.cpu().numpy(), r.cpu().numpy(),
                                   d.cpu().numpy(), w.cpu().numpy(), 3, 0.99)
    np.testing.assert_allclose(prio.cpu().numpy(), op, rtol=RTOL)
    np.testing.assert_allclose(td.cpu().numpy(), otd, rtol=RTOL, atol=1e-7)
    np.testing.assert_allclose(float(loss), float(ol), rtol=1e-4)
    sv, mv = buf._it_sum._value, buf._it_min._value
    want = po.apply_updates(before, idx.cpu().numpy(), po.pow_alpha(prio.double().cpu().numpy(), 0.6))
    assert np.abs(bits(sv[size:]).astype(np.int64) - bits(want).astype(np.int64)).max() <= 2
    assert np.array_equal(bits(sv)[1:], bits(po.build_tree(sv[size:], "sum"))[1:])
    assert np.array_equal(bits(mv)[1:], bits(po.build_tree(mv[size:], "min"))[1:])


@pytest.mark.parametrize("batch", [256, 6000])
def test_td_nan_and_bad_action_are_reported(memory, batch):
    """torch.max propagates a NaN TD error into every priority of the batch (utils.py:77) and update_priorities
    then asserts priority > 0 (memory.py:315): here every priority comes out NaN, nothing is written to the trees
    and the status word says so.  An action outside [0, nA) is reported as a bad index instead of reading another row."""
    size = 4096
    buf = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    fill_device(buf, size, (16,), rng_seed=4)
    g = torch.Generator(device="cuda").manual_seed(batch)
    q, qn, qt = (torch.randn(batch, 18, device="cuda", generator=g) for _ in range(3))
    a = torch.randint(0, 18, (batch,), device="cuda", generator=g)
    r, d, w = torch.randn(batch, device="cuda", generator=g), torch.zeros(batch, device="cuda"), torch.ones(batch, device="cuda")
    idx = torch.randint(0, size, (batch,), device="cuda", generator=g)
    before = buf._it_sum._value.copy()
    q_nan = q.clone()
    q_nan[batch // 2, int(a[batch // 2])] = float("nan")
    _, prio, _ = buf.td_priority_update(q_nan, qn, qt, a, r, d, w, 3, 0.99, idx=idx)
    assert bool(torch.isnan(prio).all())
    with pytest.raises(AssertionError, match="priority > 0"):
        buf._raise_status()
    assert np.array_equal(bits(buf._it_sum._value), bits(before))
    a_bad = a.clone()
    a_bad[3] = 18
    buf.td_priority_update(q, qn, qt, a_bad, r, d, w, 3, 0.99, idx=None)
    with pytest.raises(AssertionError, match="idx"):
        buf._raise_status()


# --------------------------------------------------------------- uniform sampler, tree views, graph replay
def test_uniform_replay_buffer_sample_matches_the_oracle(memory):
    """ReplayBuffer.sample (memory.py:184-205): indices from random.randint on the global RNG, rows stacked."""
    import random
    rng = np.random.default_rng(12)
    size, n = 500, 730  # wraps
    buf = memory.ReplayBuffer(size)
    ora = po.ReplayOracle(size, 0.0, "per")
    for i in range(n):
        row = (rng.integers(0, 256, size=(3, 4), dtype=np.uint8), int(rng.integers(0, 6)), float(rng.standard_normal()),
               rng.integers(0, 256, size=(3, 4), dtype=np.uint8), float(i % 7 == 0))
        buf.add(*row)
        ora.add(*row)
    assert len(buf) == len(ora) == size and buf._next_idx == ora.cursor == n % size
    for batch in (1, 32, 700):
        random.seed(batch)
        got = buf.sample(batch)
        random.seed(batch)
        idxes = [random.randint(0, len(ora) - 1) for _ in range(batch)]  # memory.py:204
        want = ora.gather(idxes)
        assert len(got) == 5
        for gcol, wcol in zip(got, want):
            assert gcol.shape == wcol.shape
            assert np.array_equal(gcol, np.asarray(wcol, dtype=gcol.dtype))


def test_tree_views_read_single_leaves_and_ranges(memory):
    """`buffer._it_sum[idx]`, `_it_sum.sum(start, end)`, `_it_min.min(...)` (memory.py:244,289,293) on the buffer's
    own trees, against the oracle's trees after the same adds."""
    rng = np.random.default_rng(3)
    size = 300  # 512 leaves
    buf = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    ora = po.ReplayOracle(size, 0.6, "apex")
    o = np.zeros((2, 2), dtype=np.uint8)
    for i in range(size):
        p = float(abs(rng.standard_normal()) + 1e-3)
        buf.add(o, 0, 0.0, o, 0.0, p)
        ora.add(o, 0, 0.0, o, 0.0, p)
    for k in (0, 1, 77, 299, 300, 511):
        assert bits(buf._it_sum[k]) == bits(ora.sum_tree.get(k)) and bits(buf._it_min[k]) == bits(ora.min_tree.get(k))
    with pytest.raises(AssertionError):
        buf._it_sum[512]
    for a, b in ((0, None), (0, 299), (5, 123), (100, 101), (0, -1), (17, -200)):
        assert bits(buf._it_sum.sum(a, b)) == bits(ora.sum_tree.reduce(a, b))
        assert bits(buf._it_min.min(a, b)) == bits(ora.min_tree.reduce(a, b))
    assert bits(buf._it_sum.sum()) == bits(ora.sum_tree.reduce()) and bits(buf._it_min.min()) == bits(ora.min_tree.reduce())


def test_update_priorities_replayed_from_a_cuda_graph(memory):
    """A captured per_update_priorities call on an arbitrary (unsorted, duplicated) index buffer, replayed while the
    buffer's contents change: the arbitration state a replay leaves behind must not decide the next one."""
    import per_native as pn
    size, n = 1 << 14, 900
    buf = memory.PrioritizedReplayBuffer(size, 0.6)
    fill_device(buf, size, (16,), rng_seed=6, with_prio=False)
    gen = torch.Generator(device="cuda").manual_seed(2)
    idx = torch.zeros(n, dtype=torch.int64, device="cuda")
    pr = torch.ones(n, dtype=torch.float64, device="cuda")
    rounds = []
    for r in range(4):
        i = torch.randint(0, size, (n,), dtype=torch.int64, device="cuda", generator=gen)
        if r % 2 == 0:
            i[n - 60:] = i[0]                # a leaf with many duplicates: its winner is a late item ...
        else:
            i[5] = rounds[-1][0][0]          # ... and next round the same leaf appears once, early
        rounds.append((i, torch.rand(n, dtype=torch.float64, device="cuda", generator=gen) + 1e-3))
    leaves = buf._it_sum._value[size:].copy()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            pn.check(pn.lib.per_update_priorities(buf._h, pn.ptr(idx), pn.ptr(pr), pn.PRIO_F64, n, pn.stream_ptr(side)))
    torch.cuda.current_stream().wait_stream(side)
    for i, p in rounds:
        idx.copy_(i)
        pr.copy_(p)
        g.replay()
        leaves = po.apply_updates(leaves, i.cpu().numpy(), po.pow_alpha(p.cpu().numpy(), 0.6))
    torch.cuda.synchronize()
    sv = buf._it_sum._value
    assert np.abs(bits(sv[size:]).astype(np.int64) - bits(leaves).astype(np.int64)).max() <= 2
    assert np.array_equal(bits(sv)[1:], bits(po.build_tree(sv[size:], "sum"))[1:])
