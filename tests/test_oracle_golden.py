"""Pins the oracle (oracle/per_oracle.py) to the fixtures that tests/golden/make_golden.py
produced by running the real reference (memory.py, utils.py).  CPU only."""
import hashlib
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import per_oracle as po


def load(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def bits(a):
    return np.asarray(a, dtype=np.float64).view(np.uint64)


def sha(values):
    return np.frombuffer(hashlib.sha256(np.asarray(values, dtype=np.float64).tobytes()).digest(), dtype=np.uint8)


def test_tree_point_sets_and_ranges():
    g = load("tree_ops.npz")
    cap = int(g["cap"])
    st, mt = po.HeapTree(cap, "sum"), po.HeapTree(cap, "min")
    for k, x in zip(g["set_idx"], g["set_val"]):
        st.set(int(k), float(x))
        mt.set(int(k), float(x))
    assert np.array_equal(bits(st.v), bits(g["sum_nodes"]))
    assert np.array_equal(bits(mt.v), bits(g["min_nodes"]))
    for (a, b), s, m in zip(g["ranges"], g["range_sum"], g["range_min"]):
        assert bits(st.reduce(int(a), int(b))) == bits(s)
        assert bits(mt.reduce(int(a), int(b))) == bits(m)
    for (a, b), s in zip(g["neg_ranges"], g["neg_sum"]):
        assert bits(st.reduce(int(a), int(b))) == bits(s)
    assert bits(st.reduce()) == bits(g["full_sum"]) and bits(mt.reduce()) == bits(g["full_min"])
    assert [st.descend(float(m)) for m in g["masses"]] == list(g["found"])
    assert [st.descend(float(m)) for m in g["tie_masses"]] == list(g["tie_found"])
    # vectorised forms agree with the scalar ones
    leaves = np.array(st.v[cap:])
    assert np.array_equal(bits(po.build_tree(leaves, "sum")[1:]), bits(g["sum_nodes"][1:]))
    assert np.array_equal(bits(po.build_tree(np.array(mt.v[cap:]), "min")[1:]), bits(g["min_nodes"][1:]))
    assert list(po.descend_many(np.array(st.v), g["masses"])) == list(g["found"])
    for n in range(1, cap + 1):
        assert bits(po.prefix_total(np.array(st.v), n)) == bits(st.reduce(0, n))


def test_power_of_two_assert():
    with pytest.raises(AssertionError):
        po.HeapTree(48, "sum")


def replay_rows(g, tag, variant):
    n = g[tag + "_obs"].shape[0]
    for i in range(n):
        extra = None
        if variant == "apex":
            extra = float(g[tag + "_extra"][i])
        elif variant == "aql":
            extra = g[tag + "_extra"][i]
        yield (g[tag + "_obs"][i], int(g[tag + "_act"][i]), float(g[tag + "_rew"][i]),
               g[tag + "_next"][i], float(g[tag + "_done"][i]), extra)


@pytest.mark.parametrize("variant,fname", [("per", "per_small.npz"), ("apex", "per_apex.npz"), ("aql", "per_aql.npz")])
def test_replay_sequences(variant, fname):
    g = load(fname)
    buf = po.ReplayOracle(int(g["size"]), float(g["alpha"]), variant)
    beta, batch = float(g["beta"]), int(g["batch"])
    for row in replay_rows(g, "fill", variant):
        buf.add(*row)
    assert np.array_equal(bits(buf.sum_tree.v), bits(g["after_fill_sum"]))
    assert np.array_equal(bits(buf.min_tree.v), bits(g["after_fill_min"]))
    for r in range(int(g["rounds"])):
        res = buf.sample(batch, beta, g[f"r{r}_u"])
        assert list(res[-1]) == list(g[f"r{r}_idx"])
        np.testing.assert_allclose(res[-2], g[f"r{r}_w"], rtol=1e-12)
        assert np.array_equal(res[0], g[f"r{r}_obs"]) and np.array_equal(res[3], g[f"r{r}_next"])
        assert np.array_equal(res[1], g[f"r{r}_act"])
        assert np.array_equal(res[2], g[f"r{r}_rew"]) and np.array_equal(res[4], g[f"r{r}_done"])
        if variant == "aql":
            assert np.array_equal(res[5], g[f"r{r}_amu"])
        buf.update_priorities([int(k) for k in g[f"r{r}_upd_idx"]], [float(p) for p in g[f"r{r}_upd_prio"]])
        assert np.array_equal(bits(buf.sum_tree.v), bits(g[f"r{r}_sum"]))
        assert np.array_equal(bits(buf.min_tree.v), bits(g[f"r{r}_min"]))
        assert buf.max_priority == float(g[f"r{r}_maxp"])
        for row in replay_rows(g, f"r{r}_add", variant):
            buf.add(*row)
        assert int(buf.cursor) == int(g[f"r{r}_cursor"]) and len(buf) == int(g[f"r{r}_len"])
    assert np.array_equal(bits(buf.sum_tree.v), bits(g["final_sum"]))
    assert np.array_equal(bits(buf.min_tree.v), bits(g["final_min"]))


def test_config1_vectorised_forms():
    """BASELINE config 1 (2^17 leaves, batch 512) through the numpy forms."""
    g = load("per_c1.npz")
    size, batch = int(g["size"]), int(g["batch"])
    rng = np.random.default_rng(int(g["seed"]))
    prio0 = np.abs(rng.standard_normal(size)) + 1e-6
    cap = po.next_pow2(size)
    leaves_s = np.zeros(cap)
    leaves_m = np.full(cap, np.inf)
    leaves_s[:size] = leaves_m[:size] = po.pow_alpha(prio0, 0.6)
    sv, mv = po.build_tree(leaves_s, "sum"), po.build_tree(leaves_m, "min")
    assert np.array_equal(sha(sv), g["fill_sum_sha"]) and np.array_equal(sha(mv), g["fill_min_sha"])
    for r in range(int(g["rounds"])):
        u = rng.random(batch)
        total = po.prefix_total(sv, size - 1)
        idx = po.descend_many(sv, po.stratified_masses(total, batch, u))
        assert np.array_equal(idx, g[f"r{r}_idx"])
        np.testing.assert_allclose(po.is_weights(sv, mv, idx, size, 0.4), g[f"r{r}_w"], rtol=1e-12)
        pr = np.abs(rng.standard_normal(batch)) + 1e-6
        lv = po.pow_alpha(pr, 0.6)
        leaves_s = po.apply_updates(leaves_s, idx, lv)
        leaves_m = po.apply_updates(leaves_m, idx, lv)
        sv, mv = po.build_tree(leaves_s, "sum"), po.build_tree(leaves_m, "min")
        assert np.array_equal(sha(sv), g[f"r{r}_sum_sha"]) and np.array_equal(sha(mv), g[f"r{r}_min_sha"])


@pytest.mark.parametrize("n", [1, 3, 5])
def test_nstep_assembly(n):
    g = load("nstep.npz")
    ns = po.NStepOracle(n, 0.99)
    for t in range(len(g[f"n{n}_rew"])):
        ns.add(t, float(g[f"n{n}_rew"][t]), int(g[f"n{n}_act"][t]), bool(g[f"n{n}_done"][t]), g[f"n{n}_q"][t])
    batch, prios = ns.make_batch()
    assert list(batch[0]) == list(g[f"n{n}_out_s"]) and list(batch[3]) == list(g[f"n{n}_out_sn"])
    assert list(batch[1]) == list(g[f"n{n}_out_a"])
    assert np.array_equal(bits(batch[2]), bits(g[f"n{n}_out_R"]))
    assert np.array_equal(np.asarray(batch[4], np.float32), g[f"n{n}_out_d"])
    assert np.array_equal(bits(prios), bits(g[f"n{n}_out_prio"]))


def test_nstep_done_on_empty_window_raises():
    ns = po.NStepOracle(3)
    with pytest.raises(IndexError):
        ns.add(0, 0.0, 0, True, np.zeros(2, np.float32))


@pytest.mark.parametrize("case", [0, 1, 2])
def test_td_priorities(case):
    g = load("td.npz")
    c = f"c{case}_"
    loss, prios, _ = po.td_priorities(g[c + "q"], g[c + "qn"], g[c + "qt"], g[c + "a"], g[c + "r"], g[c + "d"],
                                      g[c + "w"], int(g[c + "n"]), 0.99)
    np.testing.assert_allclose(prios, g[c + "prios"], rtol=1e-6)
    np.testing.assert_allclose(loss, g[c + "loss"], rtol=1e-5)
