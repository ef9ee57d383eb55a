"""Differential test of the oracle against the REAL reference (runs only where /root/reference exists, i.e. in
the build container; the GPU box has the committed fixtures instead).  Randomised operation sequences —
add / sample / update_priorities interleavings, duplicate indices, ring wrap-around, the three buffer variants —
go through `/root/reference/memory.py` (imported unedited, NumPy-2 shim of SURVEY.md section 8c, uniforms replayed
through its `random.random`) and through oracle/per_oracle.py; every tree node, every sampled index, every weight and
every returned row must agree (bit-exact for nodes / indices / weights: both are CPython float arithmetic)."""
import importlib.util
import os
import random
import types

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import per_oracle as po

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.exists(os.path.join(REF, "memory.py")), reason="the reference tree is only present in the build container")


def bits(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64)).view(np.uint64)


@pytest.fixture(scope="module")
def ref():
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(GOLDEN, "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    mem = mg.load_reference("memory")
    mem.np = mg.numpy1_shim()
    rnd = mg.Replay()
    mem.random = types.SimpleNamespace(random=rnd.random, randint=random.randint)
    return mem, rnd


@pytest.mark.parametrize("variant,size,seed", [("per", 37, 0), ("per", 64, 1), ("apex", 50, 2), ("apex", 129, 3), ("aql", 23, 4), ("aql", 200, 5)])
def test_random_op_sequences_agree_with_the_reference(ref, variant, size, seed):
    mem, rnd = ref
    rng = np.random.default_rng(seed)
    cls = {"per": mem.PrioritizedReplayBuffer, "apex": mem.CustomPrioritizedReplayBuffer, "aql": mem.CustomPrioritizedReplayBuffer_AQL}[variant]
    real, ora = cls(size, 0.6), po.ReplayOracle(size, 0.6, variant)
    n_ops = 0
    for step in range(400):
        op = rng.choice(["add", "add", "sample", "update"]) if len(ora) >= 2 else "add"
        if op == "add":
            for _ in range(int(rng.integers(1, 2 * size // 3 + 2))):  # runs long enough to wrap the ring several times
                row = [rng.integers(0, 256, size=(2, 3), dtype=np.uint8), int(rng.integers(0, 5)), float(rng.standard_normal()),
                       rng.integers(0, 256, size=(2, 3), dtype=np.uint8), float(rng.random() < 0.1)]
                if variant == "apex":
                    row.append(float(abs(rng.standard_normal()) + 1e-4))
                elif variant == "aql":
                    row.append(rng.standard_normal((4, 2)).astype(np.float32))
                real.add(*row)
                ora.add(*row)
        elif op == "sample":
            batch, beta = int(rng.integers(1, 40)), float(rng.uniform(0.1, 1.0))
            u = rng.random(batch)
            rnd.load(u)
            got = real.sample(batch, beta)
            want = ora.sample(batch, beta, u)
            assert list(got[-1]) == list(want[-1])
            assert np.array_equal(bits(got[-2]), bits(want[-2]))
            for gcol, wcol in zip(got[:-2], want[:-2]):
                assert np.array_equal(np.asarray(gcol), np.asarray(wcol))
        else:
            n = int(rng.integers(1, 50))
            idx = [int(k) for k in rng.integers(0, len(ora), size=n)]
            if n > 3:
                idx[1] = idx[0]  # duplicates: the later priority wins (memory.py:314-318)
                idx[-1] = idx[0]
            pr = [float(p) for p in np.abs(rng.standard_normal(n)) + 1e-5]
            real.update_priorities(idx, pr)
            ora.update_priorities(idx, pr)
        n_ops += 1
        assert len(real) == len(ora) and int(real._next_idx) == int(ora.cursor)
        assert real._max_priority == ora.max_priority
        if step % 20 == 0 or step == 399:
            assert np.array_equal(bits(real._it_sum._value), bits(ora.sum_tree.v))
            assert np.array_equal(bits(real._it_min._value), bits(ora.min_tree.v))
    assert n_ops == 400


@pytest.mark.parametrize("cap", [8, 64, 1024])
def test_tree_ranges_and_descents_agree_with_the_reference(ref, cap):
    mem, _ = ref
    rng = np.random.default_rng(cap)
    rs, rm = mem.SumSegmentTree(cap), mem.MinSegmentTree(cap)
    os_, om = po.HeapTree(cap, "sum"), po.HeapTree(cap, "min")
    for _ in range(3 * cap):
        k, v = int(rng.integers(0, cap)), float(abs(rng.standard_normal()) + 1e-9)
        rs[k] = v
        rm[k] = v
        os_.set(k, v)
        om.set(k, v)
    assert np.array_equal(bits(rs._value), bits(os_.v)) and np.array_equal(bits(rm._value), bits(om.v))
    for _ in range(300):
        a = int(rng.integers(0, cap))
        b = int(rng.integers(a + 1, cap + 1))
        assert bits(rs.sum(a, b)) == bits(os_.reduce(a, b)) and bits(rm.min(a, b)) == bits(om.reduce(a, b))
        m = float(rng.random() * os_.reduce())
        assert rs.find_prefixsum_idx(m) == os_.descend(m)
    vec = po.build_tree(np.array(os_.v[cap:]), "sum")
    assert np.array_equal(bits(vec)[1:], bits(os_.v)[1:])
    masses = rng.random(200) * vec[1]
    assert np.array_equal(po.descend_many(vec, masses), [rs.find_prefixsum_idx(float(m)) for m in masses])
    for n_leaves in (1, 2, cap // 2, cap - 1, cap):
        assert bits(po.prefix_total(vec, n_leaves)) == bits(rs.sum(0, n_leaves))


def test_nstep_oracle_agrees_with_the_reference_batchstorage(ref):
    mem, _ = ref
    rng = np.random.default_rng(9)
    for n in (1, 3, 5):
        real, ora = mem.BatchStorage(n, 0.99), po.NStepOracle(n, 0.99)
        started = False
        for t in range(300):
            done = bool(rng.random() < 0.07) and started  # a done on an empty window raises in both (Q10); avoided here
            args = (rng.standard_normal(3).astype(np.float32), float(rng.integers(-1, 2)), int(rng.integers(0, 4)), done,
                    rng.standard_normal(4).astype(np.float32))
            real.add(*args)
            ora.add(*args)
            started = not done
            if len(ora) and (done or len(ora) == 50):
                (rb, rp), (ob, op) = real.make_batch(), ora.make_batch()
                assert np.array_equal(bits(rp), bits(op))
                for rc, oc in zip(rb, ob):
                    assert len(rc) == len(oc) and all(np.array_equal(np.asarray(x), np.asarray(y)) for x, y in zip(rc, oc))
                real.reset()
                ora.reset()
