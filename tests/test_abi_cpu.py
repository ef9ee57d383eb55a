"""CPU-side checks of the boundary: the shared library loads without a GPU, exports every
symbol include/per_b200.h declares, the host helper is bit-identical to CPython's pow, and
the product classes fail loudly (no fallback) when there is no CUDA device."""
import ctypes
import math
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT

import __graft_entry__ as entry


@pytest.fixture(scope="module", autouse=True)
def built():
    entry.build()


def declared_functions():
    text = open(os.path.join(ROOT, "include", "per_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b((?:per|seg|aql|nstep)_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    import per_native
    names = declared_functions()
    assert len(names) >= 20
    raw = ctypes.CDLL(per_native.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), "libper_b200.so does not export " + n
    assert sorted(per_native.SIGNATURES) == names, "per_native.SIGNATURES and the header disagree"
    assert per_native.lib.per_version() >= 100


def test_host_pow_is_libm_pow():
    import per_native
    rng = np.random.default_rng(5)
    p = np.abs(rng.standard_normal(5000)) + 1e-6
    out = np.empty_like(p)
    for alpha in (0.6, 0.4, 1.0, 0.0, 0.123456789):
        assert per_native.lib.per_pow_alpha_host(p.ctypes.data, alpha, out.ctypes.data, p.size) == 0
        want = np.array([float(x) ** alpha for x in p])
        assert np.array_equal(out.view(np.uint64), want.view(np.uint64))
        assert np.array_equal(out.view(np.uint64), np.array([math.pow(x, alpha) for x in p]).view(np.uint64))


def test_asserts_that_need_no_gpu():
    import per_native
    h = ctypes.c_void_p()
    rc = per_native.lib.seg_create(ctypes.byref(h), 0, 48, per_native.TREE_SUM)
    assert rc == per_native.PER_ERR_ASSERT  # memory.py:34
    with pytest.raises(AssertionError):
        per_native.check(rc)
    rc = per_native.lib.per_create(ctypes.byref(h), 0, 16, -0.5, 16, 0)
    assert rc == per_native.PER_ERR_ASSERT  # memory.py:224
    assert b"alpha" in per_native.lib.per_last_error()


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    import memory
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        memory.PrioritizedReplayBuffer(128, 0.6)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        memory.SumSegmentTree(16)
    with pytest.raises(AssertionError):
        memory.SumSegmentTree(48)  # asserted before any device work, like memory.py:34


def test_batch_storage_matches_golden():
    """The actor-side BatchStorage mirror against the fixtures generated from the reference."""
    import memory
    from conftest import GOLDEN
    g = np.load(os.path.join(GOLDEN, "nstep.npz"))
    for n in (1, 3, 5):
        bs = memory.BatchStorage(n, 0.99)
        for t in range(len(g[f"n{n}_rew"])):
            bs.add(t, float(g[f"n{n}_rew"][t]), int(g[f"n{n}_act"][t]), bool(g[f"n{n}_done"][t]), g[f"n{n}_q"][t])
        batch, prios = bs.make_batch()
        assert list(batch[0]) == list(g[f"n{n}_out_s"]) and list(batch[3]) == list(g[f"n{n}_out_sn"])
        assert np.array_equal(np.asarray(batch[2]).view(np.uint64), g[f"n{n}_out_R"].view(np.uint64))
        assert np.array_equal(np.asarray(prios).view(np.uint64), g[f"n{n}_out_prio"].view(np.uint64))
        assert len(bs) == len(g[f"n{n}_out_s"])
