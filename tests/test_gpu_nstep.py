"""K7 — device n-step assembly against the golden BatchStorage runs of the reference
(tests/golden/nstep.npz) and against the oracle on random multi-stream input (episodes shorter
than n, back-to-back dones, the Q window that survives episode ends, pre-roll continuation)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import per_oracle as po

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nstep():
    import __graft_entry__ as entry
    entry.build()
    import nstep as m
    return m


def bits(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64)).view(np.uint64)


def oracle_stream(rew, act, done, q, n, gamma, first_step=0):
    ns = po.NStepOracle(n, gamma)
    for t in range(len(rew)):
        ns.add(first_step + t, float(rew[t]), int(act[t]), bool(done[t]), q[t])
    batch, prios = ns.make_batch() if len(ns) else ([[], [], [], [], []], np.zeros(0))
    return batch, prios


@pytest.mark.parametrize("n", [1, 3, 5])
def test_nstep_golden(nstep, n):
    g = np.load(os.path.join(GOLDEN, "nstep.npz"))
    dev = "cuda"
    out = nstep.nstep_assemble_device(torch.as_tensor(g[f"n{n}_rew"], device=dev), torch.as_tensor(g[f"n{n}_act"], device=dev),
                                      torch.as_tensor(g[f"n{n}_done"], device=dev), torch.as_tensor(g[f"n{n}_q"], device=dev), n, 0.99)
    assert out["first"].cpu().tolist() == list(g[f"n{n}_out_s"]) and out["next"].cpu().tolist() == list(g[f"n{n}_out_sn"])
    assert out["action"].cpu().tolist() == list(g[f"n{n}_out_a"])
    assert np.array_equal(bits(out["ret"].cpu().numpy()), bits(g[f"n{n}_out_R"]))          # returns bit-exact
    assert np.array_equal(out["done"].cpu().numpy(), g[f"n{n}_out_d"])
    assert np.array_equal(bits(out["prio"].cpu().numpy()), bits(g[f"n{n}_out_prio"]))      # actor priorities bit-exact


@pytest.mark.parametrize("n,streams,seed", [(3, 7, 0), (1, 5, 1), (8, 40, 2), (33, 9, 3), (64, 3, 4)])
def test_nstep_random_streams_vs_oracle(nstep, n, streams, seed):
    rng = np.random.default_rng(seed)
    nA, gamma = 6, 0.97
    lens = rng.integers(1, 300, size=streams)
    offs = np.concatenate([[0], np.cumsum(lens)])
    T = int(offs[-1])
    rew = rng.standard_normal(T)
    act = rng.integers(0, nA, size=T)
    q = rng.standard_normal((T, nA)).astype(np.float32)
    done = rng.random(T) < 0.08
    for s in range(streams):  # a done must never meet an empty window (the reference raises there)
        lo, hi = offs[s], offs[s + 1]
        done[lo] = False
        for t in range(lo + 1, hi):
            if done[t] and done[t - 1]:
                done[t] = False
    out = nstep.nstep_assemble_device(torch.as_tensor(rew, device="cuda"), torch.as_tensor(act, device="cuda"),
                                      torch.as_tensor(done, device="cuda"), torch.as_tensor(q, device="cuda"), n, gamma,
                                      offsets=torch.as_tensor(offs, device="cuda"))
    want = {k: [] for k in ("first", "next", "action", "ret", "done", "prio", "stream")}
    for s in range(streams):
        lo, hi = offs[s], offs[s + 1]
        batch, prios = oracle_stream(rew[lo:hi], act[lo:hi], done[lo:hi], q[lo:hi], n, gamma, first_step=lo)
        want["first"] += list(batch[0]); want["next"] += list(batch[3]); want["action"] += list(batch[1])
        want["ret"] += list(batch[2]); want["done"] += [float(x) for x in batch[4]]; want["prio"] += list(prios)
        want["stream"] += [s] * len(batch[0])
        assert int(out["count"][s]) == len(batch[0])
    for k in ("first", "next", "action", "stream"):
        assert out[k].cpu().tolist() == [int(x) for x in want[k]], k
    assert np.array_equal(bits(out["ret"].cpu().numpy()), bits(want["ret"]))
    assert np.array_equal(out["done"].cpu().numpy(), np.asarray(want["done"], dtype=np.float32))
    assert np.array_equal(bits(out["prio"].cpu().numpy()), bits(want["prio"]))


def test_nstep_preroll_and_empty_window(nstep):
    """emit_from: continue a stream from the middle with the earlier steps as pre-roll; and the
    reference's IndexError when done meets an empty window."""
    rng = np.random.default_rng(9)
    n, nA, T, cut = 4, 3, 200, 120
    rew = rng.standard_normal(T); act = rng.integers(0, nA, size=T)
    q = rng.standard_normal((T, nA)).astype(np.float32)
    done = np.zeros(T, dtype=bool); done[[30, 77, 78 + n + 3, 150]] = True
    dev = lambda a: torch.as_tensor(a, device="cuda")
    full = nstep.nstep_assemble_device(dev(rew), dev(act), dev(done), dev(q), n, 0.99)
    part = nstep.nstep_assemble_device(dev(rew), dev(act), dev(done), dev(q), n, 0.99,
                                       emit_from=torch.tensor([cut], device="cuda"))
    keep = full["next"] >= cut
    for k in ("first", "next", "action", "ret", "done", "prio"):
        assert torch.equal(part[k], full[k][keep]), k
    bad = np.zeros(T, dtype=bool); bad[[10, 11]] = True
    with pytest.raises(IndexError):
        nstep.nstep_assemble_device(dev(rew), dev(act), dev(bad), dev(q), n, 0.99)
