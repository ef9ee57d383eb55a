"""Row (b) of SURVEY.md section 8: the drop-in boundary, proven by replaying what the reference's own drivers do.

tests/golden/make_dropin_trace.py ran ApeX.py, DQN.py, AQL.py and AQL_dis.py UNMODIFIED (with their own model.py,
utils.py, batchrecorder.py, batchrecoder_AQL.py; stub gym / tensorboardX) over the reference's memory.py and recorded
every buffer call with its arguments, uniform draws and results.  The reference cannot travel to the GPU box, the
trace can: here the same call sequence goes through ape-x_b200/memory.py with the same argument types the drivers
pass (float `size`, numpy float32 priorities, Python / numpy scalars, lists of indices) and must give the reference's
answers — sampled indices exactly, importance weights to 1e-5 (float64 both sides), batch columns equal after the
conversions the drivers apply (`torch.FloatTensor(x)`, `torch.LongTensor(x)`, ApeX.py:50-55), and at the end the same
length, cursor, running max priority and — bit for bit — the same tree leaves, roots and nodes."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import per_oracle as po

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def memory():
    import __graft_entry__ as entry
    entry.build()
    import memory as m
    return m


def bits(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64)).view(np.uint64)


@pytest.mark.parametrize("driver", ["apex", "dqn", "aql", "aql_dis", "replay_server"])
def test_driver_trace_replays_on_the_dropin_module(memory, monkeypatch, driver):
    """replay_server: the trace of origin_repo/replay.py's own worker coroutines (recv_batch / recv_prios / send_batch,
    replay.py:77-146, run unmodified over in-memory sockets); here the same wire messages — pickled (batch, prios) chunks,
    (idxes, prios) pairs, batch requests — go through replay_ops.ReplayServer, the lock-free-host-work version of the
    server's three operations, and the pickled replies must carry the reference's batches."""
    import _pickle as pickle
    g = np.load(os.path.join(GOLDEN, "dropin_trace.npz"), allow_pickle=False)
    p = driver + "_"
    cls = getattr(memory, str(g[p + "cls"]))
    buf = cls(float(g[p + "size"]), float(g[p + "alpha"]))  # ApeX.py:16 / AQL_dis.py:44 pass float sizes (1e6, 1e7)
    server = None
    if driver == "replay_server":
        import replay_ops
        server = replay_ops.ReplayServer(buf, n_senders=1)
        chunk_rows = int(g["replay_server_chunk_rows"])
    queue = []
    monkeypatch.setattr(memory.random, "random", lambda: queue.pop())
    has_extra = (p + "add_extra") in g.files
    n_cols = sum(1 for k in g.files if k.startswith(p + "s_col") and k[len(p) + 5:].isdigit())
    i_add = i_s = i_u = 0
    for kind, n, beta in zip(g[p + "ev_kind"], g[p + "ev_n"], g[p + "ev_beta"]):
        n = int(n)
        if kind == 0 and server is not None:
            for lo in range(i_add, i_add + n, chunk_rows):  # one wire message per actor chunk (origin_repo/actor.py:105-109)
                sl = slice(lo, min(lo + chunk_rows, i_add + n))
                batch = [list(g[p + "add_obs"][sl]), [int(a) for a in g[p + "add_act"][sl]], [float(r) for r in g[p + "add_rew"][sl]],
                         list(g[p + "add_next"][sl]), [np.float32(d) for d in g[p + "add_done"][sl]]]
                server.push_batch(pickle.dumps((batch, g[p + "add_extra"][sl])))
            i_add += n
        elif kind == 0:
            for j in range(i_add, i_add + n):
                row = [g[p + "add_obs"][j], int(g[p + "add_act"][j]), float(g[p + "add_rew"][j]), g[p + "add_next"][j],
                       bool(g[p + "add_done"][j])]
                if has_extra:
                    e = g[p + "add_extra"][j]
                    row.append(np.float64(e) if e.ndim == 0 else e)  # actor priority (ApeX) or the candidate actions a_mu (AQL)
                buf.add(*row)
            i_add += n
        elif kind == 1:
            queue[:] = list(g[p + "s_u"][i_s:i_s + n])[::-1]
            assert len(buf) == int(g[p + "s_len"][np.searchsorted(np.cumsum(g[p + "ev_n"][g[p + "ev_kind"] == 1]), i_s + n)])
            out = buf.sample(n, float(beta)) if server is None else pickle.loads(server.sample_batch(n, float(beta)))
            assert not queue and len(out) == n_cols + 2
            assert list(out[-1]) == list(g[p + "s_idx"][i_s:i_s + n]), "sampled indices differ from the reference's"
            np.testing.assert_allclose(np.asarray(out[-2], dtype=np.float64), g[p + "s_w"][i_s:i_s + n], rtol=1e-5)
            for c in range(n_cols):
                want = g[p + "s_col%d" % c][i_s:i_s + n]
                conv = torch.LongTensor if c == 1 else torch.FloatTensor  # what the drivers do with every column
                assert torch.equal(conv(out[c]), conv(want)), "column %d" % c
            i_s += n
        else:
            idx = [int(k) for k in g[p + "u_idx"][i_u:i_u + n]]           # the drivers hand back sample()'s list
            pr = g[p + "u_prio"][i_u:i_u + n]                             # float32 numpy, as utils.py:77 produces it
            assert pr.dtype == np.dtype(str(g[p + "u_prio_dtype"]))
            if server is None:
                buf.update_priorities(idx, pr)
            else:
                server.update_prios(pickle.dumps((idx, pr)))
            i_u += n
    assert i_add == g[p + "add_act"].shape[0] and i_s == g[p + "s_idx"].shape[0] and i_u == g[p + "u_idx"].shape[0]
    assert len(buf) == int(g[p + "final_len"]) and int(buf._next_idx) == int(g[p + "final_next_idx"])
    assert buf._max_priority == float(g[p + "final_max_priority"])
    cap, n = int(g[p + "final_cap"]), int(g[p + "final_len"])
    sv, mv = buf._it_sum._value, buf._it_min._value
    assert sv.shape[0] == 2 * cap
    assert np.array_equal(bits(sv[cap:cap + n]), bits(g[p + "final_leaves_sum"]))
    assert np.array_equal(bits(mv[cap:cap + n]), bits(g[p + "final_leaves_min"]))
    assert bits(sv[1]) == bits(g[p + "final_root_sum"]) and bits(mv[1]) == bits(g[p + "final_root_min"])
    assert np.all(sv[cap + n:] == 0.0) and np.all(np.isinf(mv[cap + n:]))
    assert np.array_equal(bits(sv)[1:], bits(po.build_tree(sv[cap:], "sum"))[1:])
    assert np.array_equal(bits(mv)[1:], bits(po.build_tree(mv[cap:], "min"))[1:])
