"""Host logic of the batchrecorder mirrors on the CPU: the worker protocol (record_batch /
set_pi_weights / cleanup over a joinable queue), the n-step chunks an actor ships, and the
parent-side ingest — against the oracle's BatchStorage restatement.  gym and the reference's
model.py are not part of this repo (nor installed here), so tiny stand-ins are registered under
those module names before the workers are forked."""
import sys
import time
import types

import numpy as np
import pytest
import torch

from oracle import per_oracle as po


class ChainEnv(object):
    """Deterministic episodic env: obs = [t, seed]; the episode ends after 23 + seed steps."""

    def __init__(self, env_id, slow=False):
        self.s, self.t, self.slow = 0, 0, slow

    def seed(self, s):
        self.s = int(s)

    def reset(self):
        if self.slow:  # keeps a worker busy long enough that the pool's other task goes to the other worker
            time.sleep(0.3)
        self.t = 0
        return np.array([0.0, self.s], dtype=np.float32)

    def step(self, action):
        self.t += 1
        done = self.t >= 23 + self.s % 5
        return np.array([self.t, self.s], dtype=np.float32), float(self.t % 3 - 1), done, {}

    def close(self):
        pass


class TinyDQN(torch.nn.Module):
    def __init__(self, env):
        super().__init__()
        self.w = torch.nn.Parameter(torch.zeros(1))

    def act(self, state, epsilon):
        t = float(state[0])
        q = np.array([np.sin(t), np.cos(t), 0.1 * t + float(self.w.detach())], dtype=np.float32)
        return int(t) % 3, q


@pytest.fixture()
def stubs(monkeypatch):
    gym = types.ModuleType("gym")
    gym.make = lambda env_id: ChainEnv(env_id, slow=True)
    model = types.ModuleType("model")
    model.DuelingDQN = TinyDQN
    monkeypatch.setitem(sys.modules, "gym", gym)
    monkeypatch.setitem(sys.modules, "model", model)
    sys.modules.pop("batchrecorder", None)
    import batchrecorder
    yield batchrecorder
    sys.modules.pop("batchrecorder", None)


def expected_rows(seed, n_steps, gamma, send_interval=50):
    env, net = ChainEnv("x"), TinyDQN(None)
    env.seed(seed)
    ns = po.NStepOracle(n_steps, gamma)
    rows, state = [], env.reset()
    while True:
        a, q = net.act(torch.FloatTensor(state), 0.0)
        nxt, r, done, _ = env.step(a)
        ns.add(state, r, a, done, q)
        state = nxt
        if done or len(ns) == send_interval:
            batch, prios = ns.make_batch()
            rows += [tuple(col[i] for col in batch) + (prios[i],) for i in range(len(prios))]
            ns.reset()
        if done:
            return rows


class ListBuffer(object):
    def __init__(self):
        self.rows = []

    def add(self, s, a, r, s2, d, p):
        self.rows.append((s, a, r, s2, d, p))


class ChunkBuffer(ListBuffer):
    def add_chunk(self, states, actions, rewards, next_states, dones, prios):
        self.rows += list(zip(states, actions, rewards, next_states, dones, prios))


@pytest.mark.parametrize("buffer_cls", [ListBuffer, ChunkBuffer])
def test_record_batch_ships_reference_chunks(stubs, buffer_cls):
    br = stubs
    buf = buffer_cls()
    rec = br.BatchRecorder(env_id="chain", env_seed=3, n_workers=2, buffer=buf, n_steps=3, gamma=0.99, max_episode_length=1000)
    try:
        rec.record_batch()
        want = sorted(expected_rows(3, 3, 0.99) + expected_rows(4, 3, 0.99), key=lambda r: (r[0][1], r[0][0]))
        got = sorted(buf.rows, key=lambda r: (r[0][1], r[0][0]))
        assert len(got) == len(want) > 20
        for g, w in zip(got, want):
            assert np.array_equal(g[0], w[0]) and np.array_equal(g[3], w[3]) and g[1] == w[1]
            assert g[2] == w[2] and float(g[4]) == float(w[4]) and g[5] == w[5]  # returns and priorities bit-equal
        pi = TinyDQN(None)
        with torch.no_grad():
            pi.w.fill_(2.5)
        rec.set_worker_weights(pi)  # new weights change the Q values, hence the priorities
        buf.rows.clear()
        rec.record_batch()
        assert len(buf.rows) == len(want) and any(g[5] != w[5] for g, w in zip(sorted(buf.rows, key=lambda r: (r[0][1], r[0][0])), want))
    finally:
        rec.cleanup()
    assert [round(w.epsilon, 12) for w in rec.workers] == [round(0.4 ** 1, 12), round(0.4 ** 8, 12)]


def test_single_worker_epsilon_does_not_divide_by_zero(stubs):
    import _actor_pool
    assert _actor_pool.exploration_epsilon(0, 1) == 0.4  # the reference raises ZeroDivisionError here (Q14)
    assert _actor_pool.exploration_epsilon(7, 8) == 0.4 ** 8


# ------------------------------------------------------------------------------------------
# AQL pool (batchrecoder_AQL.py:13-138): raw (s, a, r, s', d, a_mu) rows, no n-step assembly, and the
# parent inserts every row len(state) times (batchrecoder_AQL.py:121-123).
# ------------------------------------------------------------------------------------------
class TinyAQL(torch.nn.Module):
    """Stands in for model.AQL (model.py:169-207): act(state, eps) -> (action index, a_mu [1, K, A], q)."""
    K = 5

    def __init__(self, env, propose_sample=100, uniform_sample=400, action_var=0.25, device="cpu"):
        super().__init__()
        self.w = torch.nn.Parameter(torch.zeros(1))
        self.kw = (propose_sample, uniform_sample, action_var, device)

    def act(self, state, epsilon):
        t = float(state[0])
        a_mu = (np.arange(self.K, dtype=np.float32).reshape(1, self.K, 1) + t + float(self.w.detach())) / 10.0
        return int(t) % self.K, a_mu, None


class AqlChainEnv(ChainEnv):
    """obs = [t, seed, 0]: three components, so the reference's ingest loop adds every row three times."""

    def reset(self):
        super().reset()
        return np.array([0.0, self.s, 0.0], dtype=np.float32)

    def step(self, action):
        o, r, d, info = super().step(action)
        self.last_action = np.asarray(action)
        return np.array([o[0], o[1], float(self.last_action.ravel()[0])], dtype=np.float32), r, d, info


class Writer(object):
    def __init__(self):
        self.scalars = []

    def add_scalar(self, tag, value, step):
        self.scalars.append((tag, float(value), int(step)))


class AqlListBuffer(object):
    def __init__(self):
        self.rows = []

    def add(self, s, a, r, s2, d, a_mu):
        self.rows.append((s, a, r, s2, d, a_mu))


@pytest.fixture()
def aql_stubs(monkeypatch):
    gym = types.ModuleType("gym")
    gym.make = lambda env_id: AqlChainEnv(env_id, slow=True)
    model = types.ModuleType("model")
    model.AQL = TinyAQL
    tbx = types.ModuleType("tensorboardX")
    tbx.SummaryWriter = Writer
    for name, mod in (("gym", gym), ("model", model), ("tensorboardX", tbx)):
        monkeypatch.setitem(sys.modules, name, mod)
    sys.modules.pop("batchrecoder_AQL", None)
    yield
    sys.modules.pop("batchrecoder_AQL", None)


def aql_expected(seed):
    """The rows one worker ships for one episode and what the parent inserts from them."""
    env, net = AqlChainEnv("x"), TinyAQL(None)
    env.seed(seed)
    state, rows, ep_r = env.reset(), [], 0.0
    while True:
        action, a_mu, _ = net.act(state, 0.0)
        a_mu = a_mu[0]
        nxt, r, done, _ = env.step(a_mu[action])
        rows.append((state, action, r, nxt, done, a_mu))
        state, ep_r = nxt, ep_r + r
        if done:
            return rows, ep_r


def run_aql_recorder(module):
    buf, writer = AqlListBuffer(), Writer()
    rec = module.BatchRecorder(env_id="chain", env_seed=5, n_workers=2, buffer=buf, max_episode_length=1000, writer=writer,
                               propose_sample=2, uniform_sample=3, action_var=0.5, device="cpu")
    try:
        total = rec.record_batch()
        first = list(buf.rows)
        pi = TinyAQL(None)
        with torch.no_grad():
            pi.w.fill_(1.5)
        rec.set_worker_weights(pi)
        buf.rows.clear()
        total2 = rec.record_batch()
        return total, first, total2, list(buf.rows), list(writer.scalars), [w.epsilon for w in rec.workers]
    finally:
        rec.cleanup()


def check_aql_rows(got_rows, seeds):
    want = []
    for s in seeds:
        rows, _ = aql_expected(s)
        for row in rows:
            want += [row] * len(row[0])  # Q13: each transition len(state) == 3 times
    key = lambda r: (float(r[0][1]), float(r[0][0]))
    got, want = sorted(got_rows, key=key), sorted(want, key=key)
    assert len(got) == len(want) > 60
    for g, w in zip(got, want):
        assert np.array_equal(g[0], w[0]) and np.array_equal(g[3], w[3]) and g[1] == w[1] and g[2] == w[2] and g[4] == w[4]
        assert np.array_equal(g[5], w[5]) and g[5].shape == (TinyAQL.K, 1)


def test_aql_record_batch_duplicates_rows_like_the_reference(aql_stubs):
    import batchrecoder_AQL as mirror
    total, first, total2, second, scalars, eps = run_aql_recorder(mirror)
    lens = {s: len(aql_expected(s)[0]) for s in (5, 6)}
    assert total == total2 == lens[5] + lens[6]          # record_batch() returns the summed episode lengths
    check_aql_rows(first, (5, 6))
    assert len(second) == len(first) and not all(np.array_equal(a[5], b[5]) for a, b in zip(
        sorted(second, key=lambda r: (float(r[0][1]), float(r[0][0]))), sorted(first, key=lambda r: (float(r[0][1]), float(r[0][0])))))
    tags = [t for t, _, _ in scalars]
    assert tags == ["actor/episode_reward", "actor/episode_length"] * 4 and [s for _, _, s in scalars] == [0, 0, 1, 1, 2, 2, 3, 3]
    assert sorted(v for t, v, _ in scalars[:4] if t == "actor/episode_length") == sorted(float(v) for v in lens.values())
    assert [round(e, 12) for e in eps] == [round(0.4, 12), round(0.4 ** 8, 12)]


REFERENCE = "/root/reference"


@pytest.mark.skipif(not __import__("os").path.exists(REFERENCE + "/batchrecoder_AQL.py"), reason="the reference tree is only present in the build container")
def test_aql_mirror_ships_what_the_unmodified_reference_pool_ships(aql_stubs, monkeypatch):
    """The reference's own batchrecoder_AQL.py, imported unedited (stub gym / model / tensorboardX as in the fixture),
    and the mirror feed the same buffer calls, writer calls and return values."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("reference_batchrecoder_AQL", REFERENCE + "/batchrecoder_AQL.py")
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    import batchrecoder_AQL as mirror
    r_ref, r_mir = run_aql_recorder(ref), run_aql_recorder(mirror)
    assert r_ref[0] == r_mir[0] and r_ref[2] == r_mir[2] and r_ref[5] == r_mir[5]
    key = lambda r: (float(r[0][1]), float(r[0][0]))
    for a, b in ((r_ref[1], r_mir[1]), (r_ref[3], r_mir[3])):
        a, b = sorted(a, key=key), sorted(b, key=key)
        assert len(a) == len(b)
        for x, y in zip(a, b):
            assert all(np.array_equal(np.asarray(p), np.asarray(q)) for p, q in zip(x, y))
    # (which worker's episode gets which episode index depends on the order the results arrive in)
    assert sorted((t, v) for t, v, _ in r_ref[4]) == sorted((t, v) for t, v, _ in r_mir[4])
    assert sorted(s for _, _, s in r_ref[4]) == sorted(s for _, _, s in r_mir[4])


@pytest.mark.skipif(not __import__("os").path.exists(REFERENCE + "/batchrecorder.py"), reason="the reference tree is only present in the build container")
def test_unmodified_reference_batchrecorder_runs_on_the_dropin_memory_module(stubs):
    """batchrecorder.py of the reference, imported unedited: its `from memory import BatchStorage` (batchrecorder.py:8)
    resolves to ape-x_b200/memory.py, its workers assemble n-step chunks with that BatchStorage, and the parent's
    ingest loop (batchrecorder.py:133-137) feeds buffer.add.  Rows, returns and actor priorities must be the oracle's."""
    import importlib.util
    import memory as dropin
    assert dropin.__file__.endswith("ape-x_b200/memory.py")
    tbx = types.ModuleType("tensorboardX")  # imported at batchrecorder.py:9, used only in commented-out lines
    tbx.SummaryWriter = object
    sys.modules.setdefault("tensorboardX", tbx)
    spec = importlib.util.spec_from_file_location("reference_batchrecorder", REFERENCE + "/batchrecorder.py")
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    assert ref.BatchStorage is dropin.BatchStorage
    buf = ListBuffer()
    rec = ref.BatchRecorder(env_id="chain", env_seed=3, n_workers=2, buffer=buf, n_steps=3, gamma=0.99, max_episode_length=1000)
    try:
        rec.record_batch()
    finally:
        rec.cleanup()
    want = sorted(expected_rows(3, 3, 0.99) + expected_rows(4, 3, 0.99), key=lambda r: (r[0][1], r[0][0]))
    got = sorted(buf.rows, key=lambda r: (r[0][1], r[0][0]))
    assert len(got) == len(want) > 20
    for g, w in zip(got, want):
        assert np.array_equal(g[0], w[0]) and np.array_equal(g[3], w[3]) and g[1] == w[1]
        assert g[2] == w[2] and float(g[4]) == float(w[4]) and g[5] == w[5]
