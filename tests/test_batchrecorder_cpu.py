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
