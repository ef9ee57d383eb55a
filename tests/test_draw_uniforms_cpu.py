"""memory.draw_uniforms: the one-call form of `[random.random() for _ in range(n)]` (memory.py:247) must give the same
bits and leave the generator in the same state, and must step aside for anything that is not the stock generator."""
import os
import random
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ape-x_b200")]
import memory


@pytest.mark.parametrize("n", [1, 2, 31, 512, 4096])
@pytest.mark.parametrize("seed", [0, 7, 123456789])
def test_same_values_and_same_state_as_n_calls(n, seed):
    random.seed(seed)
    random.random()  # (not at a block boundary of the generator)
    want = np.array([random.random() for _ in range(n)], dtype=np.float64)
    state_after = random.getstate()
    nxt = random.random()
    random.seed(seed)
    random.random()
    got = memory.draw_uniforms(n)
    assert got.dtype == np.float64 and got.shape == (n,)
    assert np.array_equal(got.view(np.uint64), want.view(np.uint64))
    assert random.getstate() == state_after
    assert random.random() == nxt


def test_stand_in_generators_are_called_one_by_one(monkeypatch):
    queue = [0.75, 0.5, 0.25]
    monkeypatch.setattr(memory.random, "random", lambda: queue.pop())
    assert memory.draw_uniforms(3).tolist() == [0.25, 0.5, 0.75]


def test_subclass_instances_are_not_second_guessed(monkeypatch):
    class Fixed(random.Random):
        def random(self):
            return 0.125

    monkeypatch.setattr(memory.random, "random", Fixed().random)
    assert memory.draw_uniforms(4).tolist() == [0.125] * 4


def test_zero():
    assert memory.draw_uniforms(0).shape == (0,)
