"""K6 — fused AQL candidate-Q + argmax against a plain PyTorch float32 restatement of
Q_Network.act (model.py:294-335) with NoisyLinear in training and eval mode.  Tolerance: 1e-5
relative to the Q scale (BASELINE.json north_star: TD errors / priorities within 1e-5 in
fp32); the argmax must agree wherever the reference's top-2 gap exceeds that tolerance."""
import math

import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def aql():
    import __graft_entry__ as entry
    entry.build()
    import aql as m
    return m


class Noisy(nn.Module):
    """NoisyLinear, model.py:112-164 (factorised Gaussian noise, std_init 0.4)."""

    def __init__(self, i, o):
        super().__init__()
        r = 1 / math.sqrt(i)
        self.weight_mu = nn.Parameter(torch.empty(o, i).uniform_(-r, r))
        self.weight_sigma = nn.Parameter(torch.full((o, i), 0.4 / math.sqrt(i)))
        self.bias_mu = nn.Parameter(torch.empty(o).uniform_(-r, r))
        self.bias_sigma = nn.Parameter(torch.full((o,), 0.4 / math.sqrt(o)))
        sc = lambda n: torch.randn(n).sign() * torch.randn(n).abs().sqrt()
        self.register_buffer("weight_epsilon", sc(o).ger(sc(i)))
        self.register_buffer("bias_epsilon", sc(o))

    def forward(self, x):
        if self.training:
            return F.linear(x, self.weight_mu + self.weight_sigma * self.weight_epsilon, self.bias_mu + self.bias_sigma * self.bias_epsilon)
        return F.linear(x, self.weight_mu, self.bias_mu)


class QRef(nn.Module):
    """The layers Q_Network.act touches (model.py:244-281) and its forward (model.py:294-335)."""

    def __init__(self, obs, A, continuous):
        super().__init__()
        self.continuous, self.A = continuous, A
        self.q_feature = nn.Sequential(nn.Linear(obs, 64), nn.ReLU(), nn.Linear(64, 64), nn.ReLU())
        self.action_out = (nn.Sequential(nn.Linear(A, 128), nn.ReLU(), nn.Linear(128, 64), nn.ReLU()) if continuous
                           else nn.Sequential(nn.Linear(1, 64), nn.ReLU()))
        self.advantage1, self.advantage2 = Noisy(128, 64), Noisy(64, 1)

    def act(self, state, a_mu, K):
        if self.continuous:
            a_out = self.action_out(a_mu.reshape(-1, self.A)).reshape(-1, K, 64)
        else:
            a_out = self.action_out(a_mu.reshape(-1, 1).float()).reshape(-1, K, 64)
        q_f = self.q_feature(state).repeat(1, K).reshape(-1, K, 64)
        x = F.relu(torch.cat([a_out, q_f], dim=2)).reshape(-1, 128)
        q = self.advantage2(F.relu(self.advantage1(x)))
        return q.reshape(a_mu.shape[0], K)


@pytest.mark.parametrize("continuous,A,B,K,train", [(True, 1, 3, 37, True), (True, 4, 16, 200, False), (True, 2, 256, 1000, True),
                                                   (False, 1, 5, 130, True), (False, 1, 256, 1000, False), (True, 3, 1, 1, True)])
def test_candidate_q_matches_float32_reference(aql, continuous, A, B, K, train):
    torch.manual_seed(B * 1000 + K)
    torch.backends.cuda.matmul.allow_tf32 = False
    obs = 11
    ref = QRef(obs, A, continuous).cuda()
    ref.train(train)
    state = torch.randn(B, obs, device="cuda")
    a_mu = (torch.rand(B, K, A, device="cuda") * 4 - 2) if continuous else torch.randint(0, 18, (B, K), device="cuda")
    with torch.no_grad():
        want = ref.act(state, a_mu, K)
        f = ref.q_feature(state)
    cq = aql.CandidateQ(continuous, A)
    cq.set_weights_from(ref)
    got, am = cq.forward(f, a_mu)
    torch.cuda.synchronize()
    scale = float(want.abs().max())
    err = float((got - want).abs().max())
    assert err <= 1e-5 * max(scale, 1.0), "max |dq| = %.3g at Q scale %.3g" % (err, scale)
    # argmax: agree wherever the reference's best beats its runner-up by more than the tolerance
    top2 = want.topk(min(2, K), dim=1).values
    clear = torch.ones(B, dtype=torch.bool, device="cuda") if K == 1 else (top2[:, 0] - top2[:, 1]) > 4e-5 * max(scale, 1.0)
    assert torch.equal(am[clear], want.argmax(1)[clear])
    assert torch.equal(am, got.argmax(1))  # and it is the argmax of what the kernel itself produced (first max wins)
