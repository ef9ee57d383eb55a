"""BASELINE configs[1] as a whole learner step (the metric's second half, "Ape-X learner updates/s"):
dueling Nature-CNN DQN (the reference's architecture, model.py:14-66: conv 32x8/4, 64x4/2, 64x3/1, two
128-unit heads), batch 512, n-step 3, double-Q, RMSprop, replay of 2^17 Atari-shaped transitions in HBM.
The Q-network is plain PyTorch on both sides (out of scope); what differs is the replay:

  host   the reference's loop verbatim (ApeX.py:46-63): buffer.sample() -> numpy -> torch.FloatTensor(...)
         .to(device) -> compute_loss -> update_parameters -> buffer.update_priorities(idxes, numpy prios),
         through this repo's drop-in memory module (host arrays across PCIe both ways);
  device sample_device(obs_dtype=float32) (gather fused with the uint8 -> float32 widening) -> the same
         network step -> td_priority_update (TD error -> priority -> tree update in one launch); nothing
         leaves the GPU.
Prints learner updates/s and sample+update transitions/s for both.  Usage: python tools/learner_step.py"""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ape-x_b200")]
import numpy as np
import torch
import torch.nn as nn

import __graft_entry__ as entry

entry.build()
import memory

ap = argparse.ArgumentParser()
ap.add_argument("--ring", type=int, default=1 << 17)
ap.add_argument("--batch", type=int, default=512)
ap.add_argument("--steps", type=int, default=60)
args = ap.parse_args()
dev = torch.device("cuda")
B, NA, NSTEP, GAMMA = args.batch, 18, 3, 0.99


class Dueling(nn.Module):
    def __init__(self):
        super().__init__()
        self.features = nn.Sequential(nn.Conv2d(4, 32, 8, 4), nn.ReLU(), nn.Conv2d(32, 64, 4, 2), nn.ReLU(),
                                      nn.Conv2d(64, 64, 3, 1), nn.ReLU(), nn.Flatten())
        self.advantage = nn.Sequential(nn.Linear(3136, 128), nn.ReLU(), nn.Linear(128, NA))
        self.value = nn.Sequential(nn.Linear(3136, 128), nn.ReLU(), nn.Linear(128, 1))

    def forward(self, x):
        f = self.features(x)
        a = self.advantage(f)
        return self.value(f) + a - a.mean(1, keepdim=True)


def loss_and_tables(model, tgt, states, actions, rewards, next_states, dones, weights):
    """utils.compute_loss (utils.py:64-81) without the priority line; returns the Q tables for the fused kernel."""
    q = model(states)
    qn = model(next_states)
    with torch.no_grad():
        qt = tgt(next_states)
    q_a = q.gather(1, actions.unsqueeze(1)).squeeze(1)
    nxt = qt.gather(1, qn.max(1)[1].unsqueeze(1)).squeeze(1)
    td = torch.abs((rewards + (GAMMA ** NSTEP) * nxt * (1 - dones)).detach() - q_a)
    loss = (torch.where(td < 1, 0.5 * td ** 2, td - 0.5) * weights).mean()
    return loss, td, q, qn, qt


torch.manual_seed(0)
model, tgt = Dueling().to(dev), Dueling().to(dev)
tgt.load_state_dict(model.state_dict())
opt = torch.optim.RMSprop(model.parameters(), 6.25e-5, alpha=0.95, eps=1.5e-7, centered=True)

buf = memory.CustomPrioritizedReplayBuffer(args.ring, 0.6)
gen = torch.Generator(device=dev).manual_seed(0)
for lo in range(0, args.ring, 8192):
    m = min(8192, args.ring - lo)
    obs = torch.randint(0, 256, (m, 4, 84, 84), dtype=torch.uint8, device=dev, generator=gen)
    nxt = torch.randint(0, 256, (m, 4, 84, 84), dtype=torch.uint8, device=dev, generator=gen)
    buf.add_batch_device(obs, nxt, torch.randint(0, NA, (m,), device=dev, generator=gen),
                         torch.randint(-1, 2, (m,), device=dev, generator=gen).float(),
                         (torch.rand(m, device=dev, generator=gen) < 1e-3).float(),
                         priorities=torch.randn(m, dtype=torch.float64, device=dev, generator=gen).abs() + 1e-6)
del obs, nxt


def step_host():
    states, actions, rewards, next_states, dones, weights, idxes = buf.sample(B, 0.4)
    states = torch.FloatTensor(states).to(dev)
    actions = torch.LongTensor(actions).to(dev)
    rewards = torch.FloatTensor(rewards).to(dev)
    next_states = torch.FloatTensor(next_states).to(dev)
    dones = torch.FloatTensor(dones).to(dev)
    weights = torch.FloatTensor(weights).to(dev)
    loss, td, _, _, _ = loss_and_tables(model, tgt, states, actions, rewards, next_states, dones, weights)
    prios = (0.9 * torch.max(td) + 0.1 * td + 1e-6).data.cpu().numpy()
    opt.zero_grad()
    loss.backward()
    torch.nn.utils.clip_grad_norm_(model.parameters(), 40)
    opt.step()
    buf.update_priorities(idxes, prios)


def step_device():
    d = buf.sample_device(B, 0.4, obs_dtype=torch.float32)
    loss, _, q, qn, qt = loss_and_tables(model, tgt, d["obs"], d["act"], d["rew"], d["next_obs"], d["done"], d["weights"])
    opt.zero_grad()
    loss.backward()
    torch.nn.utils.clip_grad_norm_(model.parameters(), 40)
    opt.step()
    buf.td_priority_update(q.detach(), qn.detach(), qt, d["act"], d["rew"], d["done"], d["weights"], NSTEP, GAMMA, idx=d["idx"])


def replay_only_device():
    d = buf.sample_device(B, 0.4, obs_dtype=torch.float32)
    return d


for name, fn in (("host  (reference loop, drop-in memory module)", step_host), ("device (sample_device + td_priority_update)", step_device)):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / args.steps
    print("%-48s %7.2f ms per learner step = %7.1f learner updates/s = %.3f M sample+update transitions/s"
          % (name, dt * 1e3, 1.0 / dt, B / dt / 1e6))
# the network alone on resident float32 batches, for reference
d = buf.sample_device(B, 0.4, obs_dtype=torch.float32)


def net_only():
    loss, _, _, _, _ = loss_and_tables(model, tgt, d["obs"], d["act"], d["rew"], d["next_obs"], d["done"], d["weights"])
    opt.zero_grad()
    loss.backward()
    torch.nn.utils.clip_grad_norm_(model.parameters(), 40)
    opt.step()


for _ in range(5):
    net_only()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(args.steps):
    net_only()
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / args.steps
print("%-48s %7.2f ms" % ("network step alone (PyTorch fp32, resident batch)", dt * 1e3))
