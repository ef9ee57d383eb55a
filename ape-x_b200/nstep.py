"""Device-side n-step assembly: BatchStorage (memory.py:393-478) for many actor streams at once.

`memory.BatchStorage` stays the per-step, host-side mirror an actor keeps next to its
environment.  When whole trajectory chunks are already on the GPU (vectorised actors, or a
replay server ingesting chunks from many actors, origin_repo/actor.py:79-115) this module
assembles all their n-step transitions and actor priorities in one launch, ready for
`add_batch_device(..., priorities=...)`.
"""
import numpy as np
import torch

import per_native as _n


def gamma_powers(gamma, n_steps):
    """gamma ** k for k = 0..n, evaluated like memory.py:474 / :461 (CPython float pow)."""
    return np.array([float(gamma) ** k for k in range(n_steps + 1)], dtype=np.float64)


def nstep_assemble_device(rewards, actions, dones, q_values, n_steps, gamma=0.99, offsets=None, emit_from=None):
    """rewards float64 [T], actions int64 [T], dones bool/uint8 [T], q_values float32 [T, nA] — CUDA
    tensors holding one or several streams back to back; offsets int64 [S+1] (default: one stream).
    Returns a dict of CUDA tensors: count [S]; and per emitted transition, stream after stream:
    first (step index of states[e]), next (step index of next_states[e]), action, ret (float64
    n-step return), done (float32), prio (float64 actor priority), stream (stream number)."""
    dev = rewards.device
    T = int(rewards.numel())
    if offsets is None:
        offsets = torch.tensor([0, T], dtype=torch.int64, device=dev)
    S = int(offsets.numel()) - 1
    rewards = rewards.to(torch.float64).contiguous()
    actions = actions.to(torch.int64).contiguous()
    dones = dones.to(torch.uint8).contiguous()
    q_values = q_values.to(torch.float32).contiguous()
    gp = gamma_powers(gamma, n_steps)
    first = torch.empty(T, dtype=torch.int64, device=dev)
    nxt = torch.empty(T, dtype=torch.int64, device=dev)
    act = torch.empty(T, dtype=torch.int64, device=dev)
    ret = torch.empty(T, dtype=torch.float64, device=dev)
    dn = torch.empty(T, dtype=torch.float32, device=dev)
    prio = torch.empty(T, dtype=torch.float64, device=dev)
    count = torch.zeros(S, dtype=torch.int64, device=dev)
    status = torch.zeros(1, dtype=torch.int64, device=dev)
    _n.check(_n.lib.nstep_assemble(dev.index, S, _n.ptr(offsets), _n.ptr(emit_from), _n.ptr(rewards), _n.ptr(actions),
                                   _n.ptr(dones), _n.ptr(q_values), int(q_values.shape[1]), int(n_steps), gp.ctypes.data,
                                   _n.ptr(first), _n.ptr(nxt), _n.ptr(act), _n.ptr(ret), _n.ptr(dn), _n.ptr(prio),
                                   _n.ptr(count), _n.ptr(status), _n.stream_ptr()))
    if int(status.item()) & _n.STATUS_NSTEP_EMPTY:
        raise IndexError("done arrived on an empty n-step window (memory.py:417 raises IndexError: deque index out of range)")
    # compact: slot = offsets[s] + e for e < count[s]
    pos = torch.arange(T, device=dev)
    stream_of = torch.bucketize(pos, offsets[1:], right=True).clamp_(max=S - 1)
    keep = (pos - offsets[stream_of]) < count[stream_of]
    sel = lambda t: t[keep]
    return {"count": count, "first": sel(first), "next": sel(nxt), "action": sel(act), "ret": sel(ret), "done": sel(dn),
            "prio": sel(prio), "stream": sel(stream_of)}
