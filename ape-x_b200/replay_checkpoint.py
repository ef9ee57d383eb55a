"""Snapshot / restore of a GPU replay buffer.

The reference only ever saves model weights (ApeX.py:74-80); a long Ape-X run that stops loses
its replay.  `save` writes the ring rows, both trees' leaves, the cursor, len and the running max
priority; `load` re-inserts the rows with the saved leaves, which rebuilds both trees bit for bit
(every internal node is a function of the leaves).  Works through the public entry points only:
rows leave the GPU with per_gather over consecutive slots and come back with per_add_batch.
"""
import numpy as np
import torch

import per_native as _n

_CHUNK = 4096


def save(buffer, path):
    """buffer: a memory.PrioritizedReplayBuffer (or subclass).  Writes one .npz file."""
    buffer._flush()
    n, cap = len(buffer), buffer._tree_capacity
    buffer._sync_max_priority()
    out = {"size": buffer._ring, "alpha": float(buffer._alpha), "len": n, "cursor": int(buffer._next_idx),
           "max_priority": float(buffer._max_priority), "cls": type(buffer).__name__}
    if n:
        cols = {"obs": [], "act": [], "rew": [], "next_obs": [], "done": [], "a_mu": []}
        for lo in range(0, n, _CHUNK):
            idx = torch.arange(lo, min(n, lo + _CHUNK), dtype=torch.int64, device=buffer._device)
            obs, act, rew, nxt, done, extra = buffer._gather_device(idx)
            for k, t in zip(("obs", "act", "rew", "next_obs", "done", "a_mu"), (obs, act, rew, nxt, done, extra)):
                if t is not None:
                    cols[k].append(t.cpu().numpy())
        for k, parts in cols.items():
            if parts:
                out[k] = np.concatenate(parts)
        out["leaves"] = buffer.tree_values(_n.TREE_SUM)[cap:cap + n].copy()
    np.savez(path, **out)


def load(path, cls_map, device=None):
    """Returns a new buffer equal to the saved one.  cls_map: the `memory` module (class by name)."""
    with np.load(path, allow_pickle=False) as g:
        cls = getattr(cls_map, str(g["cls"]))
        buf = cls(int(g["size"]), float(g["alpha"]), device) if device is not None else cls(int(g["size"]), float(g["alpha"]))
        n = int(g["len"])
        if n:
            dev = buf._device
            for lo in range(0, n, _CHUNK):
                hi = min(n, lo + _CHUNK)
                up = lambda k: torch.from_numpy(np.ascontiguousarray(g[k][lo:hi])).to(dev)
                extra = up("a_mu") if "a_mu" in g.files else None
                buf.add_batch_device(up("obs"), up("next_obs"), up("act"), up("rew"), up("done"), extra=extra,
                                     leaves=up("leaves"), batch_max=0.0)
            _n.check(_n.lib.per_set_cursor(buf._h, int(g["cursor"]), float(g["max_priority"]), _n.stream_ptr()))
            buf._cursor = int(g["cursor"])
        buf._max_priority = float(g["max_priority"])
        buf._max_priority_stale = False
    return buf
