"""The replay server's side of the upstream multi-node Ape-X (origin_repo/replay.py) over the GPU buffer
(SURVEY.md section 8f row 1).

origin_repo/replay.py runs N asyncio workers of three kinds (:77-146) that each do
`async with lock: await loop.run_in_executor(exe, push_batch | update_prios | sample_batch, buffer, ...)`:
ONE lock (:167) held for the whole of every operation — unpickling a 50-row chunk, 50 Python `add`s of two tree walks
each, the 512 descents of a sample, the pickling of a 29 MB batch — which is the bottleneck its README names
(origin_repo/README.md:11).  Two things are offered here:

  * `push_batch`, `update_prios`, `sample_batch` with the reference's signatures and wire format
    (pickle of (batch, prios) / (idxes, prios) / the 7-tuple), so that origin_repo/replay.py itself runs unchanged on
    top of ape-x_b200/memory.py (an actor chunk then reaches the GPU as one `add_chunk` instead of <= 50 `add`s);

  * `ReplayServer`, which is what replaces the global lock: everything that touches only the message (unpickling,
    stacking the rows, priority ** alpha, pickling the reply) runs in the calling thread with no lock at all, and a
    short `threading.Lock` is held only around the native calls that enqueue work for the buffer — the GPU stream
    serialises the operations themselves, in the order the lock was taken, which is the sequential semantics of the
    reference's lock.  Workers therefore overlap their host work, and the buffer is busy only for the device time of
    each operation (tens of microseconds for an ingest or an update, the PCIe copy for a host-bound sample).
"""
import _pickle as pickle
import random
import threading

import numpy as np


def push_batch(buffer, data):
    """origin_repo/replay.py:19-26.  data = pickle((batch, prios)) as sent by origin_repo/actor.py:105-109, batch =
    [states, actions, rewards, next_states, dones] (memory.py:466-469)."""
    batch, prios = pickle.loads(data)
    if hasattr(buffer, "add_chunk"):
        buffer.add_chunk(batch[0], batch[1], batch[2], batch[3], batch[4], prios)
    else:
        for sample in zip(*batch, prios):
            buffer.add(*sample)


def update_prios(buffer, data):
    """origin_repo/replay.py:29-35.  data = pickle((idxes, prios)) as sent by origin_repo/learner.py (prios sender)."""
    idxes, prios = pickle.loads(data)
    buffer.update_priorities(idxes, prios)


def sample_batch(buffer, batch_size, beta):
    """origin_repo/replay.py:38-45 -> pickle of the 7-tuple buffer.sample returns (memory.py:298)."""
    return pickle.dumps(buffer.sample(batch_size, beta))


class ReplayServer(object):
    """The three operations for concurrent worker threads (see the module docstring).  `log`, when given, receives
    (sequence number, kind, payload) for every operation in the order it was enqueued — the order that defines the
    result — which is how tests replay the same sequence through the oracle."""

    def __init__(self, buffer, n_senders=8, log=None):
        self.buffer = buffer
        self._enqueue = threading.Lock()
        self._seq = 0
        self._log = log
        # sample() hands out views of a ring of pinned slots: one slot per concurrent sender plus one in flight
        buffer._host.RING = max(buffer._host.RING, n_senders + 1)

    def _note(self, kind, payload):
        if self._log is not None:
            self._log.append((self._seq, kind, payload))
        self._seq += 1

    def push_batch(self, data):
        batch, prios = pickle.loads(data)                                   # no lock: message only
        if len(batch[1]) == 0:
            return
        chunk = self.buffer.prepare_chunk(batch[0], batch[1], batch[2], batch[3], batch[4], prios)  # no lock: host arithmetic
        with self._enqueue:                                                 # upload + one launch
            self.buffer.commit_chunk(chunk)
            self._note("push", chunk)

    def update_prios(self, data):
        idxes, prios = pickle.loads(data)
        idx, pr = np.ascontiguousarray(idxes, dtype=np.int64), np.ascontiguousarray(prios, dtype=np.float64)
        with self._enqueue:                                                 # validation, pow, upload, launch (stream-ordered)
            self.buffer.update_priorities(idx, pr)
            self._note("update", (idx, pr))

    def sample_batch(self, batch_size, beta):
        with self._enqueue:                                                 # draws (global RNG, memory.py:247) + kernels + PCIe copy
            u = [random.random() for _ in range(batch_size)] if self._log is not None else None
            if u is not None:
                it = iter(u)
                saved, self.buffer._draw_uniforms = self.buffer._draw_uniforms, (lambda n: np.array([next(it) for _ in range(n)]))
            try:
                out = self.buffer.sample(batch_size, beta)
            finally:
                if u is not None:
                    self.buffer._draw_uniforms = saved
            self._note("sample", (batch_size, beta, u, list(out[-1])))
        return pickle.dumps(out)                                            # no lock: the reply's 29 MB pickle
