"""Drop-in replacement for the reference's `memory.py` (same class names, constructor and
method signatures, return shapes and AssertionErrors), backed by the HBM-resident trees and
transition ring of libper_b200.so.  Put this directory first on sys.path and
`from memory import CustomPrioritizedReplayBuffer` in ApeX.py / DQN.py / AQL.py /
AQL_dis.py / batchrecorder.py resolves here.

What lives where
  * device (B200): both float64 segment trees, the observation / next-observation / action /
    reward / done / a_mu ring, every tree update, every prefix-sum descent, the importance
    weights and the row gather;
  * host (this file): argument checks the reference asserts, the `random.random()` draws
    (same global RNG, same order as memory.py:247), `priority ** alpha` through the C
    library's pow for the host-facing API (bit-identical leaves, see per_pow_alpha_host),
    and a small staging list so that one-at-a-time `add` calls reach the GPU in batches.

Each class / method cites the reference lines it mirrors.  Device-side extensions that keep
everything on the GPU (`sample_device`, `update_priorities_device`, `td_priority_update`,
`add_batch_device`) are additional methods; the reference return contracts are unchanged.
There is no CPU fallback: without the built library, importing this module raises.
"""
import operator
import random
from collections import deque

import numpy as np
import torch

import per_native as _n

_FLUSH_ROWS = 1024  # staged `add` rows that trigger an upload


def _device_index(device=None):
    if device is None:
        if not torch.cuda.is_available():
            raise RuntimeError("ape-x_b200 needs a CUDA device (B200); there is no CPU fallback")
        return torch.cuda.current_device()
    d = torch.device(device)
    return d.index if d.index is not None else torch.cuda.current_device()


class _HostStage(object):
    """Persistent pinned host buffers for the host-facing API.  Uploads: the array is copied
    into a pinned buffer and sent with an async H2D copy; an event per buffer makes the next
    reuse wait for that copy.  Downloads: results land in one of RING pinned slots and are
    handed out as numpy views (valid until RING more downloads of the same name — the
    reference's callers convert them to tensors at once, ApeX.py:50-55); set
    `memory.COPY_SAMPLES = True` to get private copies instead."""
    RING = 2

    def __init__(self, device):
        self.device = device
        self.up_bufs = {}
        self.down_bufs = {}

    def up(self, name, array):
        """One pinned buffer per (name, row shape, dtype), grown to the largest row count seen: a caller that
        uploads 1, 2, 3, ... rows at a time (per-row add interleaved with sample requests) keeps reusing it
        instead of pinning a new block per size."""
        a = np.ascontiguousarray(array)
        n = a.shape[0] if a.ndim else 1
        key = (name, a.shape[1:], a.dtype.str)
        ent = self.up_bufs.get(key)
        if ent is not None:
            ent[1].synchronize()  # the previous copy out of this buffer has finished
        if ent is None or ent[0].shape[0] < n:
            cap = max(n, 2 * ent[0].shape[0] if ent is not None else 0)
            tdt = torch.from_numpy(np.empty(0, dtype=a.dtype)).dtype
            ent = [torch.empty((cap,) + a.shape[1:], dtype=tdt, pin_memory=True), torch.cuda.Event()]
            self.up_bufs[key] = ent
        ent[0].numpy()[:n] = a.reshape((n,) + a.shape[1:])
        d = ent[0][:n].to(self.device, non_blocking=True)
        ent[1].record()
        return d.reshape(a.shape)

    def down(self, named_tensors):
        """{name: cuda tensor} -> {name: numpy array}; one synchronisation for the lot."""
        outs = {}
        for name, t in named_tensors.items():
            key = (name, tuple(t.shape), str(t.dtype))
            ent = self.down_bufs.get(key)
            if ent is None:
                ent = [[torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for _ in range(self.RING)], 0]
                self.down_bufs[key] = ent
            slot = ent[0][ent[1] % self.RING]
            ent[1] += 1
            slot.copy_(t, non_blocking=True)
            outs[name] = slot
        torch.cuda.current_stream().synchronize()
        return {k: (np.array(v.numpy()) if COPY_SAMPLES else v.numpy()) for k, v in outs.items()}


COPY_SAMPLES = False


def draw_uniforms(n):
    """The n values `random.random()` would return next (memory.py:247 draws one per stratum), as a float64 array, the
    generator left in the state n calls would leave it in.  With the stock Mersenne Twister behind `random.random` they are
    taken in ONE call: random() is ((a >> 5) * 2**26 + (b >> 6)) / 2**53 of two successive 32-bit outputs a, b
    (CPython's _randommodule.c), and getrandbits(64 n) hands out the same outputs, least significant word first — same values,
    same state, a tenth of the time of 512 Python-level calls.  Anything else behind `random.random` (a test's stand-in, a
    subclass) is called n times."""
    rnd = random.random
    inst = getattr(rnd, "__self__", None)
    if n > 0 and type(inst) is random.Random and getattr(rnd, "__name__", "") == "random":
        words = np.frombuffer(inst.getrandbits(64 * n).to_bytes(8 * n, "little"), dtype="<u4")
        a = (words[0::2] >> 5).astype(np.float64)
        b = (words[1::2] >> 6).astype(np.float64)
        return (a * 67108864.0 + b) * (1.0 / 9007199254740992.0)
    return np.array([rnd() for _ in range(n)], dtype=np.float64)


def _pow_alpha_exact(prios64, alpha):
    """leaf = priority ** alpha with libm's double pow (what CPython evaluates at memory.py:317)."""
    prios64 = np.ascontiguousarray(prios64, dtype=np.float64)
    out = np.empty_like(prios64)
    _n.check(_n.lib.per_pow_alpha_host(prios64.ctypes.data, float(alpha), out.ctypes.data, prios64.size))
    return out


# ==========================================================================================
# segment trees (memory.py:10-143)
# ==========================================================================================
class SegmentTree(object):
    """memory.py:10-91.  `operation` must be operator.add (sum tree) or min (min tree) — the
    only two the reference ever builds (memory.py:96-100, 134-138)."""

    def __init__(self, capacity, operation, neutral_element, device=None):
        assert capacity > 0 and capacity & (capacity - 1) == 0, "capacity must be positive and a power of 2."
        if operation is operator.add:
            kind = _n.TREE_SUM
        elif operation is min:
            kind = _n.TREE_MIN
        else:
            raise NotImplementedError("the CUDA tree implements operator.add and min")
        self._capacity = capacity
        self._operation = operation
        self._kind = kind
        self._dev = _device_index(device)
        self._device = torch.device("cuda", self._dev)
        h = _n.C.c_void_p()
        _n.check(_n.lib.seg_create(_n.C.byref(h), self._dev, int(capacity), kind))
        self._h = h

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and _n is not None:
            _n.lib.seg_destroy(h)
            self._h = None

    # -- batched forms (extensions) ---------------------------------------------------------
    def set_batch(self, idxes, values):
        """for i, v in zip(idxes, values): self[i] = v   — in that order (memory.py:76-87)."""
        idx = torch.as_tensor(np.asarray(idxes, dtype=np.int64)).to(self._device)
        val = torch.as_tensor(np.asarray(values, dtype=np.float64)).to(self._device)
        _n.check(_n.lib.seg_set(self._h, _n.ptr(idx), _n.ptr(val), idx.numel(), _n.stream_ptr()))
        self._raise_status()

    def _raise_status(self):
        w = _n.C.c_int64(0)
        _n.check(_n.lib.seg_status(self._h, _n.C.byref(w), _n.stream_ptr()))
        assert not (w.value & _n.STATUS_BAD_INDEX), "index out of range"
        assert not (w.value & _n.STATUS_BAD_PREFIXSUM), "prefixsum outside [0, sum + 1e-5]"

    @property
    def _value(self):
        """The node array in the reference's heap layout (memory.py:36), copied to the host."""
        out = torch.empty(2 * self._capacity, dtype=torch.float64, device=self._device)
        _n.check(_n.lib.seg_read(self._h, _n.ptr(out), _n.stream_ptr()))
        return out.cpu().numpy()

    # -- reference API ----------------------------------------------------------------------
    def reduce(self, start=0, end=None):
        """memory.py:54-74."""
        if end is None:
            end = self._capacity
        out = torch.empty(1, dtype=torch.float64, device=self._device)
        _n.check(_n.lib.seg_reduce(self._h, int(start), int(end), _n.ptr(out), _n.stream_ptr()))
        return float(out.item())

    def __setitem__(self, idx, val):
        self.set_batch([int(idx)], [float(val)])

    def __getitem__(self, idx):
        """memory.py:89-91: one leaf (8 bytes cross the bus, not the node array)."""
        assert 0 <= idx < self._capacity
        v = _n.C.c_double(0.0)
        _n.check(_n.lib.seg_get(self._h, int(idx), _n.C.byref(v), _n.stream_ptr()))
        return v.value


class SumSegmentTree(SegmentTree):
    """memory.py:94-129."""

    def __init__(self, capacity, device=None):
        super(SumSegmentTree, self).__init__(capacity=capacity, operation=operator.add, neutral_element=0.0,
                                             device=device)

    def sum(self, start=0, end=None):
        return super(SumSegmentTree, self).reduce(start, end)

    def find_prefixsum_idx_batch(self, prefixsums):
        m = torch.as_tensor(np.asarray(prefixsums, dtype=np.float64)).to(self._device)
        out = torch.empty(m.numel(), dtype=torch.int64, device=self._device)
        _n.check(_n.lib.seg_find_prefixsum(self._h, _n.ptr(m), m.numel(), _n.ptr(out), _n.stream_ptr()))
        self._raise_status()
        return out.cpu().numpy()

    def find_prefixsum_idx(self, prefixsum):
        return int(self.find_prefixsum_idx_batch([float(prefixsum)])[0])


class MinSegmentTree(SegmentTree):
    """memory.py:132-143."""

    def __init__(self, capacity, device=None):
        super(MinSegmentTree, self).__init__(capacity=capacity, operation=min, neutral_element=float('inf'),
                                             device=device)

    def min(self, start=0, end=None):
        return super(MinSegmentTree, self).reduce(start, end)


class _TreeView(object):
    """`buffer._it_sum` / `buffer._it_min`: read access to a replay buffer's device tree with
    the reference's accessor names (memory.py:231-232 make them SegmentTree objects)."""

    def __init__(self, owner, which):
        self._owner = owner
        self._which = which

    @property
    def _capacity(self):
        return self._owner._tree_capacity

    @property
    def _value(self):
        return self._owner.tree_values(self._which)

    def __getitem__(self, idx):
        """`_it_sum[idx]` (memory.py:293): one leaf."""
        assert 0 <= idx < self._capacity
        o = self._owner
        o._flush()
        if o._h is None:
            return 0.0 if self._which == _n.TREE_SUM else float("inf")
        v = _n.C.c_double(0.0)
        _n.check(_n.lib.per_tree_get(o._h, self._which, int(idx), _n.C.byref(v), _n.stream_ptr()))
        return v.value

    def reduce(self, start=0, end=None):
        """`_it_sum.sum(start, end)` / `_it_min.min(start, end)` (memory.py:54-74) on the buffer's own tree."""
        o = self._owner
        o._flush()
        if end is None:
            end = self._capacity
        if o._h is None:
            return 0.0 if self._which == _n.TREE_SUM else float("inf")
        out = torch.empty(1, dtype=torch.float64, device=o._device)
        _n.check(_n.lib.per_tree_reduce(o._h, self._which, int(start), int(end), _n.ptr(out), _n.stream_ptr()))
        return float(out.item())

    sum = reduce
    min = reduce


class _StorageView(object):
    """Stands in for the reference's `_storage` list where callers only take its len()."""

    def __init__(self, owner):
        self._owner = owner

    def __len__(self):
        return len(self._owner)


# ==========================================================================================
# replay buffers (memory.py:146-391)
# ==========================================================================================
class ReplayBuffer(object):
    """memory.py:146-205.  Rows live on the GPU; the uniform sampler draws its indices from
    Python's global `random` exactly like memory.py:204."""

    _alpha_default = 0.0
    _row_fields = 5  # obs, action, reward, next_obs, done

    def __init__(self, size, device=None):
        self._maxsize = size
        self._ring = int(size)  # the reference leaves a float size un-cast and breaks at the first wrap (memory.py:342)
        self._dev = _device_index(device)
        self._device = torch.device("cuda", self._dev)
        self._h = None
        self._len = 0
        self._cursor = 0
        self._staged = []          # rows not yet uploaded
        self._staged_prio = []     # matching priorities (apex) or None
        self._obs_shape = None
        self._obs_dtype = None
        self._extra_shape = None
        self._extra_dtype = None
        self._storage = _StorageView(self)
        self._host = _HostStage(self._device)
        self._max_priority_stale = False
        if not hasattr(self, "_alpha"):
            self._alpha = self._alpha_default
        self._max_priority = 1.0

    # -- native buffer ------------------------------------------------------------------------
    @property
    def _tree_capacity(self):
        c = 1
        while c < self._ring:
            c *= 2
        return c

    def _ensure_native(self, obs, extra=None):
        if self._h is not None:
            return
        o = np.asarray(obs)
        self._obs_shape, self._obs_dtype = o.shape, o.dtype
        extra_bytes = 0
        if extra is not None:
            e = np.asarray(extra)
            self._extra_shape, self._extra_dtype = e.shape, e.dtype
            extra_bytes = e.nbytes
        h = _n.C.c_void_p()
        _n.check(_n.lib.per_create(_n.C.byref(h), self._dev, self._ring, float(self._alpha), int(o.nbytes),
                                   int(extra_bytes)))
        self._h = h

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and _n is not None:
            _n.lib.per_destroy(h)
            self._h = None

    def __len__(self):
        return self._len

    @property
    def _next_idx(self):
        return self._cursor

    # -- insert -------------------------------------------------------------------------------
    def _stage(self, row, prio):
        """Records one add (memory.py:162-169); the upload happens in batches."""
        self._ensure_native(row[0], row[5] if len(row) > 5 else None)
        self._staged.append(row)
        self._staged_prio.append(prio)
        self._cursor = (self._cursor + 1) % self._ring
        self._len = min(self._ring, self._len + 1)
        if len(self._staged) >= min(_FLUSH_ROWS, self._ring):
            self._flush()

    def add(self, obs_t, action, reward, obs_tp1, done):
        self._stage((obs_t, action, reward, obs_tp1, done), None)

    def _leaf_for_staged(self, prios):
        """Leaf values for the staged rows.  Base ring: max_priority ** alpha with alpha 0 -> 1.0."""
        n = len(prios)
        return np.full(n, float(self._max_priority) ** float(self._alpha), dtype=np.float64), 0.0

    def _flush(self):
        if not self._staged:
            return
        rows, prios = self._staged, self._staged_prio
        self._staged, self._staged_prio = [], []
        n = len(rows)
        obs = np.stack([np.asarray(r[0], dtype=self._obs_dtype).reshape(self._obs_shape) for r in rows])
        nxt = np.stack([np.asarray(r[3], dtype=self._obs_dtype).reshape(self._obs_shape) for r in rows])
        act = np.asarray([r[1] for r in rows], dtype=np.int64)
        rew = np.asarray([r[2] for r in rows], dtype=np.float32)
        done = np.asarray([r[4] for r in rows], dtype=np.float32)
        extra = None
        if self._extra_shape is not None:
            extra = np.stack([np.asarray(r[5], dtype=self._extra_dtype).reshape(self._extra_shape) for r in rows])
        leaves, batch_max = self._leaf_for_staged(prios)
        up = self._host.up
        self.add_batch_device(up("obs", obs), up("next", nxt), up("act", act), up("rew", rew), up("done", done),
                              None if extra is None else up("extra", extra), leaves=up("leaf", leaves),
                              batch_max=batch_max, _count=False)
        assert n <= self._ring

    def add_batch_device(self, obs, next_obs, actions, rewards, dones, extra=None, priorities=None, leaves=None,
                         batch_max=0.0, _count=True):
        """Extension: n adds in one call from CUDA tensors (uint8/any [n, ...] rows, int64 actions,
        float32 rewards / dones, optional float64 priorities or leaves).  Same slot order as n
        sequential reference adds (memory.py:162-169, :334-346)."""
        n = int(actions.numel())
        if self._h is None:
            self._obs_shape, self._obs_dtype = tuple(obs.shape[1:]), np.dtype(str(obs.dtype).replace("torch.", ""))
            eb = 0
            if extra is not None:
                self._extra_shape = tuple(extra.shape[1:])
                self._extra_dtype = np.dtype(str(extra.dtype).replace("torch.", ""))
                eb = extra[0].numel() * extra.element_size()
            h = _n.C.c_void_p()
            _n.check(_n.lib.per_create(_n.C.byref(h), self._dev, self._ring, float(self._alpha),
                                       obs[0].numel() * obs.element_size(), eb))
            self._h = h
        if _count:
            self._flush()
        cut = lambda t, lo, hi: None if t is None else t[lo:hi].contiguous()
        for lo in range(0, n, self._ring):  # one native call never laps the ring
            hi = min(n, lo + self._ring)
            parts = [cut(t, lo, hi) for t in (obs, next_obs, actions, rewards, dones, extra, priorities, leaves)]
            _n.check(_n.lib.per_add_batch(self._h, *[_n.ptr(t) for t in parts], float(batch_max), hi - lo,
                                          _n.stream_ptr()))
        if _count:
            self._cursor = (self._cursor + n) % self._ring
            self._len = min(self._ring, self._len + n)
            self._max_priority_stale = self._max_priority_stale or priorities is not None
        assert _n.lib.per_len(self._h) == self._len and _n.lib.per_cursor(self._h) == self._cursor

    # -- gather -------------------------------------------------------------------------------
    def _gather_device(self, idx_t):
        b = idx_t.numel()
        dt = getattr(torch, self._obs_dtype.name)
        obs = torch.empty((b,) + tuple(self._obs_shape), dtype=dt, device=self._device)
        nxt = torch.empty_like(obs)
        act = torch.empty(b, dtype=torch.int64, device=self._device)
        rew = torch.empty(b, dtype=torch.float32, device=self._device)
        done = torch.empty(b, dtype=torch.float32, device=self._device)
        extra = None
        if self._extra_shape is not None:
            extra = torch.empty((b,) + tuple(self._extra_shape), dtype=getattr(torch, self._extra_dtype.name),
                                device=self._device)
        _n.check(_n.lib.per_gather(self._h, _n.ptr(idx_t), b, _n.ptr(obs), _n.ptr(nxt), _n.ptr(act), _n.ptr(rew),
                                   _n.ptr(done), _n.ptr(extra), _n.stream_ptr()))
        return obs, act, rew, nxt, done, extra

    def _gather_cast_device(self, idx_t, dtype, scale, channels_last=False):
        """Extension: the gather fused with the learner's uint8 -> float widening (ApeX.py:50,53); channels_last hands
        the convolution NHWC memory (torch.channels_last tensors of the usual [B, C, H, W] shape)."""
        code = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}[dtype]
        b = idx_t.numel()
        act = torch.empty(b, dtype=torch.int64, device=self._device)
        rew = torch.empty(b, dtype=torch.float32, device=self._device)
        done = torch.empty(b, dtype=torch.float32, device=self._device)
        if channels_last:
            assert len(self._obs_shape) == 3 and self._obs_shape[0] == 4, "channels_last needs [4, H, W] observations"
            obs = torch.empty((b,) + tuple(self._obs_shape), dtype=dtype, device=self._device, memory_format=torch.channels_last)
            nxt = torch.empty_like(obs)  # preserves the format
            _n.check(_n.lib.per_gather_cast_nhwc(self._h, _n.C.c_void_p(idx_t.data_ptr()), b, code, float(scale), 4,
                                                 _n.C.c_void_p(obs.data_ptr()), _n.C.c_void_p(nxt.data_ptr()), _n.ptr(act),
                                                 _n.ptr(rew), _n.ptr(done), _n.stream_ptr()))
            return obs, act, rew, nxt, done, None
        obs = torch.empty((b,) + tuple(self._obs_shape), dtype=dtype, device=self._device)
        nxt = torch.empty_like(obs)
        _n.check(_n.lib.per_gather_cast(self._h, _n.ptr(idx_t), b, code, float(scale), _n.ptr(obs), _n.ptr(nxt), _n.ptr(act),
                                        _n.ptr(rew), _n.ptr(done), _n.stream_ptr()))
        return obs, act, rew, nxt, done, None

    def _encode_sample(self, idxes):
        """memory.py:171-182: (obs, actions, rewards, next_obs, dones) as host arrays."""
        self._flush()
        idx_t = self._host.up("gather_idx", np.asarray(idxes, dtype=np.int64))
        obs, act, rew, nxt, done, extra = self._gather_device(idx_t)
        named = dict(obs=obs, act=act, rew=rew, next_obs=nxt, done=done)
        if extra is not None:
            named["a_mu"] = extra
        h = self._host.down(named)
        out = (h["obs"], h["act"], h["rew"], h["next_obs"], h["done"])
        return out + ((h["a_mu"],) if extra is not None else ())

    def sample(self, batch_size):
        """memory.py:184-205."""
        idxes = [random.randint(0, len(self) - 1) for _ in range(batch_size)]
        return self._encode_sample(idxes)


class PrioritizedReplayBuffer(ReplayBuffer):
    """memory.py:208-320."""

    def __init__(self, size, alpha, device=None):
        assert alpha >= 0
        self._alpha = alpha
        super(PrioritizedReplayBuffer, self).__init__(size, device)
        self._max_priority = 1.0
        self._max_priority_stale = False
        self._it_sum = _TreeView(self, _n.TREE_SUM)
        self._it_min = _TreeView(self, _n.TREE_MIN)

    # `_max_priority` (memory.py:233,320) is mirrored on the host; device-side updates mark it stale
    def _sync_max_priority(self):
        if self._max_priority_stale and self._h is not None:
            v = _n.C.c_double(0.0)
            _n.check(_n.lib.per_max_priority(self._h, _n.C.byref(v), _n.stream_ptr()))
            self._max_priority = max(self._max_priority, v.value)
            self._max_priority_stale = False

    def tree_values(self, which):
        """`_it_sum._value` / `_it_min._value` (memory.py:36) as a host float64 array."""
        self._flush()
        cap = self._tree_capacity
        if self._h is None:
            return np.full(2 * cap, 0.0 if which == _n.TREE_SUM else np.inf)
        out = torch.empty(2 * cap, dtype=torch.float64, device=self._device)
        _n.check(_n.lib.per_tree_read(self._h, which, _n.ptr(out), _n.stream_ptr()))
        return out.cpu().numpy()

    def _raise_status(self):
        w = _n.C.c_int64(0)
        _n.check(_n.lib.per_status(self._h, _n.C.byref(w), _n.stream_ptr()))
        assert not (w.value & _n.STATUS_BAD_PRIORITY), "priority > 0"
        assert not (w.value & _n.STATUS_BAD_INDEX), "0 <= idx < len(self._storage)"

    def add(self, *args, **kwargs):
        """memory.py:235-240: new rows enter with the running max priority."""
        self._sync_max_priority()
        row = tuple(args) + tuple(kwargs[k] for k in ("obs_t", "action", "reward", "obs_tp1", "done") if k in kwargs)
        self._stage(row, None)

    def _leaf_for_staged(self, prios):
        if all(p is None for p in prios):
            leaf = float(self._max_priority) ** float(self._alpha)
            return np.full(len(prios), leaf, dtype=np.float64), 0.0
        p = np.asarray(prios, dtype=np.float64)
        return _pow_alpha_exact(p, self._alpha), float(p.max())

    def _stage(self, row, prio):
        # a staged row's leaf depends on _max_priority at the time of the add; an update in between flushes first
        super(PrioritizedReplayBuffer, self)._stage(row, prio)

    def _draw_uniforms(self, batch_size):
        return draw_uniforms(batch_size)

    def sample_device(self, batch_size, beta, uniforms=None, gather=True, obs_dtype=None, obs_scale=1.0, channels_last=False):
        """Extension: sample() with every output left on the GPU.  Returns a dict with
        idx (int64), weights (float32), weights64, and — if gather — obs, act, rew, next_obs,
        done (and a_mu).  `uniforms`: optional float64 CUDA / host vector replacing the draws.
        obs_dtype = torch.float32 / bfloat16 / float16 widens the uint8 frames inside the gather
        (value * obs_scale), replacing the learner's torch.FloatTensor(states) (ApeX.py:50,53); channels_last=True
        returns them as torch.channels_last tensors (NHWC memory) for the convolution trunk."""
        assert beta > 0
        self._flush()
        if uniforms is None:
            uniforms = self._draw_uniforms(batch_size)
        if torch.is_tensor(uniforms):
            u = uniforms.to(self._device, dtype=torch.float64, non_blocking=True)
        else:
            u = self._host.up("u", np.asarray(uniforms, dtype=np.float64))
        idx = torch.empty(batch_size, dtype=torch.int64, device=self._device)
        w64 = torch.empty(batch_size, dtype=torch.float64, device=self._device)
        w32 = torch.empty(batch_size, dtype=torch.float32, device=self._device)
        _n.check(_n.lib.per_sample_indices(self._h, _n.ptr(u), batch_size, float(beta), _n.ptr(idx), _n.ptr(w64),
                                           _n.ptr(w32), _n.stream_ptr()))
        self._bound_idx = idx  # the native side binds this address to the sample until its update (per_b200.h): keep it allocated
        out = {"idx": idx, "weights": w32, "weights64": w64}
        if gather:
            if obs_dtype is not None and self._obs_dtype == np.uint8:
                obs, act, rew, nxt, done, extra = self._gather_cast_device(idx, obs_dtype, obs_scale, channels_last)
            else:
                obs, act, rew, nxt, done, extra = self._gather_device(idx)
            out.update(obs=obs, act=act, rew=rew, next_obs=nxt, done=done)
            if extra is not None:
                out["a_mu"] = extra
        return out

    def _sample_proportional(self, batch_size):
        """memory.py:242-250."""
        return self.sample_device(batch_size, 1.0, gather=False)["idx"].cpu().tolist()  # weights unused

    _sample_fields = ("obs", "act", "rew", "next_obs", "done")

    def sample(self, batch_size, beta):
        """memory.py:252-298 -> (obs, act, rew, next_obs, done, weights float64[B], idxes list[int]).
        One native call with host buffers (per_sample_host): uniforms up, sampling + gather kernels,
        results into pinned host arrays (handed out as numpy views, see _HostStage)."""
        assert beta > 0
        assert self._h is not None and len(self) >= 2, "sample() needs at least two stored transitions"
        self._flush()
        B = int(batch_size)
        u = self._draw_uniforms(B)
        arrs, ptrs = self._sample_slot(B)
        _n.check(_n.lib.per_sample_host(self._h, u.ctypes.data, B, float(beta), *ptrs, _n.stream_ptr()))
        out = {k: np.array(v) for k, v in arrs.items()} if COPY_SAMPLES else arrs
        return tuple([out[k] for k in self._sample_fields] + [out["weights64"], out["idx"].tolist()])

    def _sample_slot(self, B):
        """Host arrays for one sample of B rows and their addresses in per_sample_host's argument order: a ring of
        _HostStage.RING sets per batch size, built once (the frame arrays pinned: the library copies straight into them;
        the per-row scalars plain: it hands them out of its own pinned block)."""
        plans = self.__dict__.setdefault("_sample_plans", {})
        plan = plans.get(B)
        if plan is None:
            sets = []
            for _ in range(_HostStage.RING):
                odt = getattr(torch, self._obs_dtype.name)
                keep = [torch.empty((B,) + tuple(self._obs_shape), dtype=odt, pin_memory=True) for _ in range(2)]
                arrs = {"obs": keep[0].numpy(), "next_obs": keep[1].numpy(), "act": np.empty(B, np.int64),
                        "rew": np.empty(B, np.float32), "done": np.empty(B, np.float32),
                        "weights64": np.empty(B, np.float64), "idx": np.empty(B, np.int64)}
                if self._extra_shape is not None:
                    keep.append(torch.empty((B,) + tuple(self._extra_shape), dtype=getattr(torch, self._extra_dtype.name), pin_memory=True))
                    arrs["a_mu"] = keep[-1].numpy()
                P = lambda a: _n.C.c_void_p(a.ctypes.data)
                ptrs = (P(arrs["obs"]), P(arrs["next_obs"]), P(arrs["act"]), P(arrs["rew"]), P(arrs["done"]),
                        P(arrs["a_mu"]) if "a_mu" in arrs else None, P(arrs["weights64"]), P(arrs["idx"]))
                sets.append((arrs, ptrs, keep))
            plan = plans[B] = [sets, 0]
        arrs, ptrs, _ = plan[0][plan[1] % _HostStage.RING]
        plan[1] += 1
        return arrs, ptrs

    def update_priorities(self, idxes, priorities):
        """memory.py:300-320.  The reference walks the pairs in order and asserts as it goes;
        here the same conditions are checked for the whole batch first (in per_update_host), then
        the batch is applied on the GPU with the same last-write-wins result."""
        assert len(idxes) == len(priorities)
        if len(idxes) == 0:
            return
        self._flush()
        idx = np.ascontiguousarray(idxes, dtype=np.int64)
        pr = np.ascontiguousarray(priorities, dtype=np.float64)
        bmax = _n.C.c_double(0.0)
        _n.check(_n.lib.per_update_host(self._h, idx.ctypes.data, pr.ctypes.data, idx.size, _n.C.byref(bmax), _n.stream_ptr()))
        self._max_priority = max(self._max_priority, bmax.value)

    def update_priorities_device(self, idx, priorities):
        """Extension: idx int64 and priorities float32/float64 CUDA tensors; pow on the device
        (<= 2 ulp from libm's), invalid entries are skipped and reported by the next status check."""
        self._flush()
        assert idx.numel() == priorities.numel()
        kind = _n.PRIO_F64 if priorities.dtype == torch.float64 else _n.PRIO_F32
        if kind == _n.PRIO_F32:
            assert priorities.dtype == torch.float32
        _n.check(_n.lib.per_update_priorities(self._h, _n.ptr(idx), _n.ptr(priorities), kind, idx.numel(),
                                              _n.stream_ptr()))
        self._max_priority_stale = True

    def td_priority_update(self, q, q_next, q_tgt, actions, rewards, dones, weights, n_steps, gamma, idx=None,
                           want_loss=True):
        """Extension: utils.compute_loss's arithmetic after the forwards (utils.py:71-80) fused
        with update_priorities (ApeX.py:58-63) in one launch.  All arguments are CUDA tensors
        (float32 [B, nA] Q tables, int64 actions, float32 rewards / dones / weights).  Returns
        (loss[1] or None, prios float32[B], td float32[B]) on the device."""
        self._flush()
        b = int(actions.numel())
        prio = torch.empty(b, dtype=torch.float32, device=self._device)
        td = torch.empty(b, dtype=torch.float32, device=self._device)
        loss = torch.empty(1, dtype=torch.float32, device=self._device) if want_loss else None
        gamma_n = float(np.float32(gamma ** n_steps))
        _n.check(_n.lib.per_td_priority_update(self._h, _n.ptr(q), _n.ptr(q_next), _n.ptr(q_tgt), int(q.shape[1]),
                                               _n.ptr(actions), _n.ptr(rewards), _n.ptr(dones), _n.ptr(weights),
                                               gamma_n, _n.ptr(idx), b, _n.ptr(prio), _n.ptr(td), _n.ptr(loss),
                                               _n.stream_ptr()))
        if idx is not None:
            self._max_priority_stale = True
        return loss, prio, td


class CustomPrioritizedReplayBuffer(PrioritizedReplayBuffer):
    """memory.py:323-362: `add` takes the actor-computed priority."""

    def __init__(self, size, alpha, device=None):
        super(CustomPrioritizedReplayBuffer, self).__init__(size, alpha, device)

    def add(self, state, action, reward, next_state, done, priority):
        """memory.py:334-346."""
        priority = float(priority)
        self._sync_max_priority()
        if self._staged_prio and self._staged_prio[-1] is None:
            self._flush()
        self._stage((state, action, reward, next_state, done), priority)
        self._max_priority = max(self._max_priority, priority)

    def prepare_chunk(self, states, actions, rewards, next_states, dones, priorities):
        """Host half of add_chunk: stacks the rows and raises the priorities to alpha (libm pow).  Touches no buffer
        state, so any number of threads may prepare chunks at the same time (replay_ops.ReplayServer does)."""
        self._ensure_native(states[0])
        pr = np.asarray(priorities, dtype=np.float64)
        return {"n": len(actions),
                "obs": np.stack([np.asarray(s, dtype=self._obs_dtype).reshape(self._obs_shape) for s in states]),
                "next": np.stack([np.asarray(s, dtype=self._obs_dtype).reshape(self._obs_shape) for s in next_states]),
                "act": np.asarray(actions, dtype=np.int64), "rew": np.asarray(rewards, dtype=np.float32),
                "done": np.asarray(dones, dtype=np.float32), "leaf": _pow_alpha_exact(pr, self._alpha), "max": float(pr.max())}

    def commit_chunk(self, c):
        """Device half of add_chunk: one upload and one launch; one caller at a time (like every buffer call)."""
        self._sync_max_priority()
        self._flush()
        up = self._host.up
        self.add_batch_device(up("obs", c["obs"]), up("next", c["next"]), up("act", c["act"]), up("rew", c["rew"]),
                              up("done", c["done"]), leaves=up("leaf", c["leaf"]), batch_max=c["max"])
        self._max_priority = max(self._max_priority, c["max"])

    def add_chunk(self, states, actions, rewards, next_states, dones, priorities):
        """Extension: `for i: self.add(states[i], actions[i], rewards[i], next_states[i], dones[i],
        priorities[i])` for one actor chunk (batchrecorder.py:133-137) as a single upload and
        launch; same slots, same leaves (libm pow on the host), same running max."""
        if len(actions) == 0:
            return
        self.commit_chunk(self.prepare_chunk(states, actions, rewards, next_states, dones, priorities))


class CustomPrioritizedReplayBuffer_AQL(PrioritizedReplayBuffer):
    """memory.py:364-391: rows carry the K candidate actions `a_mu`; sample returns an 8-tuple."""

    def __init__(self, size, alpha, device=None):
        super(CustomPrioritizedReplayBuffer_AQL, self).__init__(size, alpha, device)

    def add(self, obs_t, action, reward, obs_tp1, done, a_mu):
        """memory.py:368-378."""
        self._sync_max_priority()
        self._stage((obs_t, action, reward, obs_tp1, done, a_mu), None)

    _sample_fields = ("obs", "act", "rew", "next_obs", "done", "a_mu")


# ==========================================================================================
# actor-side n-step assembly (memory.py:393-478)
# ==========================================================================================
class BatchStorage:
    """memory.py:393-478, host side: an actor feeds it one environment step at a time, on
    CPU, next to its environment (batchrecorder.py:54, origin_repo/actor.py:82), so the
    per-step window logic stays in Python; the arithmetic and quirks are the reference's
    (n+1-term return, single emit on done, q-window kept across episodes, plain max-Q)."""

    def __init__(self, n_steps, gamma=0.99):
        self.n_steps = n_steps
        self.gamma = gamma
        self.state_deque = deque(maxlen=n_steps)
        self.action_deque = deque(maxlen=n_steps)
        self.reward_deque = deque(maxlen=n_steps)
        self.q_values_deque = deque(maxlen=n_steps)
        self.reset()

    def reset(self):
        self.states, self.actions, self.rewards = [], [], []
        self.next_states, self.dones = [], []
        self.q_values, self.next_q_values = [], []

    def multi_step_reward(self, *rewards):
        ret = 0.
        for k, r in enumerate(rewards):
            ret += r * (self.gamma ** k)
        return ret

    def add(self, state, reward, action, done, q_values):
        if len(self.state_deque) == self.n_steps or done:
            self.states.append(self.state_deque[0])
            self.rewards.append(self.multi_step_reward(*self.reward_deque, reward))
            self.actions.append(self.action_deque[0])
            self.q_values.append(self.q_values_deque[0])
            self.next_states.append(state)
            self.next_q_values.append(q_values)
            done = np.float32(done)
            self.dones.append(done)
        if done:
            self.state_deque.clear()
            self.reward_deque.clear()
            self.action_deque.clear()
        else:
            self.state_deque.append(state)
            self.reward_deque.append(reward)
            self.action_deque.append(action)
            self.q_values_deque.append(q_values)

    def compute_priorities(self):
        actions = np.asarray(self.actions)
        rewards = np.asarray(self.rewards)
        dones = np.asarray(self.dones)
        q0 = np.stack(self.q_values)
        qn = np.stack(self.next_q_values)
        q_a = q0[(range(len(q0)), actions)]
        target = rewards + (self.gamma ** self.n_steps) * qn.max(1) * (1 - dones)
        return np.abs(target - q_a) + 1e-6

    def make_batch(self):
        prios = self.compute_priorities()
        return [self.states, self.actions, self.rewards, self.next_states, self.dones], prios

    def __len__(self):
        return len(self.states)
