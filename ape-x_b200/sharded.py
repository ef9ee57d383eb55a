"""Prioritized replay sharded over the GPUs of one box (BASELINE configs[3]).

The reference keeps one tree in one process (memory.py:231-232).  Here the replay of `size`
transitions is cut into `world` (a power of two) contiguous index ranges, one per GPU / rank:
with C = next_pow2(size) leaves in the reference's tree, rank r owns leaves
[r*C/world, (r+1)*C/world) — the subtree under the reference's node world + r — together with
the ring slots that fall into that range.  The reference tree is then exactly the `world` local
trees plus the world-1 nodes above them, and one sample() is

  1. every rank writes a 32-byte record {root sum, root min, its part of the reference's
     sum(0, len-1) chain, local len}                         (per_shard_record)
  2. ONE collective: all-gather of the records, rank order   (NCCL over NVLink; gloo on CPU)
  3. every rank runs all strata through the top log2(world) levels redundantly, keeps the
     ones that land in its slice and finishes them on its own tree        (per_sample_sharded)
  4. rank-local gather of the sampled rows; rows, indices and weights can be written straight
     into a peer GPU's batch buffers by passing peer-mapped output pointers (the gather kernel
     only sees addresses), and returned priorities are read the same way.

Every rank passes the same full-batch arrays to gather / update: indices are global, and
-1 marks a stratum owned by another rank (skipped inside the kernels).

Host logic here: geometry, routing of a global add stream to shards, cursor / len bookkeeping,
the collective.  Everything else is the C ABI (include/per_b200.h).
"""
import numpy as np
import torch

import per_native as _n

REC = 4  # doubles per shard record


class ShardGeometry(object):
    """Where the reference's global slots live.  `size` = the reference's buffer size (ring
    wraps there, memory.py:169); tree capacity = next power of two (memory.py:227-229)."""

    def __init__(self, size, world):
        size = int(size)
        assert world >= 1 and world & (world - 1) == 0 and world <= 16, "world must be a power of two <= 16"
        cap = 1
        while cap < size:
            cap *= 2
        assert cap >= world
        self.size, self.world, self.cap = size, world, cap
        self.cap_local = cap // world

    def ring_local(self, rank):
        return max(0, min(self.cap_local, self.size - rank * self.cap_local))

    def owner(self, slot):
        return slot // self.cap_local

    def route(self, cursor, n):
        """Splits n consecutive adds starting at global slot `cursor` (wrapping at size) into
        [(rank, src_lo, src_hi, local_slot)] segments in add order (memory.py:162-169)."""
        out, done = [], 0
        while done < n:
            slot = (cursor + done) % self.size
            r = slot // self.cap_local
            end = min((r + 1) * self.cap_local, self.size)
            m = min(n - done, end - slot)
            out.append((r, done, done + m, slot - r * self.cap_local))
            done += m
        return out


class PeerBuffer(object):
    """A raw device allocation that other processes' GPUs can map (CUDA IPC): passes for a tensor
    where the C ABI only needs an address, and exposes __cuda_array_interface__ views."""

    def __init__(self, ptr, nbytes, device, owned):
        self.ptr, self.nbytes, self.device, self.owned = int(ptr), int(nbytes), int(device), owned

    def data_ptr(self):
        return self.ptr

    def is_contiguous(self):
        return True

    def view(self, dtype, shape, offset=0):
        """torch tensor view (no copy) of `shape` / torch `dtype` at byte `offset`."""
        typestr = {torch.uint8: "|u1", torch.int64: "<i8", torch.float32: "<f4", torch.float64: "<f8"}[dtype]
        holder = type("_CAI", (), {})()
        holder.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (self.ptr + offset, False),
                                           "version": 2}
        holder._keep = self
        return torch.as_tensor(holder, device=torch.device("cuda", self.device))

    @staticmethod
    def alloc(nbytes, device):
        p, h = _n.C.c_void_p(), (_n.C.c_uint8 * 64)()
        _n.check(_n.lib.per_ipc_alloc(int(device), int(nbytes), _n.C.byref(p), h))
        return PeerBuffer(p.value, nbytes, device, True), bytes(h)

    @staticmethod
    def open(handle, nbytes, device):
        p = _n.C.c_void_p()
        h = (_n.C.c_uint8 * 64).from_buffer_copy(handle)
        _n.check(_n.lib.per_ipc_open(int(device), h, _n.C.byref(p)))
        return PeerBuffer(p.value, nbytes, device, False)


class LearnerBatch(object):
    """The learner GPU's batch buffers as every rank sees them: allocated once on `learner_rank` (CUDA IPC),
    mapped into all other ranks, so that each shard's gather kernel stores the rows it owns straight into the
    learner's memory over NVLink and the learner's priorities are read back the same way (SURVEY.md section 8e;
    the reference moves both as pickles over ZeroMQ, origin_repo/replay.py:29-45, origin_repo/learner.py:71-131).
    Fields (full global batch): obs, next_obs (obs_dtype [B, ...]), act int64, rew / done / weights / prio float32,
    idx int64.  `local` is True on the learner rank, where the fields are ordinary views of its own memory."""
    FIELDS = ("obs", "next_obs", "act", "rew", "done", "idx", "weights", "prio")

    def __init__(self, bufs, views, batch, learner_rank, local, group):
        self.bufs, self.views, self.batch = bufs, views, batch
        self.learner_rank, self.local, self.group = learner_rank, local, group

    def __getitem__(self, k):
        return self.views[k]

    def fence(self):
        """Everything every rank has issued so far has landed (stream sync + barrier over the ranks)."""
        import torch.distributed as dist
        torch.cuda.current_stream().synchronize()
        dist.barrier(group=self.group)


def exchange_records(rec, group=None):
    """All-gather of one 4-double record per rank (the only collective of a sample)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    out = torch.empty(world * REC, dtype=rec.dtype, device=rec.device)
    dist.all_gather_into_tensor(out, rec, group=group)
    return out


class ShardedReplay(object):
    """One rank's slice of a sharded CustomPrioritizedReplayBuffer (memory.py:323-362).
    `exchange`: callable(record tensor) -> all records in rank order; defaults to the
    torch.distributed all-gather on `group`.  Tests emulate several ranks in one process by
    passing their own exchange."""

    def __init__(self, size, alpha, rank, world, obs_shape=(4, 84, 84), obs_dtype=torch.uint8, extra_shape=None,
                 extra_dtype=torch.float32, device=None, group=None, exchange=None):
        assert alpha >= 0
        self.geo = ShardGeometry(size, world)
        self.rank, self.world, self.alpha = rank, world, alpha
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.obs_shape, self.obs_dtype = tuple(obs_shape), obs_dtype
        self.extra_shape, self.extra_dtype = extra_shape, extra_dtype
        self._exchange = exchange if exchange is not None else (lambda rec: exchange_records(rec, group))
        obs_bytes = int(np.prod(obs_shape)) * torch.empty(0, dtype=obs_dtype).element_size()
        extra_bytes = 0 if extra_shape is None else int(np.prod(extra_shape)) * torch.empty(0, dtype=extra_dtype).element_size()
        h = _n.C.c_void_p()
        _n.check(_n.lib.per_create_shard(_n.C.byref(h), self.device.index, self.geo.size, rank, world, float(alpha),
                                         obs_bytes, extra_bytes))
        self._h = h
        self._len = 0       # global len(_storage)
        self._cursor = 0    # global _next_idx
        self._rec = torch.zeros(REC, dtype=torch.float64, device=self.device)
        self._rec_fresh = False  # True while the record at _rec reflects the current trees and len
        self._peer_exchange = False
        self._attach()

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and _n is not None:
            _n.lib.per_destroy(h)
            self._h = None

    def __len__(self):
        return self._len

    # -- record exchange through peer memory (no collective in the sampling path) -------------------
    MAILBOX_BYTES = 2 * 16 * 64

    def enable_peer_exchange(self, group=None):
        """One-time set-up across processes: every rank allocates a mailbox its peers can write
        (CUDA IPC over NVLink), the 64-byte handles are all-gathered once, and from then on
        sample_device() needs no collective: the sampling kernel publishes and awaits the records
        itself (include/per_b200.h, per_shard_attach_exchange)."""
        import torch.distributed as dist
        mine, handle = PeerBuffer.alloc(self.MAILBOX_BYTES, self.device.index)
        handles = [None] * self.world
        dist.all_gather_object(handles, handle, group=group)
        peers = [mine if r == self.rank else PeerBuffer.open(handles[r], self.MAILBOX_BYTES, self.device.index)
                 for r in range(self.world)]
        self._attach_exchange(mine, peers)
        dist.barrier(group=group)

    def _attach_exchange(self, mine, peers):
        arr = (_n.C.c_void_p * self.world)(*[p.data_ptr() for p in peers])
        _n.check(_n.lib.per_shard_attach_exchange(self._h, _n.C.c_void_p(mine.data_ptr()), arr))
        self._mailbox, self._peer_mailboxes, self._peer_exchange = mine, peers, True

    @staticmethod
    def enable_local_exchange(shards):
        """All ranks in ONE process on one GPU (tests): mailboxes are ordinary local buffers.  Since
        one rank's kernel cannot wait for another's on the same GPU, the caller must run
        publish_record() on every shard before the shards sample."""
        boxes = [PeerBuffer.alloc(ShardedReplay.MAILBOX_BYTES, s.device.index)[0] for s in shards]
        for s in shards:
            s._attach_exchange(boxes[s.rank], boxes)

    def enable_early_publish(self, on=True):
        """Every priority update then ends by storing the refreshed record straight into the peers'
        mailboxes (per_tune "early_publish", include/per_b200.h): the next sample only waits for
        the records.  All ranks switch together, and nothing may modify the replay between an
        update and the next sample while it is on (the native calls refuse)."""
        _n.check(_n.lib.per_tune(self._h, b"early_publish", 1 if on else 0))

    def publish_record(self):
        _n.check(_n.lib.per_shard_publish(self._h, self._len, _n.stream_ptr()))

    def _attach(self):
        """Every tree-changing call rewrites the record for the current whole-replay len."""
        _n.check(_n.lib.per_shard_attach_record(self._h, _n.ptr(self._rec), self._len))

    def set_global_len(self, n, cursor=None):
        """For callers that fill the slices themselves (bench): the whole replay now holds n rows."""
        self._len = int(n)
        if cursor is not None:
            self._cursor = int(cursor)
        self._rec_fresh = False
        self._attach()

    @property
    def tree_capacity_local(self):
        return self.geo.cap_local

    # -- insert ---------------------------------------------------------------------------------
    def add_batch_device(self, obs, next_obs, actions, rewards, dones, priorities, extra=None):
        """The same call on every rank with the same full batch (n global adds with actor
        priorities, memory.py:334-346); each rank ingests the rows routed to its slice."""
        n = int(actions.numel())
        for r, lo, hi, _slot in self.geo.route(self._cursor, n):
            if r != self.rank:
                continue
            cut = lambda t: None if t is None else t[lo:hi].contiguous()
            parts = [cut(t) for t in (obs, next_obs, actions, rewards, dones, extra, priorities)]
            _n.check(_n.lib.per_add_batch(self._h, *[_n.ptr(t) for t in parts], None, 0.0, hi - lo, _n.stream_ptr()))
        self._cursor = (self._cursor + n) % self.geo.size
        self._len = min(self.geo.size, self._len + n)
        self._rec_fresh = False
        self._attach()

    # -- sample ---------------------------------------------------------------------------------
    def sample_device(self, batch, beta, uniforms, gather=True, out=None):
        """Global stratified sample of `batch` transitions (memory.py:252-298).  Returns full-batch
        CUDA tensors; entries of strata owned by other ranks hold idx -1 / weight 0 and their
        rows are left untouched.  `out`: optional dict of preallocated (possibly peer-mapped)
        tensors obs / next_obs / act / rew / done to gather into."""
        assert beta > 0
        u = uniforms if torch.is_tensor(uniforms) else torch.as_tensor(np.asarray(uniforms, dtype=np.float64))
        if u.device != self.device or u.dtype != torch.float64:
            u = u.to(self.device, dtype=torch.float64)
        sp = _n.stream_ptr()
        if self._peer_exchange:
            recs = None  # the sampling kernel publishes / awaits the records through the mailboxes
        else:
            if not self._rec_fresh:  # inserts since the last update: recompute; otherwise the update left it current
                _n.check(_n.lib.per_shard_record(self._h, self._len, _n.ptr(self._rec), sp))
            recs = self._exchange(self._rec)
        pre = out or {}  # `out` may also carry preallocated idx / weights / weights64 (a loop that reuses its buffers)
        idx = pre["idx"] if "idx" in pre else torch.empty(batch, dtype=torch.int64, device=self.device)
        w32 = pre["weights"] if "weights" in pre else torch.empty(batch, dtype=torch.float32, device=self.device)
        w64 = pre["weights64"] if "weights64" in pre else torch.empty(batch, dtype=torch.float64, device=self.device)
        _n.check(_n.lib.per_sample_sharded(self._h, _n.ptr(recs), _n.ptr(u), batch, float(beta), self._len, _n.ptr(idx),
                                           _n.ptr(w64), _n.ptr(w32), sp))
        self._bound_idx = idx  # the native side binds this address to the sample until its update: keep it allocated
        res = {"idx": idx, "weights": w32, "weights64": w64, "records": recs}
        if gather:
            res.update(self.gather_device(idx, out))
        return res

    def gather_device(self, idx, out=None):
        b = idx.numel()
        if out is None:
            out = {
                "obs": torch.zeros((b,) + self.obs_shape, dtype=self.obs_dtype, device=self.device),
                "next_obs": torch.zeros((b,) + self.obs_shape, dtype=self.obs_dtype, device=self.device),
                "act": torch.zeros(b, dtype=torch.int64, device=self.device),
                "rew": torch.zeros(b, dtype=torch.float32, device=self.device),
                "done": torch.zeros(b, dtype=torch.float32, device=self.device),
            }
            if self.extra_shape is not None:
                out["a_mu"] = torch.zeros((b,) + tuple(self.extra_shape), dtype=self.extra_dtype, device=self.device)
        _n.check(_n.lib.per_gather(self._h, _n.ptr(idx), b, _n.ptr(out["obs"]), _n.ptr(out["next_obs"]), _n.ptr(out["act"]),
                                   _n.ptr(out["rew"]), _n.ptr(out["done"]), _n.ptr(out.get("a_mu")), _n.stream_ptr()))
        return out

    # -- host-buffer forms (the reference's own calling convention: numpy in, numpy out) -----------------------
    def _host_stage(self, batch):
        st = getattr(self, "_hstage", None)
        if st is None or st["batch"] < batch:
            dev, pin = dict(device=self.device), dict(pin_memory=True)
            shape = (batch,) + self.obs_shape
            st = {"batch": batch,
                  "d": {"obs": torch.empty(shape, dtype=self.obs_dtype, **dev), "next_obs": torch.empty(shape, dtype=self.obs_dtype, **dev),
                        "act": torch.empty(batch, dtype=torch.int64, **dev), "rew": torch.empty(batch, dtype=torch.float32, **dev),
                        "done": torch.empty(batch, dtype=torch.float32, **dev), "idx": torch.empty(batch, dtype=torch.int64, **dev),
                        "weights": torch.empty(batch, dtype=torch.float32, **dev), "weights64": torch.empty(batch, dtype=torch.float64, **dev)},
                  "u": torch.empty(batch, dtype=torch.float64, **dev), "u_pin": torch.empty(batch, dtype=torch.float64, **pin),
                  "up_idx": torch.empty(batch, dtype=torch.int64, **dev), "up_leaf": torch.empty(batch, dtype=torch.float64, **dev),
                  "up_pin_idx": torch.empty(batch, dtype=torch.int64, **pin), "up_pin_leaf": torch.empty(batch, dtype=torch.float64, **pin),
                  "h": {"obs": torch.empty(shape, dtype=self.obs_dtype, **pin), "next_obs": torch.empty(shape, dtype=self.obs_dtype, **pin),
                        "act": torch.empty(batch, dtype=torch.int64, **pin), "rew": torch.empty(batch, dtype=torch.float32, **pin),
                        "done": torch.empty(batch, dtype=torch.float32, **pin), "idx": torch.empty(batch, dtype=torch.int64, **pin),
                        "weights64": torch.empty(batch, dtype=torch.float64, **pin)}}
            self._hstage = st
        return st

    def sample_host(self, batch, beta, uniforms):
        """PrioritizedReplayBuffer.sample (memory.py:252-298) for one rank of the sharded replay, host buffers in and out:
        `uniforms` (the global batch's draws, host float64) go up, this rank's share of the batch comes back as numpy views
        of pinned staging (valid until the next call): {"lo", "hi"} = the strata [lo, hi) this rank owns, and for those
        strata obs, next_obs, act, rew, done, weights (float64) and idx (global indices).  The union over ranks is the
        reference's batch."""
        st = self._host_stage(batch)
        st["u_pin"].numpy()[:batch] = np.asarray(uniforms, dtype=np.float64)
        st["u"][:batch].copy_(st["u_pin"][:batch], non_blocking=True)
        d = self.sample_device(batch, beta, st["u"][:batch], gather=True, out={k: v[:batch] for k, v in st["d"].items()})
        st["h"]["idx"][:batch].copy_(d["idx"], non_blocking=True)
        torch.cuda.current_stream().synchronize()
        idx_h = st["h"]["idx"].numpy()[:batch]
        own = np.flatnonzero(idx_h >= 0)
        lo, hi = (int(own[0]), int(own[-1]) + 1) if own.size else (0, 0)
        assert hi - lo == own.size, "the strata one shard owns are contiguous"
        out = {"lo": lo, "hi": hi, "idx": idx_h[lo:hi]}
        for k in ("obs", "next_obs", "act", "rew", "done", "weights64"):
            st["h"][k][: hi - lo].copy_(st["d"][k][lo:hi], non_blocking=True)
            out["weights" if k == "weights64" else k] = st["h"][k].numpy()[: hi - lo]
        torch.cuda.current_stream().synchronize()
        return out

    def update_priorities_host(self, idxes, priorities):
        """update_priorities (memory.py:300-320) with the GLOBAL batch's host arrays on every rank: entries of other
        ranks' slices are masked out here, leaves are priority ** alpha by the C library's pow (bit-identical trees)."""
        idx = np.asarray(idxes, dtype=np.int64)
        pr = np.asarray(priorities, dtype=np.float64)
        assert idx.shape == pr.shape and np.all(pr > 0) and np.all((idx >= 0) & (idx < self._len))
        n = idx.size
        st = self._host_stage(n)
        mine = (idx // self.geo.cap_local) == self.rank
        st["up_pin_idx"].numpy()[:n] = np.where(mine, idx, -1)
        leaves = np.empty(n, dtype=np.float64)
        _n.check(_n.lib.per_pow_alpha_host(np.ascontiguousarray(pr).ctypes.data, float(self.alpha), leaves.ctypes.data, n))
        st["up_pin_leaf"].numpy()[:n] = leaves
        st["up_idx"][:n].copy_(st["up_pin_idx"][:n], non_blocking=True)
        st["up_leaf"][:n].copy_(st["up_pin_leaf"][:n], non_blocking=True)
        self.update_leaves_device(st["up_idx"][:n], st["up_leaf"][:n], float(pr.max()))

    # -- learner-ward delivery over NVLink ---------------------------------------------------------
    def open_learner_batch(self, batch, learner_rank=0, group=None):
        """Collective, once: the learner rank allocates the batch buffers, everybody maps them."""
        import torch.distributed as dist
        el = torch.empty(0, dtype=self.obs_dtype).element_size()
        row = int(np.prod(self.obs_shape))
        spec = {"obs": (self.obs_dtype, (batch,) + self.obs_shape, batch * row * el),
                "next_obs": (self.obs_dtype, (batch,) + self.obs_shape, batch * row * el),
                "act": (torch.int64, (batch,), batch * 8), "rew": (torch.float32, (batch,), batch * 4),
                "done": (torch.float32, (batch,), batch * 4), "idx": (torch.int64, (batch,), batch * 8),
                "weights": (torch.float32, (batch,), batch * 4), "prio": (torch.float32, (batch,), batch * 4)}
        box = [None]
        bufs = {}
        if self.rank == learner_rank:
            handles = {}
            for k, (_, _, nb) in spec.items():
                bufs[k], handles[k] = PeerBuffer.alloc(nb, self.device.index)
            box = [handles]
        dist.broadcast_object_list(box, src=learner_rank, group=group)
        if self.rank != learner_rank:
            bufs = {k: PeerBuffer.open(box[0][k], spec[k][2], self.device.index) for k in spec}
        views = {k: bufs[k].view(spec[k][0], spec[k][1]) for k in spec}
        dist.barrier(group=group)
        return LearnerBatch(bufs, views, batch, learner_rank, self.rank == learner_rank, group)

    def sample_to_learner(self, lb, beta, uniforms):
        """One global stratified sample delivered to the learner GPU: this rank samples (the record exchange is in the
        kernel), gathers the rows it owns straight into the learner's obs / next_obs / act / rew / done, and stores its
        strata's indices and importance weights into the learner's idx / weights.  Returns this rank's own result dict
        (idx holds -1 for strata of other ranks) — keep its `idx`: update_from_learner wants that buffer back."""
        d = self.sample_device(lb.batch, beta, uniforms, gather=False)
        sp = _n.stream_ptr()
        _n.check(_n.lib.per_gather(self._h, _n.ptr(d["idx"]), lb.batch, _n.ptr(lb.bufs["obs"]), _n.ptr(lb.bufs["next_obs"]),
                                   _n.ptr(lb.bufs["act"]), _n.ptr(lb.bufs["rew"]), _n.ptr(lb.bufs["done"]), None, sp))
        _n.check(_n.lib.per_shard_deliver(self._h, _n.ptr(d["idx"]), _n.ptr(d["weights"]), lb.batch, _n.ptr(lb.bufs["idx"]),
                                          _n.ptr(lb.bufs["weights"]), sp))
        return d

    def update_from_learner(self, lb, idx):
        """The learner has written the batch's new priorities into lb["prio"] (and fenced): every rank applies the
        entries of the strata it owns, reading them out of the learner's memory over NVLink.  `idx`: the index buffer
        sample_to_learner returned on this rank."""
        _n.check(_n.lib.per_update_priorities(self._h, _n.ptr(idx), _n.ptr(lb.bufs["prio"]), _n.PRIO_F32, lb.batch, _n.stream_ptr()))
        self._rec_fresh = True

    # -- update ---------------------------------------------------------------------------------
    def update_priorities_device(self, idx, priorities):
        """Full-batch global indices (-1 = not mine) and priorities (float32 / float64 CUDA)."""
        kind = _n.PRIO_F64 if priorities.dtype == torch.float64 else _n.PRIO_F32
        _n.check(_n.lib.per_update_priorities(self._h, _n.ptr(idx), _n.ptr(priorities), kind, idx.numel(), _n.stream_ptr()))
        self._rec_fresh = True

    def update_leaves_device(self, idx, leaves, batch_max):
        _n.check(_n.lib.per_update_leaves(self._h, _n.ptr(idx), _n.ptr(leaves), float(batch_max), idx.numel(),
                                          _n.stream_ptr()))
        self._rec_fresh = True

    def td_priority_update(self, q, q_next, q_tgt, actions, rewards, dones, weights, n_steps, gamma, idx, out_prio=None,
                           out_loss=None):
        """utils.py:71-80 over the FULL batch (the 0.9*max term is batch-wide) + the tree update
        of this rank's strata.  All ranks pass the same full-batch tensors."""
        b = int(actions.numel())
        prio = out_prio if out_prio is not None else torch.empty(b, dtype=torch.float32, device=self.device)
        loss = out_loss if out_loss is not None else torch.empty(1, dtype=torch.float32, device=self.device)
        gamma_n = float(np.float32(gamma ** n_steps))
        _n.check(_n.lib.per_td_priority_update(self._h, _n.ptr(q), _n.ptr(q_next), _n.ptr(q_tgt), int(q.shape[1]),
                                               _n.ptr(actions), _n.ptr(rewards), _n.ptr(dones), _n.ptr(weights), gamma_n,
                                               _n.ptr(idx), b, _n.ptr(prio), None, _n.ptr(loss), _n.stream_ptr()))
        self._rec_fresh = True
        return loss, prio

    def tree_values(self, which):
        out = torch.empty(2 * self.geo.cap_local, dtype=torch.float64, device=self.device)
        _n.check(_n.lib.per_tree_read(self._h, which, _n.ptr(out), _n.stream_ptr()))
        return out.cpu().numpy()

    def status(self):
        w = _n.C.c_int64(0)
        _n.check(_n.lib.per_status(self._h, _n.C.byref(w), _n.stream_ptr()))
        return w.value
