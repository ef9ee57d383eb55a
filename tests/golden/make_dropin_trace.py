"""Runs the reference's four drivers UNMODIFIED and records everything they do to their replay buffer.

Run in the build container only (needs /root/reference):

    python tests/golden/make_dropin_trace.py            # writes tests/golden/dropin_trace.npz

`ApeX.py`, `DQN.py`, `AQL.py`, `AQL_dis.py` (and through them `model.py`, `utils.py`, `batchrecorder.py`,
`batchrecoder_AQL.py`) are imported from /root/reference as they are; `gym` and `tensorboardX`, which are not
installed here, come from tests/stubs/ (synthetic episodic environments, a list-backed SummaryWriter).  The drivers'
`from memory import ...` resolves to a RECORDING module built here: the reference's own memory.py (NumPy-2 shim of
SURVEY.md section 8c) with each buffer class subclassed so that every `add`, `sample`, `update_priorities` call — its
arguments, the uniforms `sample` draws from the global `random`, and what it returns — is appended to a trace.

The trace is the drop-in contract in executable form: tests/test_gpu_dropin.py replays it against
ape-x_b200/memory.py on the GPU box (where the reference does not exist) and demands the reference's answers;
tests/test_dropin_reference_cpu.py (here) re-runs the drivers and checks that the committed trace is what they do, and
that every recorded call binds to the drop-in classes' signatures.
"""
import importlib.util
import os
import random
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
TESTS = os.path.dirname(HERE)
REF = "/root/reference"
OUT = os.path.join(HERE, "dropin_trace.npz")

DRIVERS = ("apex", "dqn", "aql", "aql_dis")


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


class Trace(object):
    """Events of one buffer, in call order: ("init", cls, size, alpha) | ("add", row) | ("sample", batch, beta, uniforms,
    outputs, len) | ("update", idxes, priorities)."""

    def __init__(self):
        self.events = []


def recording_memory(trace):
    """A module object that looks like memory.py to the drivers and records into `trace`."""
    mg = _load("make_golden_for_trace", os.path.join(HERE, "make_golden.py"))
    real = mg.load_reference("memory")
    real.np = mg.numpy1_shim()
    draws = []

    def recorded_random():  # memory.py:247 draws from the global RNG; keep the stream, remember the draws
        u = random.random()
        draws.append(u)
        return u

    real.random = types.SimpleNamespace(random=recorded_random, randint=random.randint)
    mod = types.ModuleType("memory")
    mod.__file__ = os.path.join(REF, "memory.py")
    for name in ("SegmentTree", "SumSegmentTree", "MinSegmentTree", "ReplayBuffer", "BatchStorage"):
        setattr(mod, name, getattr(real, name))

    def make(cls_name):
        base = getattr(real, cls_name)

        class Recording(base):
            def __init__(self, size, alpha):
                base.__init__(self, size, alpha)
                trace.events.append(("init", cls_name, float(size), float(alpha)))

            def add(self, *args, **kwargs):
                assert not kwargs
                trace.events.append(("add", tuple(np.array(a, copy=True) if isinstance(a, np.ndarray) else a for a in args)))
                return base.add(self, *args)

            def sample(self, batch_size, beta):
                del draws[:]
                out = base.sample(self, batch_size, beta)
                trace.events.append(("sample", int(batch_size), float(beta), np.array(draws), out, len(self)))
                return out

            def update_priorities(self, idxes, priorities):
                # NumPy 2 keeps float32 ** float in float32; the reference was written against NumPy 1.16.2, where the
                # same expression is float64: hand it float64, which reproduces the pinned behaviour
                pr = np.asarray(priorities, dtype=np.float64)
                trace.events.append(("update", np.asarray(idxes, dtype=np.int64).copy(), np.asarray(priorities).copy()))
                return base.update_priorities(self, idxes, pr)

        Recording.__name__ = cls_name
        return Recording

    for name in ("PrioritizedReplayBuffer", "CustomPrioritizedReplayBuffer", "CustomPrioritizedReplayBuffer_AQL"):
        setattr(mod, name, make(name))
    return mod


def run_driver(which):
    """Imports one reference driver fresh (with the recording memory module) and runs a short training session."""
    import torch
    trace = Trace()
    for m in ("memory", "model", "utils", "batchrecorder", "batchrecoder_AQL", "ApeX", "DQN", "AQL", "AQL_dis", "gym", "tensorboardX"):
        sys.modules.pop(m, None)
    sys.modules["memory"] = recording_memory(trace)
    added = [os.path.join(TESTS, "stubs"), REF]
    sys.path[:0] = added
    cwd = os.getcwd()
    tmp = tempfile.mkdtemp(prefix="dropin_")
    os.chdir(tmp)  # the drivers save model*.pth into the working directory
    try:
        random.seed(11)
        np.random.seed(11)
        torch.manual_seed(11)
        if which == "apex":
            drv = _load("ApeX", os.path.join(REF, "ApeX.py"))
            t = drv.train_DQN("Chain-v0", seed=3, n_workers=2, batch_size=32, max_step=4, save_interval=1000, publish_param_interval=2)
            for _ in range(3):
                t.sampling_data()  # BatchRecorder.record_batch(): workers' n-step chunks -> buffer.add(..., priority)
            t.train()
            buf = t.buffer
        elif which == "dqn":
            drv = _load("DQN", os.path.join(REF, "DQN.py"))
            t = drv.train_DQN("Chain-v0", max_step=70, batch_size=32, save_interval=1000)
            t.train()
            buf = t.replay_buffer
        elif which == "aql":
            drv = _load("AQL", os.path.join(REF, "AQL.py"))
            t = drv.train_DQN("ChainCont-v0", max_step=45, batch_size=16, save_interval=1000, propose_sample=3, uniform_sample=4)
            t.train()
            buf = t.replay_buffer
        else:
            drv = _load("AQL_dis", os.path.join(REF, "AQL_dis.py"))
            t = drv.train_DQN("ChainCont-v0", max_step=2, batch_size=16, save_interval=1000, device="cpu", propose_sample=2,
                              uniform_sample=3, n_workers=2)
            t.train()
            buf = t.replay_buffer
        cap, n = buf._it_sum._capacity, len(buf)
        # the node arrays are pure functions of the leaves (memory.py:82-87), so the fixture carries the written leaves
        # and the two roots; the replay test rebuilds the nodes with the oracle (AQL_dis.py's tree has 2^24 leaves)
        final = {"leaves_sum": np.array(buf._it_sum._value[cap:cap + n], dtype=np.float64),
                 "leaves_min": np.array(buf._it_min._value[cap:cap + n], dtype=np.float64),
                 "root_sum": float(buf._it_sum._value[1]), "root_min": float(buf._it_min._value[1]), "cap": cap,
                 "len": n, "next_idx": int(buf._next_idx), "max_priority": float(buf._max_priority)}
        return trace.events, final
    finally:
        os.chdir(cwd)
        for p in added:
            sys.path.remove(p)
        for m in ("memory", "model", "utils", "batchrecorder", "batchrecoder_AQL", "ApeX", "DQN", "AQL", "AQL_dis", "gym", "tensorboardX"):
            sys.modules.pop(m, None)


def run_replay_server():
    """origin_repo/replay.py UNMODIFIED: its three worker coroutines (replay.py:77-146) with their global asyncio.Lock and
    thread-pool executor, fed through in-memory stand-ins for the ZeroMQ DEALER sockets by a script that plays the actors
    (100 chunks of 8 n-step transitions, the wire tuple of origin_repo/actor.py:105-109) and the learner (batch
    requests, priorities sent back as float32 like origin_repo/learner.py).  The buffer it holds is the recording one."""
    import asyncio
    import _pickle as pickle
    from concurrent.futures import ThreadPoolExecutor
    trace = Trace()
    mods = ("memory", "utils", "arguments", "replay", "model", "gym", "tensorboardX")
    for m in mods:
        sys.modules.pop(m, None)
    sys.modules["memory"] = recording_memory(trace)
    added = [os.path.join(TESTS, "stubs"), os.path.join(REF, "origin_repo"), REF]
    sys.path[:0] = added
    chunk_rows, n_chunks = 8, 100
    try:
        replay = _load("replay", os.path.join(REF, "origin_repo", "replay.py"))

        class FakeSocket(object):
            def __init__(self):
                self.inbox, self.sent = asyncio.Queue(), asyncio.Queue()

            def connect(self, addr):
                self.addr = addr

            async def recv_multipart(self, copy=False):
                return await self.inbox.get()

            async def send_multipart(self, parts, copy=True):
                await self.sent.put(parts)

        sockets = []

        class FakeContext(object):
            @staticmethod
            def instance():
                return FakeContext()

            def socket(self, kind):
                s = FakeSocket()
                sockets.append(s)
                return s

        replay.Context = FakeContext  # replay.py:13 `from zmq.asyncio import Context`; everything else is the file's own code
        rng = np.random.default_rng(5)
        random.seed(5)

        async def session():
            buffer = sys.modules["memory"].CustomPrioritizedReplayBuffer(2000, 0.6)
            exe, event, lock = ThreadPoolExecutor(), asyncio.Event(), asyncio.Lock()
            tasks = [asyncio.ensure_future(replay.recv_batch_worker(buffer, exe, event, lock, 500)),
                     asyncio.ensure_future(replay.recv_prios_worker(buffer, exe, event, lock)),
                     asyncio.ensure_future(replay.send_batch_worker(buffer, exe, event, lock, 64, 0.4))]
            await asyncio.sleep(0.05)  # the workers create and connect their sockets
            by_addr = {s.addr: s for s in sockets}
            batch_in, prios_in, batch_out = by_addr["ipc:///tmp/5101.ipc"], by_addr["ipc:///tmp/5102.ipc"], by_addr["ipc:///tmp/5103.ipc"]
            for c in range(n_chunks):  # actors: chunks of n-step transitions with their priorities
                states = [rng.standard_normal(4).astype(np.float32) for _ in range(chunk_rows)]
                nexts = [rng.standard_normal(4).astype(np.float32) for _ in range(chunk_rows)]
                batch = [states, [int(a) for a in rng.integers(0, 3, size=chunk_rows)], [float(r) for r in rng.standard_normal(chunk_rows)],
                         nexts, [np.float32(d) for d in (rng.random(chunk_rows) < 0.1)]]
                prios = np.abs(rng.standard_normal(chunk_rows)) + 1e-6
                await batch_in.inbox.put((b"actor", pickle.dumps((batch, prios))))
                await batch_in.sent.get()  # the ack (replay.py:94)
            await asyncio.wait_for(event.wait(), 5)  # threshold reached at the 100th chunk (replay.py:99-106)
            for step in range(6):  # learner: ask for a batch, train, send priorities back; one more chunk arrives in between
                await batch_out.inbox.put((b"learner", b""))
                ident, data = await batch_out.sent.get()
                got = pickle.loads(data)
                newp = (np.abs(rng.standard_normal(64)) + 1e-6).astype(np.float32)
                await prios_in.inbox.put((b"learner", pickle.dumps((got[-1], newp))))
                await prios_in.sent.get()
                if step % 2 == 1:
                    states = [rng.standard_normal(4).astype(np.float32) for _ in range(chunk_rows)]
                    batch = [states, [1] * chunk_rows, [0.5] * chunk_rows, states, [np.float32(0)] * chunk_rows]
                    await batch_in.inbox.put((b"actor", pickle.dumps((batch, np.full(chunk_rows, 0.25)))))
                    await batch_in.sent.get()
            for t in tasks:
                t.cancel()
            exe.shutdown(wait=False)
            return buffer

        buf = asyncio.run(session())
        cap, n = buf._it_sum._capacity, len(buf)
        final = {"leaves_sum": np.array(buf._it_sum._value[cap:cap + n], dtype=np.float64),
                 "leaves_min": np.array(buf._it_min._value[cap:cap + n], dtype=np.float64),
                 "root_sum": float(buf._it_sum._value[1]), "root_min": float(buf._it_min._value[1]), "cap": cap,
                 "len": n, "next_idx": int(buf._next_idx), "max_priority": float(buf._max_priority)}
        return trace.events, final, chunk_rows
    finally:
        for p in added:
            sys.path.remove(p)
        for m in mods:
            sys.modules.pop(m, None)


def pack(events, final, prefix, out):
    """Flattens one driver's events into arrays (no pickled objects in the fixture)."""
    kinds, a, b = [], [], []
    adds, samples, updates = [], [], []
    init = None
    for ev in events:
        if ev[0] == "init":
            init = ev
        elif ev[0] == "add":
            if kinds and kinds[-1] == 0:
                a[-1] += 1
            else:
                kinds.append(0), a.append(1), b.append(0.0)
            adds.append(ev[1])
        elif ev[0] == "sample":
            kinds.append(1), a.append(ev[1]), b.append(ev[2])
            samples.append(ev)
        else:
            kinds.append(2), a.append(len(ev[1])), b.append(0.0)
            updates.append(ev)
    p = prefix + "_"
    out[p + "cls"], out[p + "size"], out[p + "alpha"] = np.array(init[1]), init[2], init[3]
    out[p + "ev_kind"], out[p + "ev_n"], out[p + "ev_beta"] = np.array(kinds), np.array(a), np.array(b)
    out[p + "add_obs"] = np.stack([np.asarray(r[0]) for r in adds])
    out[p + "add_act"] = np.array([int(r[1]) for r in adds], dtype=np.int64)
    out[p + "add_rew"] = np.array([float(r[2]) for r in adds], dtype=np.float64)
    out[p + "add_next"] = np.stack([np.asarray(r[3]) for r in adds])
    out[p + "add_done"] = np.array([float(r[4]) for r in adds], dtype=np.float64)
    if len(adds[0]) > 5:
        extra = [np.asarray(r[5]) for r in adds]
        out[p + "add_extra"] = np.stack(extra) if extra[0].ndim else np.array([float(x) for x in extra], dtype=np.float64)
    out[p + "s_u"] = np.concatenate([s[3] for s in samples])
    out[p + "s_len"] = np.array([s[5] for s in samples])
    n_out = len(samples[0][4])
    out[p + "s_idx"] = np.concatenate([np.asarray(s[4][-1], dtype=np.int64) for s in samples])
    out[p + "s_w"] = np.concatenate([np.asarray(s[4][-2], dtype=np.float64) for s in samples])
    for col in range(n_out - 2):
        out[p + "s_col%d" % col] = np.concatenate([np.asarray(s[4][col]) for s in samples])
    out[p + "u_idx"] = np.concatenate([u[1] for u in updates])
    out[p + "u_prio"] = np.concatenate([u[2] for u in updates])
    out[p + "u_prio_dtype"] = np.array(str(updates[0][2].dtype))
    for k, v in final.items():
        out[p + "final_" + k] = v


def generate():
    import torch
    torch.set_num_threads(1)  # the actor pools fork: a forked child must not inherit a running intra-op thread pool
    out = {}
    for which in DRIVERS:
        events, final = run_driver(which)
        pack(events, final, which, out)
    events, final, chunk_rows = run_replay_server()
    pack(events, final, "replay_server", out)
    out["replay_server_chunk_rows"] = chunk_rows
    return out


if __name__ == "__main__":
    fixture = generate()
    np.savez_compressed(OUT, **fixture)
    for d in DRIVERS + ("replay_server",):
        k = fixture[d + "_ev_kind"]
        print("%-8s %-34s adds %5d  samples %3d  updates %3d  final len %d" % (
            d, str(fixture[d + "_cls"]), int(fixture[d + "_add_act"].shape[0]), int((k == 1).sum()), int((k == 2).sum()),
            int(fixture[d + "_final_len"])))
    print(OUT, os.path.getsize(OUT), "bytes")
