"""Rows SURVEY.md §8(f) marks "next": replay-server operations with the upstream wire format
(origin_repo/replay.py:19-45) and snapshot / restore of the GPU replay."""
import _pickle as pickle

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import __graft_entry__ as entry
    entry.build()
    import memory
    import replay_checkpoint
    import replay_ops
    return memory, replay_ops, replay_checkpoint


def bits(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float64)).view(np.uint64)


def actor_chunk(rng, n, shape=(4, 84, 84)):
    states = [rng.integers(0, 256, size=shape, dtype=np.uint8) for _ in range(n)]
    nexts = [rng.integers(0, 256, size=shape, dtype=np.uint8) for _ in range(n)]
    batch = [states, [int(a) for a in rng.integers(0, 18, size=n)], [float(r) for r in rng.standard_normal(n)], nexts,
             [np.float32(d) for d in (rng.random(n) < 0.1)]]
    prios = np.abs(rng.standard_normal(n)) + 1e-6
    return batch, prios


def test_replay_server_concurrent_workers_match_the_oracle(mods):
    """replay_ops.ReplayServer with real concurrency: three actor threads pushing chunks, a learner thread that samples and
    returns priorities, all at once, no global lock around the operations.  The server logs the order in which the
    operations were enqueued; the same operations in that order through the oracle's sequential port of the reference
    buffer (oracle/per_oracle.py ReplayOracle) must give the same sampled indices, the same rows and, bit for bit, the
    same trees — i.e. the short enqueue lock + stream order keep the reference's sequential semantics."""
    import threading
    from oracle import per_oracle as po
    memory, ops, _ = mods
    size = 1500
    buf = memory.CustomPrioritizedReplayBuffer(size, 0.6)
    log = []
    server = ops.ReplayServer(buf, n_senders=2, log=log)
    rngs = [np.random.default_rng(100 + t) for t in range(3)]
    warm = actor_chunk(np.random.default_rng(99), 200, shape=(2, 8))
    server.push_batch(pickle.dumps(warm))
    samples, errors = [], []

    def actor(t):
        try:
            for _ in range(12):
                server.push_batch(pickle.dumps(actor_chunk(rngs[t], 40, shape=(2, 8))))
        except Exception as exc:  # noqa: BLE001
            errors.append(exc)

    def learner():
        try:
            rng = np.random.default_rng(7)
            for _ in range(15):
                out = pickle.loads(server.sample_batch(64, 0.4))
                samples.append(out)
                server.update_prios(pickle.dumps((out[-1], (np.abs(rng.standard_normal(64)) + 1e-6).astype(np.float32))))
        except Exception as exc:  # noqa: BLE001
            errors.append(exc)

    threads = [threading.Thread(target=actor, args=(t,)) for t in range(3)] + [threading.Thread(target=learner)]
    for th in threads:
        th.start()
    for th in threads:
        th.join(timeout=120)
    assert not errors, errors
    assert [s for s, _, _ in log] == list(range(len(log))) and len(log) == 1 + 36 + 30
    ora = po.ReplayOracle(size, 0.6, "apex")
    k = 0
    for _, kind, payload in log:
        if kind == "push":
            c = payload
            for i in range(c["n"]):
                ora.add(c["obs"][i], int(c["act"][i]), float(c["rew"][i]), c["next"][i], float(c["done"][i]), 0.0)
                # the chunk carries leaves (priority ** alpha, libm); put them in as the reference's add would
                slot = (ora.cursor - 1) % size
                ora.sum_tree.set(slot, float(c["leaf"][i]))
                ora.min_tree.set(slot, float(c["leaf"][i]))
            ora.max_priority = max(ora.max_priority, c["max"])
        elif kind == "sample":
            n, beta, u, idxes = payload
            want = ora.sample(n, beta, u)
            assert list(want[-1]) == idxes == list(samples[k][-1])
            np.testing.assert_allclose(samples[k][-2], want[-2], rtol=1e-5)
            for got_col, want_col in zip(samples[k][:5], want[:5]):
                assert np.array_equal(np.asarray(got_col), np.asarray(want_col, dtype=np.asarray(got_col).dtype))
            k += 1
        else:
            idx, pr = payload
            ora.update_priorities([int(i) for i in idx], [float(p) for p in pr])
    assert k == 15 and len(buf) == len(ora) and buf._next_idx == ora.cursor
    assert np.array_equal(bits(buf._it_sum._value)[1:], bits(np.array(ora.sum_tree.v))[1:])
    assert np.array_equal(bits(buf._it_min._value)[1:], bits(np.array(ora.min_tree.v))[1:])
    assert buf._max_priority == ora.max_priority


def test_replay_server_ops_wire_format(mods):
    memory, ops, _ = mods
    rng = np.random.default_rng(0)
    a, b = memory.CustomPrioritizedReplayBuffer(400, 0.6), memory.CustomPrioritizedReplayBuffer(400, 0.6)
    for _ in range(11):  # 550 transitions: wraps the ring
        batch, prios = actor_chunk(rng, 50)
        ops.push_batch(a, pickle.dumps((batch, prios)))
        for sample in zip(*batch, prios):  # what origin_repo/replay.py:23-25 does
            b.add(*sample)
    assert len(a) == len(b) == 400 and a._next_idx == b._next_idx
    assert np.array_equal(bits(a._it_sum._value), bits(b._it_sum._value))
    import random
    random.seed(5)
    wire = ops.sample_batch(a, 64, 0.4)
    random.seed(5)
    want = b.sample(64, 0.4)
    got = pickle.loads(wire)
    assert len(got) == 7 and list(got[-1]) == list(want[-1])
    for x, y in zip(got[:-1], want[:-1]):
        assert np.array_equal(np.asarray(x), np.asarray(y))
    newp = np.abs(rng.standard_normal(64)).astype(np.float32) + 1e-6
    ops.update_prios(a, pickle.dumps((got[-1], newp)))
    b.update_priorities(want[-1], newp)
    assert np.array_equal(bits(a._it_sum._value), bits(b._it_sum._value)) and a._max_priority == b._max_priority


@pytest.mark.parametrize("cls_name,fill", [("CustomPrioritizedReplayBuffer", 700), ("CustomPrioritizedReplayBuffer", 260),
                                           ("CustomPrioritizedReplayBuffer_AQL", 333)])
def test_snapshot_restore_is_exact(mods, tmp_path, cls_name, fill):
    memory, _, ckpt = mods
    rng = np.random.default_rng(1)
    size = 500
    buf = getattr(memory, cls_name)(size, 0.6)
    for i in range(fill):
        o, o2 = rng.integers(0, 256, size=(4, 12, 12), dtype=np.uint8), rng.integers(0, 256, size=(4, 12, 12), dtype=np.uint8)
        if cls_name.endswith("_AQL"):
            buf.add(o, int(i % 5), float(i % 3), o2, float(i % 17 == 0), rng.standard_normal(6).astype(np.float32))
        else:
            buf.add(o, int(i % 5), float(i % 3), o2, float(i % 17 == 0), float(abs(rng.standard_normal()) + 1e-6))
    idx = list(rng.integers(0, len(buf), size=40))
    buf.update_priorities(idx, np.abs(rng.standard_normal(40)) + 1e-6)
    path = str(tmp_path / "replay.npz")
    ckpt.save(buf, path)
    back = ckpt.load(path, memory)
    assert type(back) is type(buf) and len(back) == len(buf) and back._next_idx == buf._next_idx
    assert back._max_priority == buf._max_priority
    assert np.array_equal(bits(back._it_sum._value), bits(buf._it_sum._value))
    assert np.array_equal(bits(back._it_min._value), bits(buf._it_min._value))
    probe = list(range(0, len(buf), 13))
    for x, y in zip(buf._encode_sample(probe), back._encode_sample(probe)):
        assert np.array_equal(x, y)
    # both continue identically: same draws, same next insert slot
    import random
    random.seed(9); s1 = buf.sample(32, 0.5)
    random.seed(9); s2 = back.sample(32, 0.5)
    assert list(s1[-1]) == list(s2[-1]) and np.array_equal(s1[-2], s2[-2])


@pytest.mark.parametrize("dtype,scale", [(torch.float32, 1.0), (torch.bfloat16, 1.0 / 255.0), (torch.float16, 1.0)])
def test_gather_cast_channels_last_equals_torch(mods, dtype, scale):
    """SURVEY section 8(f) row 3, second half: the widened batch in channels-last (NHWC) memory, as the learner's
    convolution trunk wants it (ApeX.py:50-55 -> model.py:33), against torch's own conversion of the plain uint8 gather:
    x.to(dtype) * scale, .contiguous(memory_format=torch.channels_last) — same values, same strides, same bytes."""
    memory = mods[0]
    size, batch = 3000, 257
    gen = torch.Generator(device="cuda").manual_seed(4)
    buf = memory.PrioritizedReplayBuffer(size, 0.6)
    obs = torch.randint(0, 256, (size, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=gen)
    nxt = torch.randint(0, 256, (size, 4, 84, 84), dtype=torch.uint8, device="cuda", generator=gen)
    buf.add_batch_device(obs, nxt, torch.arange(size, device="cuda"), torch.zeros(size, device="cuda"), torch.zeros(size, device="cuda"))
    u = np.random.default_rng(2).random(batch)
    plain = buf.sample_device(batch, 0.5, uniforms=u)
    d = buf.sample_device(batch, 0.5, uniforms=u, obs_dtype=dtype, obs_scale=scale, channels_last=True)
    assert torch.equal(d["idx"], plain["idx"]) and torch.equal(d["act"], plain["act"])
    for k in ("obs", "next_obs"):
        want = (plain[k].to(torch.float32) * scale).to(dtype).contiguous(memory_format=torch.channels_last)
        got = d[k]
        assert got.shape == want.shape and got.stride() == want.stride() and got.is_contiguous(memory_format=torch.channels_last)
        assert torch.equal(got, want)
        assert torch.equal(got.permute(0, 2, 3, 1).contiguous().view(-1), want.permute(0, 2, 3, 1).contiguous().view(-1))
    # and it feeds a convolution without a layout change
    conv = torch.nn.Conv2d(4, 8, 8, 4).cuda().to(dtype).to(memory_format=torch.channels_last)
    assert conv(d["obs"]).shape == (batch, 8, 20, 20)
