"""Generate the golden fixtures in this directory from the REAL reference.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

It imports `/root/reference/memory.py` and `/root/reference/utils.py` unmodified
(under private module names so they never shadow the product's own `memory`),
applies the NumPy-2 shim of SURVEY.md §8(c) (`np.array(x, copy=False)` ->
`np.asarray(x)`, the NumPy-1.16 meaning the reference was written against),
replays pre-drawn uniforms through `memory.random.random`, always hands the
reference float64 / Python-float priorities, and writes the outputs
as small .npz files.  The fixtures travel to the GPU box; the reference does not.
"""
import hashlib
import importlib.util
import os
import random
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def load_reference(name):
    spec = importlib.util.spec_from_file_location("_ref_" + name, os.path.join(REF, name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def numpy1_shim():
    proxy = types.ModuleType("numpy_shim")
    proxy.__dict__.update(np.__dict__)

    def array(obj, *a, copy=True, **k):
        if copy is False:
            return np.asarray(obj, *a, **k)
        return np.array(obj, *a, copy=copy, **k)

    proxy.array = array
    return proxy


class Replay:
    """Replaces memory.random.random with a replay of a fixed uniform vector."""

    def __init__(self):
        self.q = []

    def load(self, u):
        self.q = list(u)[::-1]

    def random(self):
        return self.q.pop()


def digest(values):
    return np.frombuffer(hashlib.sha256(np.asarray(values, dtype=np.float64).tobytes()).digest(), dtype=np.uint8)


def frames(rng, n, shape=(4, 6, 6)):
    return rng.integers(0, 256, size=(n,) + shape, dtype=np.uint8)


def gen_tree_ops(mem):
    """SegmentTree point sets, range reductions with arbitrary (start, end), descents."""
    rng = np.random.default_rng(11)
    cap = 64
    st, mt = mem.SumSegmentTree(cap), mem.MinSegmentTree(cap)
    set_idx = rng.integers(0, cap, size=200)
    set_val = np.abs(rng.standard_normal(200)) + 1e-3
    for k, x in zip(set_idx, set_val):
        st[int(k)] = float(x)
        mt[int(k)] = float(x)
    ranges, rs, rm = [], [], []
    for _ in range(300):
        a = int(rng.integers(0, cap))
        b = int(rng.integers(a + 1, cap + 1))
        ranges.append((a, b))
        rs.append(st.sum(a, b))
        rm.append(mt.min(a, b))
    # negative `end` wraps (memory.py:71-72) and end=None means the whole array
    neg = [(0, -1), (3, -5), (0, -63)]
    neg_s = [st.sum(a, b) for a, b in neg]
    masses = rng.random(256) * st.sum()
    found = [st.find_prefixsum_idx(float(m)) for m in masses]
    # exact ties: mass equal to a left child goes right (strict >)
    tie_masses = [st._value[2], st._value[4], st._value[2] + st._value[6]]
    tie_found = [st.find_prefixsum_idx(float(m)) for m in tie_masses]
    np.savez(os.path.join(HERE, "tree_ops.npz"),
             cap=cap, set_idx=set_idx, set_val=set_val,
             sum_nodes=np.array(st._value), min_nodes=np.array(mt._value),
             ranges=np.array(ranges), range_sum=np.array(rs), range_min=np.array(rm),
             neg_ranges=np.array(neg), neg_sum=np.array(neg_s), full_sum=st.sum(), full_min=mt.min(),
             masses=masses, found=np.array(found), tie_masses=np.array(tie_masses), tie_found=np.array(tie_found))


def gen_per(mem, rnd, variant, fname, size, batch, rounds, n_fill, seed):
    """add / sample / update_priorities interleavings incl. wrap-around and duplicate idx."""
    rng = np.random.default_rng(seed)
    alpha, beta = 0.6, 0.4
    cls = {"per": mem.PrioritizedReplayBuffer, "apex": mem.CustomPrioritizedReplayBuffer,
           "aql": mem.CustomPrioritizedReplayBuffer_AQL}[variant]
    buf = cls(size, alpha)
    out = dict(size=size, alpha=alpha, beta=beta, batch=batch, rounds=rounds, n_fill=n_fill, seed=seed)
    K = 5  # candidate actions per row for the AQL variant

    def add_some(n, tag):
        obs, nxt = frames(rng, n), frames(rng, n)
        act = rng.integers(0, 18, size=n)
        rew = rng.integers(-1, 2, size=n).astype(np.float64)
        done = (rng.random(n) < 0.1).astype(np.float32)
        extra = np.abs(rng.standard_normal(n)) + 1e-6 if variant == "apex" else \
            rng.standard_normal((n, K)).astype(np.float32)
        for i in range(n):
            if variant == "per":
                buf.add(obs[i], int(act[i]), float(rew[i]), nxt[i], float(done[i]))
            elif variant == "apex":
                buf.add(obs[i], int(act[i]), float(rew[i]), nxt[i], float(done[i]), float(extra[i]))
            else:
                buf.add(obs[i], int(act[i]), float(rew[i]), nxt[i], float(done[i]), extra[i])
        out.update({tag + "_obs": obs, tag + "_next": nxt, tag + "_act": act, tag + "_rew": rew,
                    tag + "_done": done, tag + "_extra": extra})

    add_some(n_fill, "fill")
    out["after_fill_sum"] = np.array(buf._it_sum._value, dtype=np.float64)
    out["after_fill_min"] = np.array(buf._it_min._value, dtype=np.float64)
    for r in range(rounds):
        u = rng.random(batch)
        rnd.load(u)
        res = buf.sample(batch, beta)
        idxes = res[-1]
        out[f"r{r}_u"] = u
        out[f"r{r}_idx"] = np.array(idxes)
        out[f"r{r}_w"] = np.asarray(res[-2], dtype=np.float64)
        out[f"r{r}_obs"] = np.array(res[0])
        out[f"r{r}_act"] = np.array(res[1])
        out[f"r{r}_rew"] = np.array(res[2])
        out[f"r{r}_next"] = np.array(res[3])
        out[f"r{r}_done"] = np.array(res[4])
        if variant == "aql":
            out[f"r{r}_amu"] = np.array(res[5])
        pr = np.abs(rng.standard_normal(batch)) + 1e-6
        upd = list(idxes)
        if r % 2 == 1:  # arbitrary, unsorted indices with forced duplicates
            upd = [int(x) for x in rng.integers(0, len(buf), size=batch)]
            upd[3] = upd[0]
            upd[-1] = upd[0]
        buf.update_priorities(upd, pr)
        out[f"r{r}_upd_idx"] = np.array(upd)
        out[f"r{r}_upd_prio"] = pr
        out[f"r{r}_sum"] = np.array(buf._it_sum._value, dtype=np.float64)
        out[f"r{r}_min"] = np.array(buf._it_min._value, dtype=np.float64)
        out[f"r{r}_maxp"] = float(buf._max_priority)
        add_some(7, f"r{r}_add")  # a few more inserts between learner steps (wraps the ring)
        out[f"r{r}_cursor"] = int(buf._next_idx)
        out[f"r{r}_len"] = len(buf)
    out["final_sum"] = np.array(buf._it_sum._value, dtype=np.float64)
    out["final_min"] = np.array(buf._it_min._value, dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, fname), **out)


def gen_c1(mem, rnd):
    """BASELINE config 1: PrioritizedReplayBuffer(100000 -> 2^17), batch 512, alpha .6, beta .4.
    Inputs are regenerated from the seed by the test; only outputs are stored."""
    seed, size, batch, rounds = 2024, 100000, 512, 4
    rng = np.random.default_rng(seed)
    buf = mem.CustomPrioritizedReplayBuffer(size, 0.6)
    prio0 = np.abs(rng.standard_normal(size)) + 1e-6
    dummy = np.zeros(1, dtype=np.uint8)
    for i in range(size):
        buf.add(dummy, 0, 0.0, dummy, 0.0, float(prio0[i]))
    out = dict(seed=seed, size=size, batch=batch, rounds=rounds,
               fill_sum_sha=digest(buf._it_sum._value), fill_min_sha=digest(buf._it_min._value),
               fill_root_sum=buf._it_sum._value[1], fill_root_min=buf._it_min._value[1])
    for r in range(rounds):
        u = rng.random(batch)
        rnd.load(u)
        idxes = buf._sample_proportional(batch)
        n = len(buf)
        p_min = buf._it_min.min() / buf._it_sum.sum()
        max_w = (p_min * n) ** (-0.4)
        w = [((buf._it_sum[i] / buf._it_sum.sum() * n) ** (-0.4)) / max_w for i in idxes]
        pr = np.abs(rng.standard_normal(batch)) + 1e-6
        buf.update_priorities(idxes, pr)
        out[f"r{r}_idx"] = np.array(idxes, dtype=np.int32)
        out[f"r{r}_w"] = np.array(w)
        out[f"r{r}_sum_sha"] = digest(buf._it_sum._value)
        out[f"r{r}_min_sha"] = digest(buf._it_min._value)
        out[f"r{r}_root_sum"] = buf._it_sum._value[1]
        out[f"r{r}_root_min"] = buf._it_min._value[1]
    np.savez_compressed(os.path.join(HERE, "per_c1.npz"), **out)


def gen_nstep(mem):
    """BatchStorage streams: several episodes of varied length, n in {1,3,5}."""
    out = {}
    for n in (1, 3, 5):
        rng = np.random.default_rng(100 + n)
        bs = mem.BatchStorage(n, 0.99)
        T, nA = 120, 6
        rew = rng.integers(-1, 2, size=T).astype(np.float64)
        act = rng.integers(0, nA, size=T)
        q = rng.standard_normal((T, nA)).astype(np.float32)
        done = np.zeros(T, dtype=bool)
        # episode ends, chosen so a done never meets an empty window (that raises IndexError, Q10)
        for t in (17, 18 + n + 2, 60, 61 + n, 119):
            done[t] = True
        state_id = np.arange(T)  # states are identified by their step number
        for t in range(T):
            bs.add(int(state_id[t]), float(rew[t]), int(act[t]), bool(done[t]), q[t])
        batch, prios = bs.make_batch()
        out[f"n{n}_rew"], out[f"n{n}_act"], out[f"n{n}_q"], out[f"n{n}_done"] = rew, act, q, done
        out[f"n{n}_out_s"] = np.array(batch[0])
        out[f"n{n}_out_a"] = np.array(batch[1])
        out[f"n{n}_out_R"] = np.array(batch[2], dtype=np.float64)
        out[f"n{n}_out_sn"] = np.array(batch[3])
        out[f"n{n}_out_d"] = np.array(batch[4], dtype=np.float32)
        out[f"n{n}_out_prio"] = np.asarray(prios, dtype=np.float64)
    np.savez_compressed(os.path.join(HERE, "nstep.npz"), **out)


def gen_td(utl):
    """utils.compute_loss driven by table-lookup 'models' so the Q values are the inputs."""
    import torch
    out = {}
    for case, (B, nA, n_steps) in enumerate([(32, 18, 3), (512, 18, 3), (64, 4, 1)]):
        g = torch.Generator().manual_seed(500 + case)
        q = torch.randn(B, nA, generator=g)
        qn = torch.randn(B, nA, generator=g)
        qt = torch.randn(B, nA, generator=g)
        s = torch.zeros(B, 1)
        s2 = torch.ones(B, 1)
        online = lambda x, q=q, qn=qn: q if float(x[0, 0]) == 0.0 else qn
        target = lambda x, qt=qt: qt
        a = torch.randint(0, nA, (B,), generator=g)
        r = torch.randint(-1, 2, (B,), generator=g).float() * 1.7
        d = (torch.rand(B, generator=g) < 0.1).float()
        w = torch.rand(B, generator=g)
        loss, prios = utl.compute_loss(online, target, (s, a, r, s2, d, w), n_steps, 0.99)
        out[f"c{case}_q"], out[f"c{case}_qn"], out[f"c{case}_qt"] = q.numpy(), qn.numpy(), qt.numpy()
        out[f"c{case}_a"], out[f"c{case}_r"], out[f"c{case}_d"], out[f"c{case}_w"] = \
            a.numpy(), r.numpy(), d.numpy(), w.numpy()
        out[f"c{case}_n"] = n_steps
        out[f"c{case}_loss"] = float(loss)
        out[f"c{case}_prios"] = np.asarray(prios)
    np.savez_compressed(os.path.join(HERE, "td.npz"), **out)


def td_aql_inputs(case):
    """Inputs of the compute_loss_AQL fixture, regenerated from a seed (numpy's Generator streams are stable),
    so that the fixture file only carries the reference's outputs.  Shared with tests/test_gpu_parity.py."""
    B, K, n_steps = [(256, 1000, 1), (64, 1000, 3)][case]
    rng = np.random.default_rng(7700 + case)
    f = np.float32
    q, qn, qt = (rng.standard_normal((B, K)).astype(f) for _ in range(3))
    a = rng.integers(0, K, size=B).astype(np.int64)
    r = (rng.integers(-1, 2, size=B) * 1.3).astype(f)
    d = (rng.random(B) < 0.1).astype(f)
    w = rng.random(B).astype(f)
    return q, qn, qt, a, r, d, w, n_steps


def gen_td_aql(utl):
    """utils.compute_loss_AQL (utils.py:44-61) at AQL's real shape [B, K = 1000]: the 'models' are table look-ups,
    so the Q values are the inputs."""
    import torch
    out = {}
    for case in range(2):
        q, qn, qt, a, r, d, w, n_steps = td_aql_inputs(case)
        B = q.shape[0]
        tq, tqn, tqt = torch.from_numpy(q), torch.from_numpy(qn), torch.from_numpy(qt)
        s, s2 = torch.zeros(B, 1), torch.ones(B, 1)
        online = lambda x, a_mu: tq if float(x[0, 0]) == 0.0 else tqn
        target = lambda x, a_mu: tqt
        batch = (s, torch.from_numpy(a), torch.from_numpy(r), s2, torch.from_numpy(d), None, torch.from_numpy(w))
        loss, prios = utl.compute_loss_AQL(online, target, batch, n_steps, 0.99)
        out[f"c{case}_loss"] = float(loss)
        out[f"c{case}_prios"] = np.asarray(prios)
    np.savez_compressed(os.path.join(HERE, "td_aql.npz"), **out)


def gen_aql():
    """Q_Network.act (model.py:294-335) of the real reference, continuous and discrete, NoisyLinear
    in training mode.  model.py imports gym at module level; a two-line stand-in satisfies it
    (Q_Network itself never touches gym)."""
    import torch
    gym = types.ModuleType("gym")
    gym.spaces = types.SimpleNamespace(Box=type("Box", (), {}))
    sys.modules.setdefault("gym", gym)
    model = load_reference("model")
    out = {}
    for case, (cont, A, B, K, obs) in enumerate([(True, 2, 6, 45, 9), (False, 7, 5, 33, 4)]):
        torch.manual_seed(900 + case)
        net = model.Q_Network(input_shape=(obs,), num_actions=A, total_sample=K, env_iscontinuous=cont, device="cpu")
        net.train()
        state = torch.randn(B, obs)
        a_mu = (torch.rand(B, K, A) * 2 - 1) if cont else torch.randint(0, A, (B, K))
        with torch.no_grad():
            _, q = net.act(state, a_mu, 0)
            f = net.q_feature(state)
        c = f"c{case}_"
        out[c + "cont"], out[c + "A"] = cont, A
        out[c + "f"], out[c + "a_mu"], out[c + "q"] = f.numpy(), a_mu.numpy(), q.numpy()
        ao = net.action_out
        out[c + "wa1"], out[c + "ba1"] = ao[0].weight.detach().numpy(), ao[0].bias.detach().numpy()
        if cont:
            out[c + "wa2"], out[c + "ba2"] = ao[2].weight.detach().numpy(), ao[2].bias.detach().numpy()
        for name, layer in (("adv1", net.advantage1), ("adv2", net.advantage2)):
            out[c + "w" + name] = (layer.weight_mu + layer.weight_sigma * layer.weight_epsilon).detach().numpy()
            out[c + "b" + name] = (layer.bias_mu + layer.bias_sigma * layer.bias_epsilon).detach().numpy()
    np.savez_compressed(os.path.join(HERE, "aql.npz"), **out)


def main():
    if "--only-aql" in sys.argv:
        gen_aql()
        return
    if "--only-td-aql" in sys.argv:
        gen_td_aql(load_reference("utils"))
        return
    mem = load_reference("memory")
    mem.np = numpy1_shim()
    rnd = Replay()
    mem.random = types.SimpleNamespace(random=rnd.random, randint=random.randint)
    gen_tree_ops(mem)
    gen_per(mem, rnd, "per", "per_small.npz", size=100, batch=16, rounds=6, n_fill=90, seed=1)
    gen_per(mem, rnd, "apex", "per_apex.npz", size=96, batch=32, rounds=6, n_fill=96, seed=2)
    gen_per(mem, rnd, "aql", "per_aql.npz", size=50, batch=8, rounds=4, n_fill=40, seed=3)
    gen_c1(mem, rnd)
    gen_nstep(mem)
    utl = load_reference("utils")
    gen_td(utl)
    gen_td_aql(utl)
    gen_aql()
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    sys.exit(main())
