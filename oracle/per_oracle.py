"""CPU oracle for the prioritized-replay hot path.  TEST INFRASTRUCTURE ONLY.

This module restates, on the CPU, the algorithm of the reference's `memory.py`
(segment trees, prioritized replay, n-step assembly) and the priority arithmetic
of `utils.compute_loss`.  It exists to CHECK the CUDA path and to be timed as the
CPU baseline; nothing in the product package imports it.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import it.

Parity pinning: the reference ships no tests or golden vectors, so
this oracle is pinned by differential runs against the reference itself:
`tests/golden/make_golden.py` imports `/root/reference/memory.py` +
`/root/reference/utils.py`, drives them with seeded inputs and replayed uniform
draws, and commits the outputs under `tests/golden/`; `tests/test_oracle_golden.py`
replays the same inputs through this file and demands bit-equality for tree
nodes / indices and 1e-5 relative for float results.  When `/root/reference` is
present (the build container), `tests/test_oracle_vs_reference.py` additionally
runs randomised op sequences (add / sample / update interleavings, duplicate indices,
ring wrap-around) through both and compares every tree node and every sampled index.

Two forms of every tree routine are given:
  * scalar, list-backed classes (`HeapTree`, `ReplayOracle`) — same data
    structures and per-element work as the reference (Python lists of Python
    floats), used for small cases and as the timed CPU baseline;
  * numpy-vectorised functions (`build_tree`, `descend_many`, ...) for sizes
    where pure-Python loops would take minutes.  They are valid because every
    internal node of the reference tree always equals op(left child, right
    child) of the *current* children (memory.py:82-87 recomputes every
    ancestor after each leaf write), so the node array is a pure function of
    the leaves, and because float64 elementwise numpy arithmetic is the same
    IEEE operation as Python float arithmetic.

All arithmetic on tree values is IEEE float64, round-to-nearest, no fused
multiply-add (CPython evaluates `a*b + c*d` as two products and one sum).
"""
import math
from collections import deque

import numpy as np

INF = float("inf")


def next_pow2(n):
    """Tree capacity rule of memory.py:227-229 (smallest power of two >= size)."""
    c = 1
    while c < n:
        c *= 2
    return c


# --------------------------------------------------------------------------------------
# scalar, list-backed tree  (memory.py:10-143)
# --------------------------------------------------------------------------------------
class HeapTree:
    """Array-backed perfect binary tree: node 1 is the root, node i has children
    2i and 2i+1, leaf k lives at `cap + k` (memory.py:36,79).  `kind` is "sum"
    (neutral 0.0, memory.py:96-100) or "min" (neutral +inf, memory.py:134-138)."""

    def __init__(self, cap, kind):
        assert cap > 0 and (cap & (cap - 1)) == 0, "capacity must be positive and a power of 2."  # memory.py:34
        self.cap = cap
        self.kind = kind
        if kind == "sum":
            self.neutral = 0.0
            self.op = lambda a, b: a + b
        elif kind == "min":
            self.neutral = INF
            self.op = min
        else:
            raise ValueError(kind)
        self.v = [self.neutral] * (2 * cap)

    def set(self, k, val):
        """Point write + recompute every ancestor as left (op) right (memory.py:76-87)."""
        node = int(k) + self.cap
        v = self.v
        v[node] = val
        node >>= 1
        while node:
            v[node] = self.op(v[node + node], v[node + node + 1])
            node >>= 1

    def get(self, k):
        assert 0 <= k < self.cap  # memory.py:90
        return self.v[self.cap + k]

    def _fold(self, lo, hi, node, node_lo, width):
        """op over leaves [lo, hi] (inclusive) inside `node`, which spans `width`
        leaves starting at `node_lo`.  Same split rule, hence same association,
        as memory.py:39-52: an exactly covered node is taken whole; a range that
        straddles the midpoint is op(left part, right part)."""
        if lo == node_lo and hi == node_lo + width - 1:
            return self.v[node]
        half = width >> 1
        right_lo = node_lo + half
        if hi < right_lo:
            return self._fold(lo, hi, 2 * node, node_lo, half)
        if lo >= right_lo:
            return self._fold(lo, hi, 2 * node + 1, right_lo, half)
        a = self._fold(lo, right_lo - 1, 2 * node, node_lo, half)
        b = self._fold(right_lo, hi, 2 * node + 1, right_lo, half)
        return self.op(a, b)

    def reduce(self, start=0, end=None):
        """Half-open [start, end) like memory.py:54-74 (end None -> capacity,
        negative end wraps).  `end == start` would recurse for ever in the
        reference; here it raises instead."""
        if end is None:
            end = self.cap
        if end < 0:
            end += self.cap
        last = end - 1
        if last < start:
            raise RecursionError("empty range: the reference recurses without bound here (memory.py:73-74)")
        return self._fold(start, last, 1, 0, self.cap)

    def descend(self, mass):
        """Prefix-sum descent (memory.py:106-129): go left iff left child > mass
        (strict), otherwise subtract the left child and go right."""
        assert 0 <= mass <= self.reduce() + 1e-5  # memory.py:121
        v = self.v
        node = 1
        while node < self.cap:
            left = v[2 * node]
            if left > mass:
                node = 2 * node
            else:
                mass -= left
                node = 2 * node + 1
        return node - self.cap


# --------------------------------------------------------------------------------------
# scalar replay oracle (memory.py:146-391)
# --------------------------------------------------------------------------------------
class ReplayOracle:
    """PrioritizedReplayBuffer and its two variants in one class.

    variant: "per"   -> PrioritizedReplayBuffer            (memory.py:208-320)
             "apex"  -> CustomPrioritizedReplayBuffer      (memory.py:323-362), add takes a priority
             "aql"   -> CustomPrioritizedReplayBuffer_AQL  (memory.py:364-391), rows carry a_mu
    Rows are kept by reference in a Python list exactly like memory.py:155,166-168.
    Uniform draws are passed in explicitly (the reference pulls them from the
    global `random` module, memory.py:247)."""

    def __init__(self, size, alpha, variant="per"):
        assert alpha >= 0  # memory.py:224
        self.maxsize = size
        self.alpha = alpha
        self.variant = variant
        self.rows = []
        self.cursor = 0
        cap = next_pow2(size)
        self.sum_tree = HeapTree(cap, "sum")
        self.min_tree = HeapTree(cap, "min")
        self.max_priority = 1.0  # memory.py:233

    def __len__(self):
        return len(self.rows)

    def _store(self, row):
        k = self.cursor
        if k >= len(self.rows):
            self.rows.append(row)
        else:
            self.rows[k] = row
        nxt = (k + 1) % self.maxsize
        # memory.py:342 leaves the cursor un-cast (float size -> float cursor);
        # memory.py:376 (AQL) and :169 with an int size give ints.
        self.cursor = int(nxt) if self.variant == "aql" else nxt
        return k

    def add(self, obs, action, reward, next_obs, done, extra=None):
        """variant per: extra unused, leaf = max_priority**alpha (memory.py:235-240)
        variant apex: extra = priority (memory.py:334-346)
        variant aql:  extra = a_mu, leaf = max_priority**alpha (memory.py:368-378)"""
        if self.variant == "apex":
            prio = extra
            k = self._store((obs, action, reward, next_obs, done))
            leaf = prio ** self.alpha
            self.sum_tree.set(k, leaf)
            self.min_tree.set(k, leaf)
            self.max_priority = max(self.max_priority, prio)
        else:
            row = (obs, action, reward, next_obs, done, extra) if self.variant == "aql" \
                else (obs, action, reward, next_obs, done)
            k = self._store(row)
            leaf = self.max_priority ** self.alpha
            self.sum_tree.set(k, leaf)
            self.min_tree.set(k, leaf)
        return k

    def mass_total(self):
        """p_total of memory.py:244: sum over leaves [0, len-2] — the newest slot is
        left out and the range sum keeps memory.py:39-52's association (Q2)."""
        return self.sum_tree.reduce(0, len(self.rows) - 1)

    def sample_indices(self, batch, uniforms):
        """memory.py:242-250 with the i-th `random.random()` replaced by uniforms[i]."""
        total = self.mass_total()
        seg = total / batch
        out = []
        for i in range(batch):
            mass = uniforms[i] * seg + i * seg  # two products, one sum
            out.append(self.sum_tree.descend(mass))
        return out

    def weights_for(self, idxes, beta):
        """memory.py:288-296: full-root sum / min, N = len."""
        assert beta > 0  # memory.py:284
        n = len(self.rows)
        root_sum = self.sum_tree.reduce()
        p_min = self.min_tree.reduce() / root_sum
        max_w = (p_min * n) ** (-beta)
        w = []
        for k in idxes:
            p = self.sum_tree.get(k) / root_sum
            w.append(((p * n) ** (-beta)) / max_w)
        return np.array(w)

    def gather(self, idxes):
        """memory.py:171-182 (and :380-391 for aql): stack the sampled rows.  With
        `lists_like_reference` set on an "apex" oracle the observations come back the way
        CustomPrioritizedReplayBuffer._encode_sample returns them (memory.py:348-362): two Python
        lists of references to the stored frames, no copy — that is the form bench.py times as the
        CPU baseline, since it is the class ApeX.py uses."""
        if self.variant == "apex" and getattr(self, "lists_like_reference", False):
            obs, act, rew, nxt, done = [], [], [], [], []
            for k in idxes:
                o, a, r, n, d = self.rows[k]
                obs.append(o)
                act.append(np.asarray(a))
                rew.append(np.asarray(r))
                nxt.append(n)
                done.append(np.asarray(d))
            return (obs, np.array(act), np.array(rew), nxt, np.array(done))
        cols = list(zip(*[self.rows[k] for k in idxes]))
        return tuple(np.array([np.asarray(x) for x in col]) for col in cols)

    def sample(self, batch, beta, uniforms):
        idxes = self.sample_indices(batch, uniforms)
        w = self.weights_for(idxes, beta)
        return self.gather(idxes) + (w, idxes)

    def update_priorities(self, idxes, priorities):
        """memory.py:300-320: strictly sequential (duplicates: last wins);
        the running max never decays."""
        assert len(idxes) == len(priorities)
        for k, p in zip(idxes, priorities):
            assert p > 0
            assert 0 <= k < len(self.rows)
            leaf = p ** self.alpha
            self.sum_tree.set(k, leaf)
            self.min_tree.set(k, leaf)
            self.max_priority = max(self.max_priority, p)


# --------------------------------------------------------------------------------------
# numpy-vectorised forms for large sizes
# --------------------------------------------------------------------------------------
def build_tree(leaves, kind):
    """Node array (length 2*cap, node 0 = neutral like the reference's unused slot)
    from a float64 leaf vector whose length is a power of two."""
    leaves = np.asarray(leaves, dtype=np.float64)
    cap = leaves.shape[0]
    assert cap > 0 and (cap & (cap - 1)) == 0
    v = np.empty(2 * cap, dtype=np.float64)
    v[0] = 0.0 if kind == "sum" else np.inf
    v[cap:] = leaves
    width = cap
    while width > 1:
        half = width >> 1
        kids = v[width:2 * width]
        if kind == "sum":
            v[half:width] = kids[0::2] + kids[1::2]
        else:
            v[half:width] = np.minimum(kids[0::2], kids[1::2])
        width = half
    return v


def prefix_total(v, n_leaves):
    """sum over leaves [0, n_leaves) of a sum-tree node array `v`, with the
    association of memory.py:39-52 for start=0: walking down from the root, each
    time the range covers the left child entirely and spills into the right one
    the result is  left + (rest) ; an exactly covered node ends the chain.
    Hence t1 + (t2 + (... + tk)) — evaluated innermost first."""
    cap = v.shape[0] // 2
    assert 1 <= n_leaves <= cap
    terms = []
    node, width, need = 1, cap, n_leaves
    while True:
        if need == width:
            terms.append(float(v[node]))
            break
        half = width >> 1
        if need <= half:
            node, width = 2 * node, half
        else:
            terms.append(float(v[2 * node]))
            node, width, need = 2 * node + 1, half, need - half
    acc = terms[-1]
    for t in reversed(terms[:-1]):
        acc = t + acc
    return acc


def stratified_masses(total, batch, uniforms):
    """memory.py:245-247 for all i at once (float64, unfused)."""
    u = np.asarray(uniforms, dtype=np.float64)
    seg = np.float64(total) / np.float64(batch)
    i = np.arange(batch, dtype=np.float64)
    return u * seg + i * seg


def descend_many(v, masses):
    """memory.py:122-129 for a vector of masses."""
    cap = v.shape[0] // 2
    m = np.array(masses, dtype=np.float64, copy=True)
    node = np.ones(m.shape[0], dtype=np.int64)
    width = cap
    while width > 1:
        left = v[2 * node]
        go_left = left > m
        m = np.where(go_left, m, m - left)
        node = np.where(go_left, 2 * node, 2 * node + 1)
        width >>= 1
    return node - cap


def is_weights(sum_v, min_v, idxes, n_rows, beta):
    """memory.py:289-295 vectorised (float64 out)."""
    cap = sum_v.shape[0] // 2
    root_sum = sum_v[1]
    p_min = min_v[1] / root_sum
    max_w = (p_min * n_rows) ** (-beta)
    p = sum_v[cap + np.asarray(idxes, dtype=np.int64)] / root_sum
    return ((p * n_rows) ** (-beta)) / max_w


def apply_updates(leaves, idxes, leaf_values):
    """Sequential last-write-wins leaf scatter (memory.py:314-318) on a leaf vector."""
    out = np.array(leaves, dtype=np.float64, copy=True)
    for k, val in zip(idxes, leaf_values):
        out[int(k)] = val
    return out


def pow_alpha(priorities, alpha):
    """`priority ** alpha` with the C library's double pow, one element at a time —
    what CPython's float.__pow__ calls (memory.py:317)."""
    return np.array([math.pow(float(p), alpha) for p in priorities], dtype=np.float64)


# --------------------------------------------------------------------------------------
# learner-side priority arithmetic  (utils.py:64-81 after the three forwards)
# --------------------------------------------------------------------------------------
def td_priorities(q, q_next, q_tgt, actions, rewards, dones, weights, n_steps, gamma):
    """float32 restatement of utils.py:71-80.  Inputs: q, q_next (online net on s and
    s'), q_tgt (target net on s') as float32 [B, nA]; actions int64 [B]; rewards,
    dones, weights float32 [B].  Returns (loss, prios[B] float32, td[B] float32).
    Python-float scalars multiply float32 arrays in float32, as in torch."""
    f = np.float32
    q = np.asarray(q, f)
    q_next = np.asarray(q_next, f)
    q_tgt = np.asarray(q_tgt, f)
    rows = np.arange(q.shape[0])
    q_a = q[rows, np.asarray(actions, np.int64)]
    a_star = q_next.argmax(axis=1)                                   # utils.py:72, first max wins
    boot = q_tgt[rows, a_star]                                       # utils.py:73
    disc = f(gamma ** n_steps)
    y = np.asarray(rewards, f) + (disc * boot) * (f(1) - np.asarray(dones, f))  # utils.py:74
    td = np.abs(y - q_a)                                             # utils.py:76
    prios = (f(0.9) * td.max() + f(0.1) * td) + f(1e-6)              # utils.py:77  
    hub = np.where(td < f(1), f(0.5) * td * td, td - f(0.5))         # utils.py:79
    loss = (hub * np.asarray(weights, f)).mean(dtype=f)              # utils.py:80
    return loss, prios.astype(f), td.astype(f)


# --------------------------------------------------------------------------------------
# actor-side n-step assembly  (memory.py:393-478)
# --------------------------------------------------------------------------------------
class NStepOracle:
    """BatchStorage restated.  Keeps every quirk of SURVEY Q10/Q11: the return has
    n+1 terms when the window is full, one emit on `done`, the q-value window is
    NOT cleared on done (memory.py:432-435), plain max-Q bootstrap."""

    def __init__(self, n_steps, gamma=0.99):
        self.n = n_steps
        self.gamma = gamma
        self.win_s = deque(maxlen=n_steps)
        self.win_a = deque(maxlen=n_steps)
        self.win_r = deque(maxlen=n_steps)
        self.win_q = deque(maxlen=n_steps)
        self.reset()

    def reset(self):
        self.states, self.actions, self.rewards = [], [], []
        self.next_states, self.dones, self.q0, self.qn = [], [], [], []

    def discounted(self, *rs):
        """memory.py:471-475: ret += r_k * gamma**k, k ascending, starting from 0.0."""
        acc = 0.0
        for k, r in enumerate(rs):
            acc += r * (self.gamma ** k)
        return acc

    def add(self, state, reward, action, done, q_values):
        if len(self.win_s) == self.n or done:
            self.states.append(self.win_s[0])          # IndexError on an empty window, as in the reference
            self.rewards.append(self.discounted(*self.win_r, reward))
            self.actions.append(self.win_a[0])
            self.q0.append(self.win_q[0])
            self.next_states.append(state)
            self.qn.append(q_values)
            done = np.float32(done)
            self.dones.append(done)
        if done:
            self.win_s.clear()
            self.win_r.clear()
            self.win_a.clear()
        else:
            self.win_s.append(state)
            self.win_r.append(reward)
            self.win_a.append(action)
            self.win_q.append(q_values)

    def __len__(self):
        return len(self.states)

    def priorities(self):
        """memory.py:451-464."""
        a = np.asarray(self.actions)
        r = np.asarray(self.rewards)
        d = np.asarray(self.dones)
        q0 = np.stack(self.q0)
        qn = np.stack(self.qn)
        q_a = q0[(range(len(q0)), a)]
        boot = qn.max(1)
        y = r + (self.gamma ** self.n) * boot * (1 - d)
        return np.abs(y - q_a) + 1e-6

    def make_batch(self):
        return [self.states, self.actions, self.rewards, self.next_states, self.dones], self.priorities()
