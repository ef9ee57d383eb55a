"""CPU restatement of the sharded-replay protocol.  TEST INFRASTRUCTURE ONLY.

The reference has a single tree (memory.py:231-232); the claim checked with this file is that
`world` contiguous slices of that tree plus a 4-double record per slice reproduce
`_sample_proportional` (memory.py:242-250) bit for bit.  The functions take the node arrays
of the local trees (numpy, heap layout) and mirror, step for step, what
per_shard_record / per_sample_sharded do on the GPU; tests compare both with the single-tree
oracle (oracle/per_oracle.py).
"""
import numpy as np

from oracle import per_oracle as po


def shard_record(sum_v, min_v, rank, global_len):
    """{root sum, root min, local part of the sum(0, len-1) chain, unused} for one slice."""
    cap = sum_v.shape[0] // 2
    need_local = min(max((global_len - 1) - rank * cap, 0), cap)
    partial = po.prefix_total(sum_v, need_local) if 0 < need_local < cap else 0.0
    return np.array([sum_v[1], min_v[1], partial, 0.0], dtype=np.float64)


def top_tree(records):
    """Nodes 1..world-1 above the shard roots, pairwise like memory.py:82-87; index world+r = root r."""
    world = records.shape[0]
    top = np.zeros(2 * world, dtype=np.float64)
    top[world:] = records[:, 0]
    for i in range(world - 1, 0, -1):
        top[i] = top[2 * i] + top[2 * i + 1]
    return top


def global_mass_total(records, cap_local, global_len):
    """The reference's p_total = sum(0, len-1) (memory.py:244) from the records alone."""
    world = records.shape[0]
    top = top_tree(records)
    need = global_len - 1
    need_hi, need_lo = need // cap_local, need % cap_local
    acc, first = 0.0, True
    if need_lo:
        acc, first = float(records[need_hi, 2]), False
    b = 0
    while (1 << b) <= need_hi:
        if (need_hi >> b) & 1:
            t = float(top[(world >> b) + ((need_hi >> (b + 1)) << 1)])
            acc = t if first else t + acc
            first = False
        b += 1
    return acc


def sharded_sample(records, sum_v_local, rank, uniforms, batch, global_len):
    """Indices this rank produces: global index for its own strata, -1 elsewhere."""
    records = np.asarray(records, dtype=np.float64).reshape(-1, 4)
    world = records.shape[0]
    cap_local = sum_v_local.shape[0] // 2
    top = top_tree(records)
    total = global_mass_total(records, cap_local, global_len)
    masses = po.stratified_masses(total, batch, uniforms)
    out = np.full(batch, -1, dtype=np.int64)
    mine, resid = [], []
    for i, m in enumerate(masses):
        node = 1
        m = float(m)
        while node < world:
            left = float(top[2 * node])
            if left > m:
                node = 2 * node
            else:
                m -= left
                node = 2 * node + 1
        if node - world == rank:
            mine.append(i)
            resid.append(m)
    if mine:
        out[np.array(mine)] = rank * cap_local + po.descend_many(sum_v_local, np.array(resid))
    return out


def slice_of(global_v, rank, world):
    """Node array of slice `rank` cut out of a single tree's node array (for cross-checks):
    local node j at depth d, position p  <->  global node ((world + rank) << d) + p."""
    cap = global_v.shape[0] // 2
    cap_l = cap // world
    out = np.empty(2 * cap_l, dtype=np.float64)
    out[0] = global_v[0]
    d = 0
    while (1 << d) <= cap_l:
        lo = (world + rank) << d
        out[(1 << d):(2 << d)] = global_v[lo:lo + (1 << d)]
        d += 1
    return out
