"""seg_set_noalloc — the batched tree write for callers that are not allowed to allocate
device memory (a process whose framework already owns the whole GPU through its own pool, or
a loop that cannot take the device-wide synchronisation of cudaMalloc/cudaFree).  It uses a
static 8 MiB stamp array inside the library instead of the per-handle allocation of
seg_create, and works on node arrays the caller owns.

Collected last on purpose (file name): a run that stops at the first failure has then
already run every other test.  The check: same batches through the existing entry point
(seg_create + seg_set) and through the new one, called through the C ABI, two calls in a row
on the same arrays (the second call works in the array the first one used); the node
arrays must be identical bit for bit — the tolerance the other tree tests use.  Further cases: trees of
different sizes taking turns on the shared array, a call without a status word, and the same call replayed
from a CUDA graph over changing, duplicated indices (the shared array must be back to all-zero after every call).
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def native():
    import __graft_entry__ as entry
    entry.build()
    import per_native
    return per_native


def bits(t):
    return t.cpu().numpy().view(np.uint64)


def reference_nodes(pn, cap, kind, batches, dev):
    """Existing path: seg_create -> seg_set (one call per batch) -> seg_read."""
    h = C.c_void_p()
    pn.check(pn.lib.seg_create(C.byref(h), dev.index, cap, kind))
    try:
        for idx, val in batches:
            pn.check(pn.lib.seg_set(h, pn.ptr(idx), pn.ptr(val), idx.numel(), pn.stream_ptr()))
        out = torch.empty(2 * cap, dtype=torch.float64, device=dev)
        pn.check(pn.lib.seg_read(h, pn.ptr(out), pn.stream_ptr()))
        torch.cuda.synchronize()
        return out
    finally:
        pn.lib.seg_destroy(h)


@pytest.mark.parametrize("cap,n", [(1 << 12, 300), (1 << 16, 2048), (1 << 20, 512), (1 << 20, 50000)])
def test_noalloc_matches_seg_set_twice_in_a_row(native, cap, n):
    pn = native
    dev = torch.device("cuda", torch.cuda.current_device())
    gen = torch.Generator(device=dev).manual_seed(cap + n)
    batches = []
    for _ in range(2):  # two calls in a row
        idx = torch.randint(0, cap, (n,), dtype=torch.int64, device=dev, generator=gen)
        idx[: n // 8] = idx[n // 2: n // 2 + n // 8]  # duplicates: the later write must win
        val = torch.rand(n, dtype=torch.float64, device=dev, generator=gen) + 1e-9
        batches.append((idx, val))
    want_sum = reference_nodes(pn, cap, pn.TREE_SUM, batches, dev)
    want_min = reference_nodes(pn, cap, pn.TREE_MIN, batches, dev)

    # caller-owned node arrays, allocated BEFORE the calls under test
    sum_nodes = torch.zeros(2 * cap, dtype=torch.float64, device=dev)
    min_nodes = torch.full((2 * cap,), float("inf"), dtype=torch.float64, device=dev)
    status = torch.zeros(1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(dev)
    free_before = torch.cuda.mem_get_info(dev)[0]  # driver's view: catches a cudaMalloc inside the library
    for idx, val in batches:
        rc = pn.lib.seg_set_noalloc(dev.index, pn.ptr(sum_nodes), pn.ptr(min_nodes), cap, pn.ptr(idx), pn.ptr(val),
                                    idx.numel(), pn.ptr(status), pn.stream_ptr())
        pn.check(rc)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(dev) == before
    assert torch.cuda.mem_get_info(dev)[0] >= free_before
    assert int(status.item()) == 0
    assert np.array_equal(bits(sum_nodes), bits(want_sum))
    assert np.array_equal(bits(min_nodes), bits(want_min))


def test_noalloc_single_tree_and_bad_index(native):
    pn = native
    dev = torch.device("cuda", torch.cuda.current_device())
    cap = 256
    sum_nodes = torch.zeros(2 * cap, dtype=torch.float64, device=dev)
    status = torch.zeros(1, dtype=torch.int64, device=dev)
    idx = torch.tensor([3, 999, 7], dtype=torch.int64, device=dev)  # 999 is out of range: skipped and reported
    val = torch.tensor([1.5, 9.0, 2.5], dtype=torch.float64, device=dev)
    pn.check(pn.lib.seg_set_noalloc(dev.index, pn.ptr(sum_nodes), None, cap, pn.ptr(idx), pn.ptr(val), 3,
                                    pn.ptr(status), pn.stream_ptr()))
    torch.cuda.synchronize()
    assert int(status.item()) == pn.STATUS_BAD_INDEX
    assert float(sum_nodes[1]) == 4.0 and float(sum_nodes[cap + 3]) == 1.5 and float(sum_nodes[cap + 7]) == 2.5


def test_noalloc_rejects_oversize_and_bad_shapes(native):
    pn = native
    dev = torch.device("cuda", torch.cuda.current_device())
    one = torch.zeros(4, dtype=torch.float64, device=dev)
    idx = torch.zeros(1, dtype=torch.int64, device=dev)
    too_big = 2 * pn.SEG_NOALLOC_MAX_CAPACITY
    rc = pn.lib.seg_set_noalloc(dev.index, pn.ptr(one), None, too_big, pn.ptr(idx), pn.ptr(one), 1, None, pn.stream_ptr())
    assert rc == pn.PER_ERR_ARG
    msg = pn.lib.per_last_error().decode()
    assert "SEG_NOALLOC_MAX_CAPACITY" in msg and str(pn.SEG_NOALLOC_MAX_CAPACITY) in msg
    rc = pn.lib.seg_set_noalloc(dev.index, pn.ptr(one), None, 48, pn.ptr(idx), pn.ptr(one), 1, None, pn.stream_ptr())
    assert rc == pn.PER_ERR_ASSERT  # capacity must be a power of two, like memory.py:34
    rc = pn.lib.seg_set_noalloc(dev.index, None, None, 64, pn.ptr(idx), pn.ptr(one), 1, None, pn.stream_ptr())
    assert rc == pn.PER_ERR_ARG


def test_noalloc_trees_of_different_sizes_take_turns(native):
    """Two trees of different capacity share the one static array: calls alternate between them, with
    index sets that overlap in their low leaves; each tree must match its own seg_set reference."""
    pn = native
    dev = torch.device("cuda", torch.cuda.current_device())
    caps = (1 << 10, 1 << 14)
    gen = torch.Generator(device=dev).manual_seed(5)
    batches = {c: [] for c in caps}
    for r in range(3):
        for c in caps:
            n = 700
            idx = torch.randint(0, 1 << 10, (n,), dtype=torch.int64, device=dev, generator=gen)  # same low leaves in both trees
            idx[:100] = idx[300:400]
            batches[c].append((idx, torch.rand(n, dtype=torch.float64, device=dev, generator=gen) + 1e-9))
    nodes = {c: torch.zeros(2 * c, dtype=torch.float64, device=dev) for c in caps}
    for r in range(3):
        for c in caps:  # alternate: a, b, a, b, ...
            idx, val = batches[c][r]
            pn.check(pn.lib.seg_set_noalloc(dev.index, pn.ptr(nodes[c]), None, c, pn.ptr(idx), pn.ptr(val), idx.numel(),
                                            None, pn.stream_ptr()))  # no status word: the library's own fallback word is used
    torch.cuda.synchronize()
    for c in caps:
        assert np.array_equal(bits(nodes[c]), bits(reference_nodes(pn, c, pn.TREE_SUM, batches[c], dev)))


def test_noalloc_replayed_from_a_cuda_graph(native):
    """One captured call, replayed while the contents of its index / value buffers change between replays
    (duplicates included).  Every replay must behave like a fresh call: nothing a replay leaves in the
    shared array may decide the next one."""
    pn = native
    dev = torch.device("cuda", torch.cuda.current_device())
    cap, n = 1 << 12, 600
    gen = torch.Generator(device=dev).manual_seed(11)
    idx = torch.zeros(n, dtype=torch.int64, device=dev)
    val = torch.zeros(n, dtype=torch.float64, device=dev)
    nodes = torch.zeros(2 * cap, dtype=torch.float64, device=dev)
    status = torch.zeros(1, dtype=torch.int64, device=dev)

    def call(stream):
        pn.check(pn.lib.seg_set_noalloc(dev.index, pn.ptr(nodes), None, cap, pn.ptr(idx), pn.ptr(val), n, pn.ptr(status),
                                        pn.stream_ptr(stream)))

    rounds = []
    for r in range(4):
        i = torch.randint(0, cap, (n,), dtype=torch.int64, device=dev, generator=gen)
        # round r: a leaf that had many duplicates (late winner) in round r-1 now appears once, early in the batch
        i[n - 50:] = i[0] if r % 2 == 0 else i[n - 50:]
        if r % 2 == 1:
            i[3] = rounds[-1][0][0]
        rounds.append((i, torch.rand(n, dtype=torch.float64, device=dev, generator=gen) + 1e-9))
    idx.copy_(rounds[0][0])
    val.copy_(rounds[0][1])
    call(torch.cuda.current_stream())  # first call on the device happens outside the capture (address look-up)
    nodes.zero_()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    g = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            call(side)
    torch.cuda.current_stream().wait_stream(side)
    for i, v in rounds:
        idx.copy_(i)
        val.copy_(v)
        g.replay()
    torch.cuda.synchronize()
    assert int(status.item()) == 0
    assert np.array_equal(bits(nodes), bits(reference_nodes(pn, cap, pn.TREE_SUM, rounds, dev)))
