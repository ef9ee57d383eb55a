"""CPU half of the drop-in proof (needs /root/reference, i.e. the build container): the committed trace
tests/golden/dropin_trace.npz is what the UNMODIFIED reference drivers do — re-run here and compared — and every call
they make on their buffer binds to the signatures of the drop-in classes in ape-x_b200/memory.py.  The GPU half
(tests/test_gpu_dropin.py) replays the trace on the B200."""
import importlib.util
import inspect
import os

import numpy as np
import pytest

from conftest import GOLDEN

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.exists(os.path.join(REF, "ApeX.py")), reason="the reference tree is only present in the build container")


@pytest.fixture(scope="module")
def tracer():
    spec = importlib.util.spec_from_file_location("make_dropin_trace", os.path.join(GOLDEN, "make_dropin_trace.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    import torch
    torch.set_num_threads(1)
    return m


@pytest.mark.parametrize("driver", ["dqn", "aql", "apex", "aql_dis"])
def test_committed_trace_is_what_the_unmodified_driver_does(tracer, driver):
    g = np.load(os.path.join(GOLDEN, "dropin_trace.npz"), allow_pickle=False)
    events, final = tracer.run_driver(driver)
    out = {}
    tracer.pack(events, final, driver, out)
    keys = [k for k in g.files if k.startswith(driver + "_") and not (driver == "aql" and k.startswith("aql_dis_"))]
    assert sorted(keys) == sorted(out.keys())
    if driver in ("dqn", "aql"):  # one process: the run is reproducible to the bit
        for k in keys:
            assert np.array_equal(np.asarray(out[k]), g[k]), k
    else:
        # two actor processes share one task queue (batchrecorder.py:129-132): which worker picks up which task, hence
        # which episodes are played, differs from run to run in the reference itself; the shape of the session does not
        for k in ("_cls", "_size", "_alpha", "_u_prio_dtype"):
            assert str(out[driver + k]) == str(g[driver + k]), k
        assert list(out[driver + "_ev_kind"][out[driver + "_ev_kind"] > 0]) == list(g[driver + "_ev_kind"][g[driver + "_ev_kind"] > 0])
        assert out[driver + "_add_obs"].shape[1:] == g[driver + "_add_obs"].shape[1:] and out[driver + "_add_obs"].dtype == g[driver + "_add_obs"].dtype
        assert all(out[k].dtype == g[k].dtype and out[k].shape[1:] == g[k].shape[1:] for k in keys if k.startswith(driver + "_s_col"))


def test_committed_trace_is_what_the_unmodified_replay_server_does(tracer):
    """origin_repo/replay.py's worker coroutines, run unmodified over in-memory sockets (one worker of each kind, a
    scripted actor / learner).  Its sender seeds the global RNG from the clock (replay.py:128-129), so the draws — and
    with them the sampled batches and the updates that answer them — differ from run to run in the reference itself;
    the ingest and the shape of the session do not."""
    g = np.load(os.path.join(GOLDEN, "dropin_trace.npz"), allow_pickle=False)
    events, final, chunk_rows = tracer.run_replay_server()
    out = {}
    tracer.pack(events, final, "replay_server", out)
    assert chunk_rows == int(g["replay_server_chunk_rows"])
    for k, v in out.items():
        v = np.asarray(v)
        if "_add_" in k or "_ev_" in k or k.endswith(("_cls", "_size", "_alpha", "_final_len", "_final_next_idx", "_final_cap")):
            assert np.array_equal(v, g[k]), k
        else:
            assert v.shape == g[k].shape and v.dtype == g[k].dtype, k


def test_every_recorded_call_binds_to_the_dropin_signatures():
    """Same constructor / add / sample / update_priorities signatures as memory.py: the exact
    positional and keyword shapes the drivers use must bind (no GPU needed to check that)."""
    import memory as dropin
    assert dropin.__file__.endswith(os.path.join("ape-x_b200", "memory.py"))
    spec = importlib.util.spec_from_file_location("_ref_memory_sig", os.path.join(REF, "memory.py"))
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    for cls in ("SegmentTree", "SumSegmentTree", "MinSegmentTree", "ReplayBuffer", "PrioritizedReplayBuffer",
                "CustomPrioritizedReplayBuffer", "CustomPrioritizedReplayBuffer_AQL", "BatchStorage"):
        rc, dc = getattr(ref, cls), getattr(dropin, cls)
        for name, fn in inspect.getmembers(rc, predicate=inspect.isfunction):
            if name.startswith("_") and name not in ("__init__", "__len__", "__setitem__", "__getitem__", "_encode_sample", "_sample_proportional"):
                continue
            assert hasattr(dc, name), "%s.%s is missing from the drop-in" % (cls, name)
            rp = inspect.signature(fn).parameters
            dp = inspect.signature(getattr(dc, name)).parameters
            if any(p.kind in (p.VAR_POSITIONAL, p.VAR_KEYWORD) for p in rp.values()):
                continue
            # every reference parameter exists under the same name and in the same position (the drop-in may append optional ones)
            names = [n for n in rp]
            assert list(dp)[:len(names)] == names, "%s.%s%s vs drop-in %s" % (cls, name, tuple(rp), tuple(dp))
            for extra in list(dp)[len(names):]:
                assert dp[extra].default is not inspect.Parameter.empty, "%s.%s: new parameter %s must be optional" % (cls, name, extra)
    # the drivers' call shapes (ApeX.py:32,49,63; DQN.py:33,46,72,86; AQL.py:40,57,103,121; AQL_dis.py:44,64,100; batchrecorder.py:137)
    inspect.signature(dropin.CustomPrioritizedReplayBuffer.__init__).bind(None, size=1e6, alpha=0.6)
    inspect.signature(dropin.PrioritizedReplayBuffer.__init__).bind(None, 100000, alpha=0.6)
    inspect.signature(dropin.CustomPrioritizedReplayBuffer_AQL.__init__).bind(None, 1e7, alpha=0.6)
    inspect.signature(dropin.CustomPrioritizedReplayBuffer.add).bind(None, 0, 1, 2.0, 3, False, 0.5)
    inspect.signature(dropin.CustomPrioritizedReplayBuffer_AQL.add).bind(None, 0, 1, 2.0, 3, False, [0.1])
    inspect.signature(dropin.PrioritizedReplayBuffer.sample).bind(None, 32, 0.4)
    inspect.signature(dropin.PrioritizedReplayBuffer.update_priorities).bind(None, [1], [0.5])
