"""The sharded replay on more than one REAL GPU: one process per GPU, the record
exchange through peer-memory mailboxes inside the sampling kernel and the early publish from the update
kernel's tail, checked against the single-tree oracle by tests/multi_gpu_parity.py.  Skipped on a box
with one GPU (tests/test_gpu_sharded.py covers the same kernels with all ranks emulated on one device);
bench.py --gpus N runs the same check at start-up."""
import os
import socket
import subprocess
import sys

import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least two GPUs")
def test_sharded_replay_on_real_peers_matches_the_single_tree_oracle(tmp_path):
    import time
    n = torch.cuda.device_count()
    world = 1 << (n.bit_length() - 1)  # largest power of two
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(free_port()), WORLD_SIZE=str(world))
    logs = [open(tmp_path / ("rank%d.log" % r), "w+") for r in range(world)]
    procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "multi_gpu_parity.py")],
                              env=dict(env, RANK=str(r), LOCAL_RANK=str(r)), stdout=logs[r], stderr=subprocess.STDOUT, text=True)
             for r in range(world)]
    # a rank that fails leaves its peers waiting in a collective: watch all of them and stop the lot at the first failure
    deadline = time.time() + 420
    while time.time() < deadline and any(p.poll() is None for p in procs) and not any(p.poll() not in (None, 0) for p in procs):
        time.sleep(0.5)
    for p in procs:
        if p.poll() is None:
            p.kill()
        p.wait()
    outs = []
    for f in logs:
        f.seek(0)
        outs.append(f.read())
        f.close()
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, "rank %d failed:\n%s" % (r, o[-3000:])
        assert "rank %d ok" % r in o and o.count("multi-GPU parity ok") == 3 and o.count("P2P delivery ok") == 1
