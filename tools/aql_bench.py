"""BASELINE configs[2]: AQL candidate-Q at B=256, K=1000 — fused tcgen05 kernel vs the stock
PyTorch ops the reference runs (model.py:294-335) on the same B200.  CUDA-event timing."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "ape-x_b200"), os.path.join(ROOT, "tests")]
import torch

import __graft_entry__ as entry

entry.build()
import aql
from test_gpu_aql import QRef

torch.backends.cuda.matmul.allow_tf32 = False


def timeit(fn, iters=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


for continuous, A in ((True, 1), (True, 4), (False, 1)):
    B, K, obs = 256, 1000, 24
    ref = QRef(obs, A, continuous).cuda()
    state = torch.randn(B, obs, device="cuda")
    a_mu = torch.rand(B, K, A, device="cuda") if continuous else torch.randint(0, 18, (B, K), device="cuda")
    cq = aql.CandidateQ(continuous, A)
    cq.set_weights_from(ref)
    with torch.no_grad():
        f = ref.q_feature(state)
        t_ref = timeit(lambda: ref.act(state, a_mu, K).argmax(1))
        t_ours = timeit(lambda: cq.forward(f, a_mu))
        t_all = timeit(lambda: (cq.set_weights_from(ref), cq.forward(ref.q_feature(state), a_mu)))
        t_noarg = timeit(lambda: cq.forward(f, a_mu, want_argmax=False))
        # the same three launches replayed from a CUDA graph: GPU time without the Python / ctypes launch path
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            cq.forward(f, a_mu)
            with torch.cuda.graph(g, stream=side):
                for _ in range(10):
                    keep = cq.forward(f, a_mu)
        torch.cuda.current_stream().wait_stream(side)
        t_graph = timeit(g.replay, iters=10) / 10
        import time
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(50):
            cq.forward(f, a_mu)
        t_cpu = (time.perf_counter() - t0) / 50 * 1e6
        torch.cuda.synchronize()
    print("   fused without argmax %.1f us | graph replay %.1f us per forward | host launch path %.1f us per forward" % (t_noarg, t_graph, t_cpu))
    with torch.no_grad():
        pass
    import per_native as pn
    st = (pn.C.c_int64 * 8)()
    pn.lib.aql_trace(cq._h, st, pn.stream_ptr())
    cq.forward(f, a_mu)
    pn.lib.aql_trace(cq._h, st, pn.stream_ptr())
    tiles = max(st[5], 1)
    print("   CTA 0 cycles per tile: L1+issue %d | wait MMA2 %d | epi2+issue %d | wait MMA3 %d | epi3 %d | tiles %d | entry->first tile %d | entry->exit %d"
          % (st[0] // tiles, st[1] // tiles, st[2] // tiles, st[3] // tiles, st[4] // tiles, st[5], st[6], st[7]))
    flop = 2 * B * K * ((A * 128 + 128 * 64 if continuous else 64) + 64 * 64 + 64) + 2 * B * 64 * 64
    print("continuous=%s A=%d: pytorch fp32 %.1f us | fused %.1f us (%.1fx) | fused incl. weight prep + q_feature %.1f us | "
          "%.2f GFLOP -> %.1f TFLOP/s fp32-equivalent (x3 TF32 MMAs issued)" % (continuous, A, t_ref, t_ours, t_ref / t_ours, t_all,
                                                                             flop / 1e9, flop / t_ours / 1e6))
