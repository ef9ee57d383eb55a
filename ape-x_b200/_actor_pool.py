"""Process pool shared by batchrecorder.py and batchrecoder_AQL.py: N actor processes, each
owning an environment and a CPU copy of the policy, driven by three commands on a joinable
queue — the wire protocol of the reference pools (batchrecorder.py:79-98,128-152):
  {"desc": "record_batch"}                       run one episode, put the result on the result queue
  {"desc": "set_pi_weights", "pi_state_dict": …} load new policy weights
  {"desc": "cleanup"}                            close the environment
The environment loop itself is not part of the accelerated path; what
matters here is that its output reaches the GPU replay in batches instead of one Python call
per transition (batchrecorder.py:133-137).
"""
import multiprocessing as mp
import random

import numpy as np
import torch


def exploration_epsilon(i, n_workers):
    """0.4 ** (1 + 7 i / (N - 1))  (batchrecorder.py:121, origin_repo/actor.py:69).  The reference
    divides by zero for a single worker; a lone worker gets the i = 0 value here."""
    return 0.4 ** (1 + (i / (n_workers - 1) * 7 if n_workers > 1 else 0.0))


class ActorProcess(mp.Process):
    """Subclasses provide: make_env_and_model() called in __init__, rollout() -> result,
    load_weights(state_dict)."""

    def __init__(self, seed, task_queue, result_queue):
        mp.Process.__init__(self)
        self.seed = seed
        self.task_queue = task_queue
        self.buffer = result_queue  # the reference calls the result queue `buffer` (batchrecorder.py:20)

    def set_all_seeds(self):
        self.env.seed(self.seed)
        np.random.seed(self.seed)
        random.seed(self.seed)
        torch.manual_seed(self.seed)
        if torch.cuda.is_available():
            torch.cuda.manual_seed_all(self.seed)

    def run(self):
        while True:
            task = self.task_queue.get(block=True)
            kind = task["desc"]
            if kind == "record_batch":
                self.buffer.put(self.rollout())
            elif kind == "set_pi_weights":
                self.load_weights(task["pi_state_dict"])
            elif kind == "cleanup":
                self.env.close()
            self.task_queue.task_done()


class ActorPool(object):
    def __init__(self, workers, task_queue, result_queue):
        self.workers, self.task_queue, self.res_queue = workers, task_queue, result_queue
        for w in self.workers:
            w.start()

    def broadcast(self, task, wait=True):
        for _ in self.workers:
            self.task_queue.put(task)
        if wait:
            self.task_queue.join()

    def results(self):
        return [self.res_queue.get() for _ in self.workers]

    def set_weights(self, pi):
        pi.to("cpu")
        self.broadcast({"desc": "set_pi_weights", "pi_state_dict": pi.state_dict()})

    def shutdown(self):
        self.broadcast({"desc": "cleanup"}, wait=False)
        for w in self.workers:
            w.terminate()
