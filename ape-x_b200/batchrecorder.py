"""Drop-in for the reference's `batchrecorder.py` (same `Worker` / `BatchRecorder` signatures and
methods: record_batch(), set_worker_weights(pi), cleanup()), feeding the GPU replay.

Differences a maintainer should know: the parent ingests each actor chunk (<= 50 n-step
transitions + their actor priorities, memory.py:466-469) with ONE `buffer.add_chunk` call when
the buffer offers it (ape-x_b200/memory.py does) instead of a Python `buffer.add` per
transition (batchrecorder.py:133-137); slot order and priorities are identical.  `gym` and
`model.DuelingDQN` are imported from the caller's environment exactly as the reference does.
"""
import multiprocessing as mp

import numpy as np
import torch

from _actor_pool import ActorPool, ActorProcess, exploration_epsilon
from memory import BatchStorage


class Worker(ActorProcess):
    """One actor: an environment, a CPU DuelingDQN, a BatchStorage (batchrecorder.py:13-98)."""

    def __init__(self, worker_id, env_id, seed, epsilon, size, lock, n_steps, gamma, send_interval, task_queue, buffer,
                 max_episode_length):
        ActorProcess.__init__(self, seed, task_queue, buffer)
        import gym
        from model import DuelingDQN
        self.worker_id = worker_id
        self.env = gym.make(env_id)
        self.max_episode_length = max_episode_length
        self.send_interval = send_interval
        self.storage = BatchStorage(n_steps, gamma)
        self.size = size
        self.lock = lock
        self.memory = []
        self.model = DuelingDQN(self.env)
        self.set_all_seeds()
        self.epsilon = epsilon

    def load_weights(self, state_dict):
        self.model.load_state_dict(state_dict)

    update_weights = load_weights

    def rollout(self):
        """One episode; chunks of n-step transitions with their priorities, cut every
        `send_interval` emissions and at the episode end (batchrecorder.py:42-78)."""
        state = self.env.reset()
        self.storage.reset()
        self.memory = []
        steps = 0
        while True:
            action, q_values = self.model.act(torch.FloatTensor(np.array(state)), self.epsilon)
            next_state, reward, done, _ = self.env.step(action)
            self.storage.add(state, reward, action, done, q_values)
            state = next_state
            steps += 1
            truncated = steps >= self.max_episode_length
            if done or truncated:
                state = self.env.reset()
                steps = 0
            if (done or len(self.storage) == self.send_interval) and len(self.storage):
                batch, prios = self.storage.make_batch()
                self.memory.append((*batch, prios))
                self.storage.reset()
            if done:
                return self.memory

    record_batch = rollout


class BatchRecorder(object):
    """batchrecorder.py:100-152."""

    def __init__(self, env_id, env_seed, n_workers, buffer, n_steps, gamma, max_episode_length):
        self.env_id, self.env_seed, self.n_workers, self.buffer = env_id, env_seed, n_workers, buffer
        self.size = 1e5
        self.task_queue = mp.JoinableQueue()
        self.res_queue = mp.Queue()
        lock = mp.Lock()
        self.workers = [Worker(worker_id=i, env_id=env_id, seed=env_seed + i, epsilon=exploration_epsilon(i, n_workers),
                               size=self.size // n_workers, lock=lock, n_steps=n_steps, gamma=gamma, send_interval=50,
                               task_queue=self.task_queue, buffer=self.res_queue, max_episode_length=max_episode_length)
                        for i in range(n_workers)]
        self._pool = ActorPool(self.workers, self.task_queue, self.res_queue)

    def record_batch(self):
        self._pool.broadcast({"desc": "record_batch"})
        for mem in self._pool.results():
            for states, actions, rewards, next_states, dones, prios in mem:
                ingest_chunk(self.buffer, states, actions, rewards, next_states, dones, prios)

    def set_worker_weights(self, pi):
        self._pool.set_weights(pi)

    def cleanup(self):
        self._pool.shutdown()


def ingest_chunk(buffer, states, actions, rewards, next_states, dones, prios):
    """buffer.add(s, a, R, s', d, prio) for every row of one actor chunk, in order."""
    if hasattr(buffer, "add_chunk"):
        buffer.add_chunk(states, actions, rewards, next_states, dones, prios)
    else:
        for i in range(len(states)):
            buffer.add(states[i], actions[i], rewards[i], next_states[i], dones[i], prios[i])
