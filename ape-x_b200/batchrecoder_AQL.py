"""Drop-in for the reference's `batchrecoder_AQL.py` [sic] (same `Worker` / `BatchRecorder`
signatures; record_batch() -> total episode length, set_worker_weights(pi), cleanup()).  No n-step
assembly here: actors ship raw (s, a, r, s', d, a_mu) rows and the parent inserts each row
len(state) times, as the reference does (batchrecoder_AQL.py:121-123)."""
import multiprocessing as mp

from _actor_pool import ActorPool, ActorProcess, exploration_epsilon


class Worker(ActorProcess):
    """batchrecoder_AQL.py:13-80."""

    def __init__(self, worker_id, env_id, seed, epsilon, task_queue, buffer, max_episode_length, propose_sample=100,
                 uniform_sample=400, action_var=0.25, device="cuda"):
        ActorProcess.__init__(self, seed, task_queue, buffer)
        import gym
        from model import AQL
        self.worker_id = worker_id
        self.env = gym.make(env_id)
        self.max_episode_length = max_episode_length
        self.memory = []
        self.model = AQL(self.env, propose_sample=propose_sample, uniform_sample=uniform_sample, action_var=action_var,
                         device=device)
        self.set_all_seeds()
        self.epsilon = epsilon

    def load_weights(self, state_dict):
        self.model.load_state_dict(state_dict)

    update_weights = load_weights

    def rollout(self):
        state = self.env.reset()
        self.memory = []
        self.episode_reward, self.episode_length = 0, 0
        while True:
            action, a_mu, _ = self.model.act(state, self.epsilon)
            a_mu = a_mu[0]
            next_state, reward, done, _ = self.env.step(a_mu[action])
            self.memory.append((state, action, reward, next_state, done, a_mu))
            state = next_state
            self.episode_reward += reward
            self.episode_length += 1
            if done or self.episode_length >= self.max_episode_length:
                self.env.reset()
                return self.memory, self.episode_reward, self.episode_length

    record_batch = rollout


class BatchRecorder(object):
    """batchrecoder_AQL.py:82-138."""

    def __init__(self, env_id, env_seed, n_workers, buffer, max_episode_length, writer, propose_sample=100,
                 uniform_sample=400, action_var=0.25, device="cuda"):
        self.env_id, self.env_seed, self.n_workers = env_id, env_seed, n_workers
        self.buffer, self.writer, self.episode_idx = buffer, writer, 0
        self.task_queue = mp.JoinableQueue()
        self.res_queue = mp.Queue()
        self.workers = [Worker(worker_id=i, env_id=env_id, seed=env_seed + i, epsilon=exploration_epsilon(i, n_workers),
                               max_episode_length=max_episode_length, task_queue=self.task_queue, buffer=self.res_queue,
                               propose_sample=propose_sample, uniform_sample=uniform_sample, action_var=action_var,
                               device="cpu") for i in range(n_workers)]
        self._pool = ActorPool(self.workers, self.task_queue, self.res_queue)

    def record_batch(self):
        self._pool.broadcast({"desc": "record_batch"})
        total = 0
        for mem, ep_reward, ep_len in self._pool.results():
            total += ep_len
            self.writer.add_scalar("actor/episode_reward", ep_reward, self.episode_idx)
            self.writer.add_scalar("actor/episode_length", ep_len, self.episode_idx)
            self.episode_idx += 1
            for state, action, reward, next_state, done, a_mu in mem:
                for _ in range(len(state)):  # the reference's duplication, kept
                    self.buffer.add(state, action, reward, next_state, done, a_mu)
        return total

    def set_worker_weights(self, pi):
        self._pool.set_weights(pi)

    def cleanup(self):
        self._pool.shutdown()
