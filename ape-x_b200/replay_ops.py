"""The three buffer operations of the upstream replay server with its pickle wire format
(origin_repo/replay.py:19-45), for a server that holds the GPU buffer.

origin_repo/replay.py itself runs unchanged on top of ape-x_b200/memory.py; these versions only
differ in that an actor chunk goes to the GPU as one `add_chunk` instead of <= 50 `add` calls."""
import _pickle as pickle


def push_batch(buffer, data):
    """data = pickle((batch, prios)) as sent by origin_repo/actor.py:105-109, batch =
    [states, actions, rewards, next_states, dones] (memory.py:466-469)."""
    batch, prios = pickle.loads(data)
    if hasattr(buffer, "add_chunk"):
        buffer.add_chunk(batch[0], batch[1], batch[2], batch[3], batch[4], prios)
    else:
        for sample in zip(*batch, prios):
            buffer.add(*sample)


def update_prios(buffer, data):
    """data = pickle((idxes, prios)) as sent by origin_repo/learner.py (prios sender)."""
    idxes, prios = pickle.loads(data)
    buffer.update_priorities(idxes, prios)


def sample_batch(buffer, batch_size, beta):
    """-> pickle of the 7-tuple buffer.sample returns (memory.py:298)."""
    return pickle.dumps(buffer.sample(batch_size, beta))
