"""Stand-in for `tensorboardX` (not installed here): a SummaryWriter that keeps its scalars in a list.
TEST INFRASTRUCTURE ONLY."""


class SummaryWriter(object):
    def __init__(self, *args, **kwargs):
        self.scalars = []

    def add_scalar(self, tag, value, step=None):
        self.scalars.append((tag, float(value), step))

    def close(self):
        pass
