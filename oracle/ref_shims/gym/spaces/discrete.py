class Discrete:
    """Discrete(n, start): integers start..start+n-1 (only n/start/contains are used)."""

    def __init__(self, n, start=0):
        self.n = int(n)
        self.start = int(start)

    def contains(self, x):
        return isinstance(x, (int,)) and self.start <= x < self.start + self.n

    __contains__ = contains
