"""gymnasium.utils.seeding.np_random: Generator(PCG64(SeedSequence(seed)))."""
import numpy as np


def np_random(seed=None):
    if seed is not None and not (isinstance(seed, (int, np.integer)) and seed >= 0):
        raise ValueError(f"Seed must be a non-negative integer or omitted, not {seed!r}")
    seq = np.random.SeedSequence(None if seed is None else int(seed))
    return np.random.Generator(np.random.PCG64(seq)), seq.entropy
