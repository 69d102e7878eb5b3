"""Generator timing at a few sizes (run with RAGEN_GEN_REFILL=k to probe the refill threshold)."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ragen_b200.env.sokoban import SokobanBatch  # noqa: E402

b = SokobanBatch(1, (10, 10), 2, 100, 300, device=0)
b.generate(np.arange(64))
torch.cuda.synchronize()
for n in (16384, 65536, 262144):
    seeds = np.arange(n)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    f, s, st, ex = b.generate(seeds)
    t1.record()
    torch.cuda.synchronize()
    print(os.environ.get("RAGEN_GEN_REFILL", "default"), n, round(t0.elapsed_time(t1), 2), "ms", round(n / t0.elapsed_time(t1) * 1e3), "rooms/s",
          float((st == 0).float().mean()), flush=True)
