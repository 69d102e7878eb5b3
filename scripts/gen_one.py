"""One generator launch of N rooms 10x10 / 2 boxes / depth 300 (for ncu)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ragen_b200.env.sokoban import SokobanBatch  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
b = SokobanBatch(1, (10, 10), 2, 100, 300, device=0)
b.generate(np.arange(n))  # workspace
torch.cuda.synchronize()
f, s, st, ex = b.generate(np.arange(n))
torch.cuda.synchronize()
print(n, float((st == 0).float().mean()), float(ex.float().mean()))
