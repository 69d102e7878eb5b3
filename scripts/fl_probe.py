"""FrozenLake step probe (GPU): back-to-back launches at n envs, for CUDA-event timing and as an ncu target."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from ragen_b200.env.frozen_lake import FrozenLakeBatch  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 200
b = FrozenLakeBatch(n, 4, 0.8, True, device=0)
b.reset(np.arange(n), check=False)
g = torch.Generator(device="cuda").manual_seed(2)
acts = torch.randint(1, 5, (4, n, 1), dtype=torch.uint8, device="cuda", generator=g)
for i in range(3):
    b.step(acts[i % 4])
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for i in range(reps):
    b.step(acts[i % 4])
e.record()
torch.cuda.synchronize()
us = 1e3 * s.elapsed_time(e) / reps
print(f"frozenlake 4x4 n={n}: {us:.1f} us/launch, {n / us * 1e6:.3e} env-steps/s, {n * 109 / us * 1e-3:.0f} GB/s on 109 B/env-step")
