"""GRPO at many rows (GPU): back-to-back calls for CUDA-event timing and as an ncu target."""
import sys

import torch

sys.path.insert(0, ".")
from ragen_b200 import ops  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
T = int(sys.argv[2]) if len(sys.argv) > 2 else 2047
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
g = torch.Generator(device="cuda").manual_seed(0)
mask = (torch.rand(B, T, device="cuda", generator=g) < 0.5).to(torch.uint8)
rewards = torch.randn(B, T, device="cuda", generator=g)
gi = torch.arange(B, device="cuda", dtype=torch.int32) // 16
for _ in range(3):
    ops.grpo_outcome(rewards, mask, gi, B // 16)
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(reps):
    ops.grpo_outcome(rewards, mask, gi, B // 16)
e.record()
torch.cuda.synchronize()
print(f"grpo B={B} T={T}: {1e3 * s.elapsed_time(e) / reps:.1f} us per call")
