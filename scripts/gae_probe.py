"""GAE timing probe at many rows (GPU): back-to-back launches, for CUDA-event timing and as an ncu target."""
import sys

import torch

sys.path.insert(0, ".")
from ragen_b200 import ops  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
T = int(sys.argv[2]) if len(sys.argv) > 2 else 2047
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
g = torch.Generator(device="cuda").manual_seed(0)
mask = (torch.rand(B, T, device="cuda", generator=g) < 0.5).to(torch.uint8)
values = torch.randn(B, T, device="cuda", generator=g) * mask
rewards = torch.randn(B, T, device="cuda", generator=g)
# bi-level: a reward at the end of each of ~5 turns per row
r2 = rewards * (torch.rand(B, T, device="cuda", generator=g) < 5.0 / T) * mask
r2[:, -1] = 1.0
mask[:, -1] = 1
base = torch.randn(B, device="cuda", generator=g)
adv0, _, mom0 = ops.gae_raw(rewards, values, mask, 1.0, 1.0, exact=False)
for fn, name in [(lambda: ops.gae_raw(rewards, values, mask, 1.0, 1.0), "gae"),
                 (lambda: ops.gae_raw(rewards, values, mask, 1.0, 1.0, exact=False), "gae-scan"),
                 (lambda: ops.whiten_(adv0, mom0, check=False), "whiten"),
                 (lambda: ops.bilevel_gae_raw(r2, values, mask, 1.0, 1.0, 0.95), "bilevel"),
                 (lambda: ops.bilevel_gae_raw(r2, values, mask, 1.0, 1.0, 0.95, exact=False), "bilevel-scan"),
                 (lambda: ops.discounted_returns(rewards, mask, 1.0), "rfpp"),
                 (lambda: ops.discounted_returns(rewards, mask, 1.0, exact=False), "rfpp-scan"),
                 (lambda: ops.remax(rewards, mask, base), "remax"),
                 (lambda: ops.remax(rewards, mask, base, exact=False), "remax-scan")]:
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps):
        fn()
    e.record()
    torch.cuda.synchronize()
    print(f"{name} B={B} T={T}: {1e3 * s.elapsed_time(e) / reps:.1f} us per call")
