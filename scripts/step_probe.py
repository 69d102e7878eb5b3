"""Step-kernel timing probe (GPU): back-to-back launches of one room shape, for CUDA-event timing and as an ncu target.

    python scripts/step_probe.py H W num_boxes n K packed reps     (packed: 0 bytes, 1 nibbles, 3 three-bit rows, 4 three-bit rows + compact rooms)
"""
import sys

import torch

sys.path.insert(0, ".")
from ragen_b200 import synth  # noqa: E402
from ragen_b200.env.sokoban import SokobanBatch  # noqa: E402


def main():
    a = [int(x) for x in sys.argv[1:]] + [6, 6, 1, 1 << 20, 1, 1, 200][len(sys.argv) - 1:]
    H, W, nb, n, K, packed, reps = a
    fixed, state = synth.synthetic_rooms(n, H, W, nb, seed=4)
    compact = packed == 4
    packed = 3 if compact else packed
    b = SokobanBatch(n, (H, W), nb, 100, device=0, compact=compact)
    b.load_rooms(fixed, state)
    g = torch.Generator(device="cuda").manual_seed(1)
    acts = torch.randint(1, 5, (4, n, K), dtype=torch.uint8, device="cuda", generator=g)
    for i in range(5):
        b.step(acts[i % 4], packed=(3 if packed == 3 else bool(packed)))
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(reps):
        b.step(acts[i % 4], packed=(3 if packed == 3 else bool(packed)))
    e.record()
    torch.cuda.synchronize()
    us = 1e3 * s.elapsed_time(e) / reps
    moved = (H * W if compact else 2 * H * W) + 16 + 11 + (((3 * H * W + 15) // 16) * 2 if packed == 3 else H * (W + 1) // 2 if packed else H * (W + 1)) + K
    print(f"{H}x{W} n={n} K={K} packed={packed}{" compact" if compact else ""}: {us:.1f} us/launch, {n * K / us * 1e6:.3e} env-steps/s, "
          f"{n * moved / us * 1e-3:.0f} GB/s on the {moved} B/env this launch moves")


main()
