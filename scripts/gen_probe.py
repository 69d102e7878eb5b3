"""Generator timing probe (GPU): distribution of search lengths and where the launch time goes."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from ragen_b200.env.sokoban import SokobanBatch  # noqa: E402


def timed(b, seeds):
    torch.cuda.synchronize()
    a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    out = b.generate(seeds)
    e.record()
    torch.cuda.synchronize()
    return a.elapsed_time(e), out


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    b = SokobanBatch(16, (10, 10), 2, 100)
    b.search_depth = 300
    timed(b, list(range(64)))
    ms, (_, _, status, explored) = timed(b, list(range(n)))
    ex = explored.cpu().numpy()
    print(f"n={n} ms={ms:.1f} mean={ex.mean():.0f} max={ex.max()} p50={np.percentile(ex, 50):.0f} p90={np.percentile(ex, 90):.0f} "
          f"p99={np.percentile(ex, 99):.0f} p999={np.percentile(ex, 99.9):.0f} n_cap={(ex >= 299999).sum()} sum={ex.sum()}")
    order = np.argsort(-ex)
    for k in (1, 8, 148):
        ms_k, (_, _, _, e2) = timed(b, [int(s) for s in order[:k]])
        print(f"top-{k} rooms alone: {ms_k:.1f} ms, last-search explored max {int(e2.max())} -> {1e6 * ms_k / max(int(e2.max()), 1):.0f} ns/state")
    easy = [int(s) for s in np.argsort(ex)[:64]]
    for k in (1, 32, 64):
        ms_k, (_, _, _, e2) = timed(b, easy[:k])
        print(f"easiest {k} rooms alone: {ms_k:.2f} ms, explored {e2.cpu().numpy()[:4]}")
    small = [int(s) for s in np.argsort(ex)[: n // 2]]
    ms_s, (_, _, _, e3) = timed(b, small)
    print(f"easiest half ({len(small)} rooms): {ms_s:.1f} ms, explored sum {int(e3.sum())} max {int(e3.max())}")


main()
