"""Latency of one ContextManager._parse_batch call at C5's 128 responses, back to back and after the GPU sat idle."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ragen_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)
tables = ops.ActionTables([{1: "Up", 2: "Down", 3: "Left", 4: "Right"}], dev)
responses = ["<think>plan %d</think><answer>Up || Down || Left</answer>" % i for i in range(128)]
tid = [0] * 128
for idle_ms in (0, 1, 7, 30):
    ts = {"stage": 0.0, "launch": 0.0, "slab.cpu": 0.0, "host loop": 0.0}
    for it in range(30):
        time.sleep(idle_ms * 1e-3)
        t0 = time.perf_counter()
        res = ops.parse_actions(responses, tables, tid, True, "||", 5, dev)
        t1 = time.perf_counter()
        h = res.slab.cpu()
        t2 = time.perf_counter()
        out = res.to_host()
        t3 = time.perf_counter()
        if it >= 5:
            ts["launch"] += t1 - t0
            ts["slab.cpu"] += t2 - t1
            ts["host loop"] += t3 - t2
    print(idle_ms, {k: round(1e3 * v / 25, 3) for k, v in ts.items()}, flush=True)
