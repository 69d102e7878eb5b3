"""Golden solution paths for the Sokoban solver: the UNMODIFIED get_shortest_action_path (ragen/env/sokoban/utils.py:11-101)
run in this container on the golden rooms of sokoban_rooms.npz (and on a few hand-made edge rooms).

    python tests/golden/make_golden_paths.py        # writes tests/golden/sokoban_paths.json

The reference BFS deep-copies a board per child, so 10x10 rooms take a long time; every room gets a CPU budget and rooms that
exceed it are left out (recorded in `skipped`)."""
import json
import os
import signal
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import ref_loader  # noqa: E402


class Timeout(Exception):
    pass


def _alarm(signum, frame):
    raise Timeout()


def main():
    U = ref_loader.sokoban_utils()
    rooms = np.load(os.path.join(HERE, "sokoban_rooms.npz"))
    out, skipped = [], []
    signal.signal(signal.SIGALRM, _alarm)

    def run(tag, H, W, fixed, state, max_depth=100, budget=60):
        f = np.asarray(fixed, dtype=np.int64).reshape(H, W)
        s = np.asarray(state, dtype=np.int64).reshape(H, W)
        signal.alarm(budget)
        try:
            path = U.get_shortest_action_path(f, s, MAX_DEPTH=max_depth)
        except Timeout:
            skipped.append(tag)
            return
        finally:
            signal.alarm(0)
        out.append({"tag": tag, "H": H, "W": W, "max_depth": max_depth, "fixed": f.reshape(-1).tolist(), "state": s.reshape(-1).tolist(),
                    "path": [int(a) for a in path]})

    for name, take in (("s6", 40), ("s6d30", 12), ("s6b3", 8), ("s7x9", 8), ("s8", 24), ("s10", 8)):
        cfg = rooms[f"{name}_cfg"]
        H, W = int(cfg[0]), int(cfg[1])
        ok = rooms[f"{name}_ok"]
        n = 0
        for j in np.flatnonzero(ok):
            if n >= take:
                break
            run(f"{name}/{int(rooms[f'{name}_seeds'][j])}", H, W, rooms[f"{name}_fixed"][j], rooms[f"{name}_state"][j],
                budget=240 if name == "s10" else 60)
            n += 1
    # depth limit: the same rooms with MAX_DEPTH below / at the solution length
    for e in [x for x in list(out) if x["tag"].startswith("s6/") and len(x["path"]) >= 3][:6]:
        L = len(e["path"])
        for md in (L - 1, L):
            run(e["tag"] + f"@{md}", e["H"], e["W"], e["fixed"], e["state"], max_depth=md)
    # edge rooms: already solved, unsolvable (box in a corner), no box at all
    solved = [0, 0, 0, 0, 0,  0, 5, 3, 1, 0,  0, 1, 1, 1, 0,  0, 0, 0, 0, 0]
    run("edge/solved", 4, 5, [0, 0, 0, 0, 0,  0, 1, 2, 1, 0,  0, 1, 1, 1, 0,  0, 0, 0, 0, 0], solved)
    run("edge/corner", 4, 5, [0, 0, 0, 0, 0,  0, 1, 1, 1, 0,  0, 1, 1, 2, 0,  0, 0, 0, 0, 0], [0, 0, 0, 0, 0,  0, 4, 1, 1, 0,  0, 1, 5, 2, 0,  0, 0, 0, 0, 0])
    run("edge/nobox", 4, 5, [0, 0, 0, 0, 0,  0, 1, 1, 1, 0,  0, 1, 1, 1, 0,  0, 0, 0, 0, 0], [0, 0, 0, 0, 0,  0, 5, 1, 1, 0,  0, 1, 1, 1, 0,  0, 0, 0, 0, 0])
    json.dump({"rooms": out, "skipped": skipped, "made_by": "tests/golden/make_golden_paths.py (unmodified get_shortest_action_path)"},
              open(os.path.join(HERE, "sokoban_paths.json"), "w"))
    print(len(out), "paths;", "skipped:", skipped)
    print("lengths:", sorted(len(x["path"]) for x in out))


if __name__ == "__main__":
    main()
