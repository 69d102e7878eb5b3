#!/usr/bin/env python
"""Per-kernel measurement of every config of SURVEY.md 8(d) (C1..C5) on ONE B200, one JSON line per row.

    python bench.py --rows [--quick] [--only sokoban,generator,frozenlake,parse,c5,manager,rollout,rgb]  > profiles/rNN_kernels.jsonl

bench.py is the contract line (C4 step); this script is the supporting evidence for the other rows: each line carries
the kernel's device time (CUDA events on the launching stream, after warm-up, median of the repetitions), the
algorithmic bytes per unit, achieved GB/s against MEASURED_PEAKS.json, and the CPU port of the
reference (oracle/, single core, bounded sample) timed beside it - the same `cpu_baseline` role oracle/ has in bench.py,
which is the only way in (`bench.py --rows`).  Rows whose working set is a few hundred KB are
launch-latency-bound: they are reported as microseconds per call and their roofline fraction is not a quality claim.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from ragen_b200 import _lib, ops  # noqa: E402
from ragen_b200 import synth  # noqa: E402


def peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    return float(json.load(open(p))["hbm_gbs"]) if os.path.exists(p) else 6650.0


def dev_time(fn, reps=20, warm=3, inner=1):
    """median milliseconds of fn() by CUDA events on the current stream.  inner > 1 times that many back-to-back calls per
    event pair and divides (bandwidth-regime rows: a lone ~50 us launch otherwise carries ~10 us of launch/event gap)."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(inner):
            fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / inner)
    global _LAST_GRAPH_MS
    _LAST_GRAPH_MS = graph_time(fn, reps, inner) if inner > 1 else None
    return float(np.median(ts))


_LAST_GRAPH_MS = None


def graph_time(fn, reps, inner):
    """The same back-to-back calls captured once as a CUDA graph and replayed: GPU time only.  An op of two 15 us kernels issued from
    Python is bound by its ~45 us of host work; the replay shows what the kernels take.  None when the op cannot be captured."""
    try:
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            fn()
        cur.wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for _ in range(inner):
                fn()
        g.replay()
        torch.cuda.synchronize()
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            g.replay()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b) / inner)
        return float(np.median(ts))
    except Exception:
        torch.cuda.synchronize()
        return None


def cpu_time(fn, budget_s=5.0, min_reps=1):
    t0 = time.perf_counter()
    reps = 0
    while reps < min_reps or time.perf_counter() - t0 < budget_s:
        fn()
        reps += 1
        if reps >= 50:
            break
    return (time.perf_counter() - t0) / reps


def line(row, unit, units, ms, bytes_per_unit, cpu_units_per_s=None, cpu_sample=None, **extra):
    ach = bytes_per_unit * units / (ms * 1e-3) / 1e9 if bytes_per_unit else None
    out = {"row": row, "unit": unit, "units_per_launch": units, "ms": ms, "units_per_s": units / (ms * 1e-3),
           "bytes_per_unit": bytes_per_unit, "achieved_gbs": ach, "peak_gbs": peak_gbs(), "frac": ach / peak_gbs() if ach else None,
           "cpu_port_units_per_s": cpu_units_per_s, "cpu_sample": cpu_sample}
    global _LAST_GRAPH_MS
    if _LAST_GRAPH_MS is not None:  # the row's last timed op, replayed from a CUDA graph (GPU time without the host's share)
        out["ms_graph"] = _LAST_GRAPH_MS
        if bytes_per_unit:
            out["frac_graph"] = bytes_per_unit * units / (_LAST_GRAPH_MS * 1e-3) / 1e9 / peak_gbs()
        _LAST_GRAPH_MS = None
    out.update(extra)
    print(json.dumps(out), flush=True)
    return out


# ---- rgb_array observations: sprite blit ------------------------------------------------------------------------------
def rgb_rows(quick):
    from oracle import sokoban as S
    from ragen_b200.env.sokoban import SokobanBatch

    rows = []
    for (H, W, nb, n, tag) in [(10, 10, 2, 16384, "rgb_array blit 10x10, 16 Ki envs (1.26 GB of images)"),
                               (6, 6, 1, 16384, "rgb_array blit 6x6, 16 Ki envs"),
                               (6, 6, 1, 128, "rgb_array blit 6x6, C1's 128 envs")]:
        if quick and n > 4096:
            n = 4096
        fixed, state = synth.synthetic_rooms(n, H, W, nb, seed=4)
        b = SokobanBatch(n, (H, W), nb, 100, device=0)
        b.load_rooms(fixed, state)
        surfaces = S.synthetic_surfaces(16, seed=1)
        b.set_atlas(surfaces)
        out = torch.empty((n, H * 16, W * 16, 3), dtype=torch.uint8, device="cuda")

        def blit():
            _lib.check(b.lib.ragen_sokoban_render_rgb(b.ctx.handle, n, H, W, 16, _lib.ptr(b.state), _lib.ptr(b.meta),
                                                      _lib.ptr(b.atlas), None, 0, _lib.ptr(out), b.ctx.stream()))
        ms = dev_time(blit, reps=30, inner=8 if n > 1000 else 1)
        m = min(n, 64)

        def cpu():
            for j in range(m):
                S.room_to_rgb(state[j].reshape(H, W), fixed[j].reshape(H, W), surfaces)
        cs = cpu_time(cpu, 2.0)
        rows.append(line(tag, "image", n, ms, H * W * 768 + H * W + 8, m / cs, f"{m} rooms, oracle room_to_rgb (numpy), 1 core",
                         envs=n, kernel="sokoban_rgb_vec_kernel<16>", note="write-only traffic; sprites are synthetic"))
        del b, out
    return rows


# ---- C4(ii) / C1 step shapes: Sokoban step at several sizes -----------------------------------------------------------
def sokoban_rows(quick):
    from oracle import sokoban as S
    from ragen_b200.env.sokoban import SokobanBatch

    rows = []
    # packed: False byte rows, True nibble rows, 3 three-bit rows, 4 three-bit rows + rooms resting as nibbles (step_compact)
    for (H, W, nb, n, K, packed, tag) in [(10, 10, 2, 1 << 20, 1, 4, "C4 step 10x10 K=1, nibble rooms + three-bit glyph rows (bench default)"),
                                          (10, 10, 2, 1 << 20, 1, 3, "C4 step 10x10 K=1, three-bit glyph rows"),
                                          (10, 10, 2, 1 << 20, 5, 4, "C4 turn-fused 10x10 K=5, nibble rooms + three-bit glyph rows"),
                                          (8, 8, 2, 1 << 20, 1, 4, "8x8 K=1 at 1Mi envs, nibble rooms + three-bit glyph rows"),
                                          (6, 6, 1, 1 << 20, 1, 4, "6x6 K=1 at 1Mi envs, nibble rooms + three-bit glyph rows"),
                                          (10, 10, 2, 1 << 20, 1, True, "C4 step 10x10 K=1, packed glyph rows"),
                                          (10, 10, 2, 1 << 20, 1, False, "C4 step 10x10 K=1, byte glyph rows"),
                                          (10, 10, 2, 1 << 20, 5, True, "C4 turn-fused 10x10 K=5, packed glyph rows"),
                                          (6, 6, 1, 1 << 20, 1, True, "6x6 K=1 at 1Mi envs, packed glyph rows"),
                                          (6, 6, 1, 1 << 20, 1, False, "6x6 K=1 at 1Mi envs, byte glyph rows"),
                                          (6, 6, 1, 128, 5, True, "C1 shape 6x6 128 envs K=5")]:
        if quick and n > 1 << 18:
            n = 1 << 18
        fixed, state = synth.synthetic_rooms(n, H, W, nb, seed=4)
        compact = packed == 4
        fmt = 3 if compact else packed
        b = SokobanBatch(n, (H, W), nb, 100, device=0, compact=compact)
        b.load_rooms(fixed, state)
        g = torch.Generator(device="cuda").manual_seed(1)
        acts = torch.randint(1, 5, (4, n, K), dtype=torch.uint8, device="cuda", generator=g)
        i = [0]

        def step():
            b.step(acts[i[0] % 4], packed=fmt)
            i[0] += 1
        ms = dev_time(step, reps=50 if n > 1000 else 200, inner=8 if n > 1000 else 1)
        per_step = 3 * H * W + 1 + 10 + 8 + 1 + 1 + H * (W + 1)
        per_turn = per_step if K == 1 else (per_step - 1 + K)      # "(170+K)/K per env-step" variant: one render per turn
        # CPU port: the reference's per-env Python loop on a bounded sample
        m = 2048 if n > 2048 else n
        envs = []
        for j in range(m):
            e = S.SokobanOracle((H, W), nb, 100)
            e.load(fixed[j], state[j])
            envs.append(e)
        a_host = acts[0, :m].cpu().numpy()

        def cpu():
            for e, row in zip(envs, a_host):
                for a in row:
                    e.step(int(a))
        cs = cpu_time(cpu, 3.0)
        rows.append(line(tag, "env-step", n * K, ms, per_turn / K, m * K / cs, f"{m} envs x {K} actions, oracle/sokoban.py, 1 core",
                         envs=n, K=K, kernel="sokoban_tile_kernel", rooms="nibbles" if compact else "bytes",
                         glyph_rows="three-bit" if fmt == 3 and fmt is not True else "packed" if fmt else "bytes",
                         bytes_moved_per_unit=((H * W if compact else 2 * H * W) + 16 + 11 + (((3 * H * W + 15) // 16) * 2 if fmt == 3 and fmt is not True
                                                                                          else H * (W + 1) // 2 if fmt else H * (W + 1))) / K + 1))
        del b
    return rows


# ---- C4(i): room generator -------------------------------------------------------------------------------------------
def generator_rows(quick):
    from ragen_b200.env.sokoban import SokobanBatch

    rows = []
    for (H, W, nb, depth, n, tag) in [(6, 6, 1, 300, 8192, "6x6/1 box depth 300"), (8, 8, 2, 10, 8192, "8x8/2 boxes depth 10"),
                                      (10, 10, 2, 300, 2048 if quick else 16384, "C4 reset 10x10/2 boxes depth 300")]:
        b = SokobanBatch(1, (H, W), nb, 100, depth, device=0)
        seeds = list(range(n))
        t0 = time.perf_counter()
        _, _, st, explored = b.generate(seeds)
        torch.cuda.synchronize()
        first = time.perf_counter() - t0
        ms = dev_time(lambda: b.generate(seeds), reps=3, warm=0)
        ok = float((st == 0).float().mean())
        # solvability by the reference's own criterion (get_shortest_action_path, restated as oracle solve_bfs)
        # and mean solution length, on a bounded sample of the rooms this launch produced
        from oracle import sokoban as S
        f_all, s_all, _, _ = b.generate(seeds[:256])
        f_np, s_np, st_np = f_all.cpu().numpy(), s_all.cpu().numpy(), st[:256].cpu().numpy()
        lens, t_bfs = [], time.perf_counter()
        for j in np.flatnonzero(st_np == 0):
            if time.perf_counter() - t_bfs > (10.0 if quick else 40.0):
                break
            lens.append(len(S.solve_bfs(f_np[j], s_np[j], H, W)))
        solved = sum(1 for x in lens if x > 0)
        # the same check on the device for EVERY room of the launch (ragen_sokoban_solve = get_shortest_action_path)
        from ragen_b200.env.sokoban import solve_rooms
        f_gen, s_gen, st_gen, _ = b.generate(seeds)
        good = st_gen == 0
        fg, sg = f_gen[good].contiguous(), s_gen[good].contiguous()
        ms_solve = dev_time(lambda: solve_rooms(fg, sg, (H, W), nb), reps=2, warm=1)
        sst, slen, _, sexp = solve_rooms(fg, sg, (H, W), nb)
        rows.append(line(tag + " - device solver over the generated rooms", "room", int(good.sum()), ms_solve, 2 * H * W + 8 + 100, None, None,
                         solvable_fraction=float((sst == 0).float().mean()), mean_solution_len=float(slen.float().mean()),
                         max_solution_len=int(slen.max()), mean_states=float(sexp.float().mean()), kernel="sokoban_solve_kernel",
                         note="get_shortest_action_path on the device, reference BFS order (paths equal the reference's); no roofline claim"))
        rows.append(line(tag, "room", n, ms, 2 * H * W + 2, None, "reference: 4 ms/room (6x6), 1.0 s/room (10x10) on one core (SURVEY.md section 0)",
                         valid_fraction=ok, mean_explored=float(explored.float().mean()), first_call_s=first, kernel="sokoban_generate_kernel",
                         bfs_sample=len(lens), bfs_solvable=solved, bfs_mean_solution_len=(float(np.mean([x for x in lens if x > 0])) if solved else None),
                         note="no roofline claim: on-chip search, output bytes only; bfs_* = oracle solve_bfs (get_shortest_action_path contract, "
                              "MAX_DEPTH 100) on the first rooms of the launch within a CPU time budget"))
        del b
    return rows


# ---- C2: FrozenLake ----------------------------------------------------------------------------------------------------
def frozenlake_rows(quick):
    from oracle import toy_envs as TE
    from ragen_b200.env.frozen_lake import FrozenLakeBatch

    rows = []
    for n, tag in [(1024, "C2 FrozenLake 4x4 slippery, 1024 envs"), (1 << 18 if quick else 1 << 20, "FrozenLake 4x4 slippery at 1Mi envs")]:
        b = FrozenLakeBatch(n, 4, 0.8, True, device=0)
        seeds = np.arange(n)
        ms_reset = dev_time(lambda: b.reset(seeds, check=False), reps=5, warm=1)
        g = torch.Generator(device="cuda").manual_seed(2)
        acts = torch.randint(1, 5, (4, n, 1), dtype=torch.uint8, device="cuda", generator=g)
        i = [0]

        def step():
            b.step(acts[i[0] % 4])
            i[0] += 1
        ms = dev_time(step, reps=100, inner=8 if n > 1024 else 1)
        m = min(n, 1024)
        envs = [TE.FrozenLakeOracle() for _ in range(m)]
        for j, e in enumerate(envs):
            e.reset(j)
        a_host = (acts[0, :m, 0].cpu().numpy() - 1).tolist()

        def cpu():
            for e, a in zip(envs, a_host):
                e.step(a + 1)
                e.render()
        cs = cpu_time(cpu, 2.0)
        rows.append(line(tag, "env-step", n, ms, 109, m / cs, f"{m} envs x 1 step + render, oracle/toy_envs.py, 1 core", envs=n,
                         reset_ms=ms_reset, kernel="frozenlake_kernel<STEP>"))
    return rows


# ---- C3: parse + Bandit / Countdown scoring ----------------------------------------------------------------------------
def parse_rows(quick):
    from oracle import managers as M
    from oracle import toy_envs as TE

    nb = nc = 2048 if quick else 8192
    corpus = synth.synthetic_responses(nb, nc, seed=3)
    responses = corpus["responses"] if isinstance(corpus, dict) else corpus[0]
    dev = torch.device("cuda", 0)
    raw, off_np, text, offsets = ops.stage_responses(responses, dev)
    tables = ops.ActionTables([{1: "Phoenix", 2: "Dragon"}], dev)
    tid = torch.tensor([0] * nb + [-1] * nc, dtype=torch.int32, device=dev)
    bufs = ops.parse_actions_device(text, offsets, tables, tid, True, "||", 1)
    ms = dev_time(lambda: ops.parse_actions_device(text, offsets, tables, tid, True, "||", 1, out=bufs), reps=30)
    ms_e2e = dev_time(lambda: ops.parse_actions(responses, tables, tid, True, "||", 1, dev).to_host(), reps=3, warm=1)
    m = 2048
    cs = cpu_time(lambda: [M.parse_response(r, True, "||", 1) for r in responses[:m]], 3.0)
    n = len(responses)
    return [line("C3 parse + action lookup, 16k responses", "response", n, ms, len(raw) / n + 24, m / cs,
                 f"{m} responses, re-based oracle/managers.py:parse_response, 1 core", mean_response_bytes=len(raw) / n,
                 e2e_ms_with_staging_and_host_strings=ms_e2e, kernel="parse_kernel")]


# ---- C5: trajectory tensors + advantage ----------------------------------------------------------------------------------
def c5_rows(quick):
    from oracle import advantage as A
    from oracle import managers as M

    rows = []
    for B, L, tag in [(128, 2048, "C5 B=128 L=2048"), (2048 if quick else 8192, 2048, "B=8192 L=2048 (bandwidth regime)")]:
        INNER = 8 if B > 128 else 1  # B = 128 rows report the latency of one call, the bandwidth rows back-to-back launches
        rng = np.random.default_rng(5)
        ids = np.full((B, L), 151643, dtype=np.int64)
        scores = np.zeros((B, 5))
        nt = np.zeros(B, dtype=np.int32)
        for b in range(B):
            seq = [151644] + [7] * 20 + [151645, 198]
            k = int(rng.integers(1, 6))
            for _ in range(k):
                seq += [151644] + [9] * int(rng.integers(100, 180)) + [151645, 198]
                seq += [151644] + [11] * int(rng.integers(20, 160)) + [151645, 198]
            seq = seq[-L:]
            ids[b, L - len(seq):] = seq
            nt[b] = k
            scores[b, :k] = rng.integers(-2, 11, k) * 0.1
        d_ids, d_sc, d_nt = torch.from_numpy(ids).cuda(), torch.from_numpy(scores).cuda(), torch.from_numpy(nt).cuda()
        ms = dev_time(lambda: ops.masks_scores(d_ids, d_sc, d_nt, 151644, 151645, False, True, True), reps=50, inner=INNER)
        t_ids = torch.from_numpy(ids[:128])
        sc_list = [scores[b, :nt[b]].tolist() for b in range(128)]
        cs = cpu_time(lambda: M.masks_and_scores(t_ids, sc_list, 151644, 151645), 2.0)
        rows.append(line(f"masks+scores {tag}", "token", B * L, ms, 14, 128 * L / cs, "128 rows, torch CPU oracle (all host threads)", kernel="masks_scores_kernel"))
        sc, lm, rm = ops.masks_scores(d_ids, d_sc, d_nt, 151644, 151645, False, True, True)
        T = L - 1
        mask = lm.float()
        rewards = sc.clone()
        values = torch.randn(B, T, device="cuda") * mask
        ms_raw = dev_time(lambda: ops.gae_raw(rewards, values, lm, 1.0, 1.0), reps=50, inner=INNER)

        def full():
            adv, ret, mom = ops.gae_raw(rewards, values, lm, 1.0, 1.0)
            ops.whiten_(adv, mom, check=False)
        ms_full = dev_time(full, reps=50, inner=INNER)
        r_c, v_c, m_c = rewards[:128].cpu(), values[:128].cpu(), mask[:128].cpu()
        cs = cpu_time(lambda: A.gae(r_c, v_c, m_c, 1.0, 1.0), 3.0)
        rows.append(line(f"GAE+whiten {tag}", "token", B * T, ms_full, 25, 128 * T / cs, "128 rows, torch CPU loop as the reference (all host threads)",
                         gae_only_ms=ms_raw, kernel="gae_kernel + whiten_kernel (bit-exact sequential order)"))
        ms_scan = dev_time(lambda: ops.gae_raw(rewards, values, lm, 1.0, 1.0, exact=False), reps=50, inner=INNER)

        def full_scan():
            adv, ret, mom = ops.gae_raw(rewards, values, lm, 1.0, 1.0, exact=False)
            ops.whiten_(adv, mom, check=False)
        ms_scan_full = dev_time(full_scan, reps=50, inner=INNER)
        rows.append(line(f"GAE (parallel scan)+whiten {tag}", "token", B * T, ms_scan_full, 25, 128 * T / cs,
                         "128 rows, torch CPU loop as the reference (all host threads)", gae_only_ms=ms_scan,
                         gae_only_frac_17B=17 * B * T / (ms_scan * 1e-3) / 1e9 / peak_gbs(),
                         kernel="gae_scan_kernel + whiten_kernel (within 1e-5 of the sequential order)"))
        # bi-level needs a reward on the last loss-mask position
        r2 = torch.zeros_like(rewards)
        last = (mask * torch.arange(1, T + 1, device="cuda")).argmax(dim=1)
        r2[torch.arange(B), last] = 1.0
        ms_bi = dev_time(lambda: ops.bilevel_gae_raw(r2, values, lm, 1.0, 1.0, 0.95), reps=50, inner=INNER)
        r2c = r2[:16].cpu()
        cs = cpu_time(lambda: A.bilevel_gae(r2c, v_c[:16], m_c[:16], 1.0, 1.0, 0.95), 3.0)
        rows.append(line(f"bi-level GAE {tag}", "token", B * T, ms_bi, 17, 16 * T / cs, "16 rows, oracle/advantage.py loop as the reference", kernel="bilevel_gae_kernel"))
        ms_bis = dev_time(lambda: ops.bilevel_gae_raw(r2, values, lm, 1.0, 1.0, 0.95, exact=False), reps=50, inner=INNER)
        rows.append(line(f"bi-level GAE (block scans) {tag}", "token", B * T, ms_bis, 17, 16 * T / cs, "16 rows, oracle/advantage.py loop as the reference",
                         kernel="bilevel_scan_bulk_kernel (within 1e-5 of the sequential order)"))
        gi = torch.arange(B, device="cuda", dtype=torch.int32) // 16
        ms_g = dev_time(lambda: ops.grpo_outcome(rewards, lm, gi, B // 16), reps=50, inner=INNER)
        rows.append(line(f"GRPO {tag}", "token", B * T, ms_g, 9, None, None, kernel="grpo_scores_kernel + grpo_apply_kernel"))
        ms_r = dev_time(lambda: ops.outcome_advantage(rewards, lm, gi, B // 16, "rloo"), reps=50, inner=INNER)
        rows.append(line(f"RLOO {tag}", "token", B * T, ms_r, 9, None, None, kernel="grpo_scores_kernel + outcome_apply_kernel"))

        def rfpp():
            ret, m8, mom = ops.discounted_returns(rewards, lm, 1.0)
            ops.whiten_masked(ret, m8, mom, check=False)
        ms_p = dev_time(rfpp, reps=50, inner=INNER)
        rows.append(line(f"REINFORCE++ {tag}", "token", B * T, ms_p, 4 + 1 + 4 + 4 + 1 + 4, None, None, kernel="returns_kernel<0> + whiten_masked_kernel"))

        def rfpp_scan():
            ret, m8, mom = ops.discounted_returns(rewards, lm, 1.0, exact=False)
            ops.whiten_masked(ret, m8, mom, check=False)
        ms_ps = dev_time(rfpp_scan, reps=50, inner=INNER)
        rows.append(line(f"REINFORCE++ (block scans) {tag}", "token", B * T, ms_ps, 4 + 1 + 4 + 4 + 1 + 4, None, None,
                         kernel="returns_scan_bulk_kernel<0> + whiten_masked_kernel (within 1e-5 of the sequential order)"))
        base = torch.randn(B, device="cuda")
        ms_x = dev_time(lambda: ops.remax(rewards, lm, base), reps=50, inner=INNER)
        rows.append(line(f"ReMax {tag}", "token", B * T, ms_x, 4 + 1 + 4 + 4, None, None, kernel="returns_kernel<1>"))
        ms_xs = dev_time(lambda: ops.remax(rewards, lm, base, exact=False), reps=50, inner=INNER)
        rows.append(line(f"ReMax (block scans) {tag}", "token", B * T, ms_xs, 4 + 1 + 4 + 4, None, None,
                         kernel="returns_scan_bulk_kernel<1> (within 1e-5 of the sequential order)"))
        lp = -torch.rand(B, T, device="cuda")
        rf = lp + 0.1 * torch.randn(B, T, device="cuda")
        ms_k = dev_time(lambda: ops.kl_rewards(rewards, lp, rf, lm, "low_var_kl", 0.05), reps=50, inner=INNER)
        rows.append(line(f"KL-in-reward {tag}", "token", B * T, ms_k, 4 * 3 + 1 + 4, None, None, kernel="kl_rewards_kernel"))
        ms_f = dev_time(lambda: ops.group_filter_stats(sc, B // 16, 16), reps=50, inner=INNER)
        rows.append(line(f"rollout filter stats {tag}", "token", B * T, ms_f, 4, None, None, kernel="filter_stats_kernel"))
        pen = torch.zeros(B, device="cuda")
        ms_n = dev_time(lambda: ops.group_norm(sc[:, -1], pen, gi, B // 16, "mean_std"), reps=50, inner=INNER)
        rows.append(line(f"group reward norm {tag}", "row", B, ms_n, 16, None, None, kernel="norm_stats_kernel + norm_apply_kernel"))
    return rows


# ---- C1: the manager seam (host + device) ----------------------------------------------------------------------------------
def manager_rows(quick):
    """EnvStateManager.reset / step through the public API at C1's size and at 16 Ki / 1 Mi envs, wall clock (host + device), with
    the CPU port of the reference's manager loop (oracle/managers.py: one Python env object per env, serial step loop with a text
    render per step and per turn, es_manager.py:93-95, 149-169) timed beside it on a bounded sample of the same inputs."""
    from oracle import cfg as C
    from oracle import managers as M
    from oracle import sokoban as S
    from ragen_b200.llm_agent.es_manager import EnvStateManager

    names = ["Up", "Down", "Left", "Right"]
    for groups, tag in [(8, "C1 EnvStateManager 128 envs x 5 turns"), (1024, "EnvStateManager 16 Ki envs x 5 turns"),
                        (16384 if quick else 65536, "EnvStateManager 1 Mi envs x 2 turns")]:
        cfg = C.base_config(tags=("SimpleSokoban",), n_groups=(groups,), group_size=16)
        mgr = EnvStateManager(cfg, mode="train")
        n = groups * 16
        turns = 5 if n <= 16384 else 2
        t0 = time.perf_counter(); mgr.reset(seed=123); torch.cuda.synchronize(); first = time.perf_counter() - t0
        t0 = time.perf_counter(); cache = mgr.reset(seed=123); torch.cuda.synchronize(); reset_s = time.perf_counter() - t0
        b = mgr.blocks[0].batch
        fixed0, state0 = b.fixed.cpu().numpy(), b.state.cpu().numpy()
        rng = np.random.default_rng(1)
        alive = [c["env_id"] for c in cache]
        turn_s, steps, script = [], 0, []
        for _ in range(turns):
            inputs = []
            for e in alive:
                acts = [names[int(rng.integers(0, 4))] if rng.random() > 0.1 else "Jump" for _ in range(int(rng.integers(1, 4)))]
                resp = "<think>t</think><answer>" + " || ".join(acts) + "</answer>"
                inputs.append({"env_id": e, "llm_raw_response": resp, "llm_response": resp, "actions": acts})
            script.append(inputs)
            t0 = time.perf_counter()
            out = mgr.step(inputs)
            turn_s.append(time.perf_counter() - t0)
            steps += sum(len(i["actions"]) for i in inputs)
            alive = [c["env_id"] for c in out]
            if not alive:
                break
        # the port on the first m envs, same rooms, same action strings
        m = min(n, 2048)
        lookup = {1: "Up", 2: "Down", 3: "Left", 4: "Right"}
        envs = [{"tag": "SimpleSokoban", "env": S.SokobanOracle((6, 6), 1, 100), "max_actions_per_traj": 10, "action_lookup": lookup}
                for _ in range(m)]
        port = M.ManagerOracle(envs, 16)
        port.reset(0, reset_fn=lambda i, e, seed: e["env"].load(fixed0[i], state0[i]))
        cpu_turn_s = []
        for inputs in script:
            sub = [i for i in inputs if i["env_id"] < m]
            t0 = time.perf_counter()
            port.step(sub)
            cpu_turn_s.append(time.perf_counter() - t0)
        print(json.dumps({"row": tag + " (wall clock, host + device)", "envs": n, "reset_s": reset_s, "first_reset_s": first,
                          "turn_ms": [1e3 * t for t in turn_s], "env_steps": steps,
                          "cpu_port_turn_ms": [1e3 * t for t in cpu_turn_s], "cpu_port_envs": m,
                          "cpu_port_turn_ms_scaled_to_envs": [1e3 * t * n / m for t in cpu_turn_s],
                          "cpu_sample": f"{m} envs, oracle/managers.py ManagerOracle over oracle/sokoban.py envs, 1 core (the reference's execution model)",
                          "note": "the Python history bookkeeping the API promises (one dict per env per turn) is what remains on the host"}), flush=True)
        del mgr


# ---- C5: rollout -> advantage through the public API, scripted policy -----------------------------------------------------
def rollout_rows(quick):
    """8 groups x 16 SimpleSokoban envs, 5 turns, LLMAgentProxy.rollout with the scripted stand-in policy of tests/fake_tokenizer.py
    (the LLM is out of scope), then compute_advantage (GAE and GRPO).  Wall clock per phase; the policy's own time is reported
    separately so that what remains is this package's share of a rollout.  Three arms: assembly="device" (token trajectories
    gathered on the GPU from pre-tokenised segments, llm_agent/assembly.py), assembly="host" (the reference's string path kept as
    the explicit alternative), and the CPU port of the whole reference loop (oracle/rollout.py, pinned on the reference-made
    rollouts.json) on the same rooms with the same policy."""
    from fake_tokenizer import FakeQwenTokenizer, ScriptedWg
    from oracle import cfg as C
    from oracle.rollout import RolloutPort
    from ragen_b200.env.sokoban import SokobanBatch
    from ragen_b200.llm_agent.agent_proxy import LLMAgentProxy
    from ragen_b200.llm_agent.ctx_manager import DataProto
    from ragen_b200.trainer import AdvantageEstimator, compute_advantage

    class Timed(ScriptedWg):
        spent = 0.0

        def generate_sequences(self, lm_inputs):
            t0 = time.perf_counter()
            out = super().generate_sequences(lm_inputs)
            Timed.spent += time.perf_counter() - t0
            return out

    cfg = C.base_config(tags=("SimpleSokoban",), n_groups=(8,), group_size=16, reward_normalization={"grouping": "state", "method": "mean_std"})
    tok = FakeQwenTokenizer()
    import gc

    gc.collect()
    gc.freeze()  # all three arms: the interpreter's full collections (~8 ms each over torch's object graph) otherwise land at random inside a 20 ms rollout
    arms = {}
    for assembly in ("device", "host"):
        proxy = LLMAgentProxy(cfg, Timed(tok, DataProto), tok, assembly=assembly)
        res = []
        for rep in range(6):
            Timed.spent = 0.0
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            dp = proxy.rollout(DataProto(batch=None, non_tensor_batch=None, meta_info={}), val=False, seed=123 + rep)
            torch.cuda.synchronize()
            t_roll = time.perf_counter() - t0
            batch = dp.batch
            B, T = batch["rm_scores"].shape
            batch["response_mask"] = batch["loss_mask"].float()
            batch["token_level_rewards"] = batch["rm_scores"].clone()
            batch["values"] = torch.randn(B, T) * batch["response_mask"]
            dp.non_tensor_batch["uid"] = dp.non_tensor_batch["group_ids"]
            t0 = time.perf_counter()
            compute_advantage(dp, AdvantageEstimator.GAE, gamma=1.0, lam=1.0)
            t_gae = time.perf_counter() - t0
            t0 = time.perf_counter()
            compute_advantage(dp, AdvantageEstimator.GRPO, multi_turn=True)
            t_grpo = time.perf_counter() - t0
            res.append({"rollout_s": t_roll, "policy_stub_s": Timed.spent, "ours_s": t_roll - Timed.spent, "gae_s": t_gae, "grpo_s": t_grpo,
                        "B": B, "T": T, "success": float(dp.meta_info["metrics"].get("SimpleSokoban/success", 0.0))})
        arms[assembly] = res
    # CPU port of the reference loop: same rooms (generated once per group seed), same scripted policy
    gen = SokobanBatch(1, (6, 6), 1, 100, 300, device=0)
    port_res = []
    for rep in range(3):
        f, s, st, _ = gen.generate([123 + rep + g for g in range(8)])
        assert not bool(st.any())
        f, s = f.cpu().numpy(), s.cpu().numpy()
        fixed = np.stack([f[i // 16] for i in range(128)])
        state = np.stack([s[i // 16] for i in range(128)])
        port = RolloutPort(tok, 8, 16, env_instruction=C.SOKOBAN_INSTRUCTION, max_tokens=100, norm_method="mean_std")
        wg = ScriptedWg(tok, DataProto)
        spent = [0.0]

        def policy(lm):
            t0 = time.perf_counter()
            ids = np.array(lm["env_ids"], dtype=object)
            out = wg.generate_sequences(DataProto(batch={"input_ids": lm["input_ids"]}, non_tensor_batch={"env_ids": ids, "group_ids": ids}))
            spent[0] += time.perf_counter() - t0
            return out.non_tensor_batch["response_texts"]

        t0 = time.perf_counter()
        final = port.rollout(fixed, state, policy)
        t_roll = time.perf_counter() - t0
        port_res.append({"rollout_s": t_roll, "policy_stub_s": spent[0], "port_s": t_roll - spent[0], "B": int(final["input_ids"].shape[0]),
                         "T": int(final["input_ids"].shape[1]) - 1})
    med = lambda xs: float(np.median(xs))  # noqa: E731
    print(json.dumps({"row": "C5 rollout -> advantage through LLMAgentProxy / compute_advantage, 8 groups x 16 SimpleSokoban, 5 turns, scripted policy "
                             "(wall clock, CPU tensors in and out as the reference's driver has them; the first repetition of an arm includes warm-up "
                             "and the one-off tokenizer decomposability check; gc.freeze() before all arms)",
                      "ours_s_device_assembly_median": med([r["ours_s"] for r in arms["device"][1:]]),
                      "ours_s_host_assembly_median": med([r["ours_s"] for r in arms["host"][1:]]),
                      "cpu_port_s_median": med([r["port_s"] for r in port_res]),
                      "cpu_sample": "oracle/rollout.py RolloutPort: the reference's loop (per-env Python envs, whole history re-tokenised per turn), "
                                    "rooms supplied, same scripted policy, 1 core; the byte-level stand-in tokenizer is Python, a Rust tokenizer "
                                    "would make the host legs faster",
                      "device": arms["device"], "host": arms["host"], "cpu_port": port_res}), flush=True)


def main(quick=False, only=""):
    a = argparse.Namespace(quick=quick, only=only)
    torch.cuda.set_device(0)
    todo = {"sokoban": sokoban_rows, "generator": generator_rows, "frozenlake": frozenlake_rows, "parse": parse_rows, "c5": c5_rows,
            "manager": manager_rows, "rollout": rollout_rows, "rgb": rgb_rows}
    for name, fn in todo.items():
        if a.only and name not in a.only.split(","):
            continue
        fn(a.quick)
