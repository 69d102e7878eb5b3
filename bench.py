#!/usr/bin/env python
"""bench.py — Sokoban env-steps/sec (step + render + reward), BASELINE.json's metric.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (config 4 of BASELINE.json, the one the metric is quoted on): Sokoban 10x10, 2 boxes,
max_steps 100; 1 Mi envs PER GPU (weak scaling: env groups are independent, no data-path collective;
at N=8 that is 8 Mi envs); one "step" = every env takes one action: state update + reward + done/info
flags + text glyph render, one kernel launch.  Two such batches are resident and stepped alternately, so
the distinct lines between two touches of a line (~226 MB with the default formats) exceed the 126 MB L2
and no L2 flush is needed between timed iterations.  The rooms rest in HBM two cells per byte (--rooms
nibbles, the default: ragen_sokoban_step_compact; --rooms bytes = one cell per byte).

`value`   : device-timed (CUDA events, max over ranks), inputs resident in HBM.
`e2e`     : same step through SokobanBatch.step_host — the actions come from pinned HOST memory and the
            reward / done / flags / n_executed / glyph rows go back to pinned HOST memory inside the timed region.
            The glyph rows travel as three bits per cell (--obs packed3, the default; --obs packed = two 4-bit codes per byte, --obs bytes = one code per byte).
`roofline`: algorithmic bytes per env-step (431 B for 10x10) x envs / mean launch time.
`cpu_baseline` / `--impl reference`: the reference's execution model (one Python env object per env,
            serial step loop, text render per step — oracle/sokoban.py, a port because gym_sokoban is not
            installable here) on all host cores via multiprocessing, on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

H, W, NUM_BOXES, MAX_STEPS = 10, 10, 2, 100
BYTES_PER_STEP = 3 * H * W + 1 + 10 + 8 + 1 + 1 + H * (W + 1)  # = 431, SURVEY.md 8(d) C4(ii)
MOVED_PER_STEP = 2 * H * W + 2 * 8 + 1 + 8 + 1 + 1 + 1 + H * (W + 1)  # = 338: state r+w, 8-byte record r+w, action, results, glyph rows
TRAFFIC_FILE = "traffic_r2.json"  # DRAM bytes per launch of the bench's kernels, summed over >= 100 steady-state launches by ncu
METRIC = "sokoban_env_steps_per_sec"
UNIT = "env-steps/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--envs", type=int, default=1 << 20, help="envs per GPU")
    ap.add_argument("--ref-envs", type=int, default=0, help="envs per reference step (0 = auto from core count)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-reset", action="store_true", help="skip the batched-reset leg (room generator + solver at config 4 size)")
    ap.add_argument("--no-chains", action="store_true", help="`value` from one chain of launches instead of one chain per resident batch")
    ap.add_argument("--no-graph", action="store_true", help="issue the launches of a step from Python instead of replaying a CUDA graph")
    ap.add_argument("--launches-per-step", type=int, default=128,
                    help="back-to-back step launches (alternating batches) that make ONE bench step; env-steps per step = envs x this")
    ap.add_argument("--reset-groups", type=int, default=65536, help="group rooms of the reset leg, over all ranks (0 = skip)")
    ap.add_argument("--reset-distinct", type=int, default=1 << 20, help="distinct rooms of the reset leg's group_size=1 variant, over all ranks (0 = skip)")
    ap.add_argument("--obs", default="packed3", choices=["packed3", "packed", "bytes"],
                    help="glyph-row format of the render: three bits per cell without separators (packed3), two 4-bit codes per byte "
                         "(packed, what the managers consume) or one code per byte (bytes)")
    ap.add_argument("--rooms", default="glyph3", choices=["glyph3", "nibbles", "bytes"],
                    help="how the rooms rest in HBM: one cell per byte, two (ragen_sokoban_step_compact), or as three-bit glyph codes that "
                         "are their own rendered observation (ragen_sokoban_step_glyph3); the last two need --obs packed3")
    ap.add_argument("--rows", action="store_true", help="per-kernel lines for every SURVEY.md 8(d) config instead of the contract line")
    ap.add_argument("--quick", action="store_true", help="with --rows: smaller batches")
    ap.add_argument("--only", default="", help="with --rows: comma list of sokoban,generator,frozenlake,parse,c5,manager,rollout")
    return ap.parse_args()


def line_bytes_per_env(obs, rooms):
    """Bytes of distinct cache lines one launch touches per env: room, 8-byte record, action, results, separate glyph rows."""
    return {"nibbles": 50, "bytes": 100, "glyph3": 38}[rooms] + 8 + 1 + 11 + (0 if rooms == "glyph3" else {"packed3": 38, "packed": 55, "bytes": 110}[obs])


def resident_batches(envs_per_gpu, obs, rooms):
    """How many batches are stepped in rotation so that the lines between two touches of a line exceed the 126 MB L2 by 1.6x."""
    per_batch = envs_per_gpu * line_bytes_per_env(obs, rooms)
    need = -(-int(1.6 * 126 * (1 << 20)) // max(per_batch, 1))
    return max(2, min(need + (need & 1), 16))  # even, so that graph capture and the pipelined legs pair up


def workload_config(n_gpus, envs_per_gpu, obs="packed", rooms="bytes", launches_per_step=128):
    rows = {"packed3": "three bits per cell, no separators (38 B per env; ragen_sokoban_step_packed3)",
            "packed": "two 4-bit glyph codes per byte (55 B per env)", "bytes": "one glyph code per byte (110 B per env)"}[obs]
    return {"glyph_rows": rows, "rooms_in_hbm": {"nibbles": "two cells per byte (ragen_sokoban_step_compact)", "bytes": "one cell per byte",
                                                "glyph3": "three-bit glyph codes, the room is its own rendered observation (ragen_sokoban_step_glyph3)"}[rooms],
            "workload": "sokoban_10x10_2box_step_render_reward", "dim_room": [H, W], "num_boxes": NUM_BOXES,
            "max_steps": MAX_STEPS, "envs_per_gpu": envs_per_gpu, "envs_total": envs_per_gpu * n_gpus,
            "actions_per_launch": 1, "launches_per_step": launches_per_step, "env_steps_per_step": envs_per_gpu * n_gpus * launches_per_step,
            "rooms": "synthetic random-walk rooms (ragen_b200/synth.py)",
            "actions": "uniform 1..4", "l2": "inputs larger than L2: %d batches of envs_per_gpu envs stepped in rotation, %d MB of distinct lines between two touches of a line (L2 is 126 MB)" % (resident_batches(envs_per_gpu, obs, rooms), (envs_per_gpu * resident_batches(envs_per_gpu, obs, rooms) * line_bytes_per_env(obs, rooms)) >> 20),
            "parallelism": f"env-sharded x{n_gpus}"}


# ------------------------------------------------------------------------------------------------
# reference arm: the CPU execution model of the reference on all host cores
# ------------------------------------------------------------------------------------------------
_W = {}


def _ref_init(seed, n):
    from oracle import sokoban as S  # the CPU port IS what this arm measures
    from ragen_b200.synth import synthetic_rooms
    import numpy as np

    fixed, state = synthetic_rooms(n, H, W, NUM_BOXES, seed)
    envs = []
    for i in range(n):
        e = S.SokobanOracle((H, W), NUM_BOXES, MAX_STEPS)
        e.load(fixed[i], state[i])
        envs.append(e)
    _W["envs"] = envs
    _W["rng"] = np.random.default_rng(seed)
    return n


def _ref_run(spec):
    """warmup + steps serial passes over this worker's envs; returns the seconds the timed passes took."""
    warmup, steps = spec
    envs, rng = _W["envs"], _W["rng"]
    spent = 0.0
    for it in range(warmup + steps):
        acts = rng.integers(1, 5, len(envs)).tolist()
        t0 = time.perf_counter()
        tot = 0.0
        for e, a in zip(envs, acts):  # es_manager.py:149-169: serial loop, one env.step (with text render) each
            obs, r, d, info = e.step(a)
            tot += r
        if it >= warmup:
            spent += time.perf_counter() - t0
    return spent


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import multiprocessing as mp

    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    per_core = (args.ref_envs // cores) if args.ref_envs else 4096
    per_core = max(per_core, 1)
    ctx = mp.get_context("fork")
    pools = [ctx.Pool(1) for _ in range(cores)]  # one single-worker pool per core = one env shard per core
    try:
        for r in [p.apply_async(_ref_init, (100 + i, per_core)) for i, p in enumerate(pools)]:
            r.get()
        # every worker runs ALL its steps on its own (no per-step round trip through the parent inside the timed region);
        # the job's time is the slowest worker's
        times = [r.get() for r in [p.apply_async(_ref_run, ((args.warmup, args.steps),)) for p in pools]]
    finally:
        for p in pools:
            p.terminate()
    total = max(times)
    n = per_core * cores
    value = n * args.steps / total
    sample = f"{n} envs ({per_core}/core x {cores} procs) x {args.steps} steps, 1 action each, text render per step"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": workload_config(args.gpus, args.envs, args.obs, args.rooms, max(1, args.launches_per_step)),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.rows = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
            except (ValueError, IndexError):
                continue
            for nm, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def physical_gpu_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        ids = [x for x in vis.split(",") if x]
        if local_rank < len(ids) and ids[local_rank].isdigit():
            return int(ids[local_rank])
    return local_rank


# ------------------------------------------------------------------------------------------------
# native arm
# ------------------------------------------------------------------------------------------------
def cpu_baseline_subprocess(args):
    """Runs the reference arm (bounded sample) in a fresh process BEFORE any CUDA work so the host cores are idle."""
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "5", "--warmup", "1"]
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE")}
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env).stdout.strip().splitlines()
        return json.loads(out[-1])["cpu_baseline"]
    except Exception as e:  # the bench line must still be printed
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": f"failed: {e!r}"}


def _claim_stdout():
    """Keeps the real stdout for the ONE JSON line; whatever libraries print to fd 1 meanwhile (NCCL's version banner at
    communicator creation, for one) goes to stderr."""
    sys.stdout.flush()
    keep = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return keep


def _reset_leg(args, rank, world, local_rank, barrier, dist, torch, np):
    """C4(i) of SURVEY.md 8(d) / BASELINE config 4: the batched reset.  65 536 group rooms (group seed g, every env of a group shares
    the room: es_manager.py:80-95) split over the ranks by group_id % world, then the group_size = 1 variant (1 Mi distinct rooms
    in all).  Rooms are the reference generator's, seed for seed (tests/test_gpu_sokoban.py::test_generator_reproduces_reference_rooms);
    validity = status 0 of generate_room's four tries; the solvability check is the reference's own BFS (get_shortest_action_path)
    on the device over every valid group room."""
    from ragen_b200.env.sokoban import SokobanBatch, solve_rooms

    dev = torch.device("cuda", local_rank)
    gen = SokobanBatch(1, (H, W), NUM_BOXES, MAX_STEPS, 300, device=local_rank)
    out = {"kernels": ["sokoban_generate_kernel", "sokoban_solve_kernel"], "dim_room": [H, W], "num_boxes": NUM_BOXES, "search_depth": 300}
    for name, total in (("group_rooms", args.reset_groups), ("distinct_rooms", args.reset_distinct)):
        if total <= 0:
            continue
        seeds = np.arange(rank, total, world, dtype=np.int64)  # group g lives on rank g % world
        gen.generate(seeds[:min(seeds.size, 1 << 17)])  # workspace allocation (it grows with the rooms in flight) and first-launch costs stay outside the timed region
        barrier()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        fixed, state, status, explored = gen.generate(seeds)
        t1.record()
        barrier()
        ms = torch.tensor([t0.elapsed_time(t1)], dtype=torch.float64, device=dev)
        ok = status == 0
        stats = torch.tensor([float(ok.sum()), float(explored.sum()), float(seeds.size)], dtype=torch.float64, device=dev)
        mx = torch.tensor([float(explored.max())], dtype=torch.float64, device=dev)
        rec = {}
        if name == "group_rooms":
            fg, sg = fixed[ok].contiguous(), state[ok].contiguous()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record()
            sst, slen, _, _ = solve_rooms(fg, sg, (H, W), NUM_BOXES, max_depth=100, device=local_rank, want_path=False)
            s1.record()
            torch.cuda.synchronize()
            solved = sst == 0
            sol = torch.tensor([float(solved.sum()), float(slen[solved].sum()), s0.elapsed_time(s1)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(sol[:2])
                smax = sol[2:].clone()
                dist.all_reduce(smax, op=dist.ReduceOp.MAX)
                sol[2] = smax[0]
            rec.update({"solver_ms": float(sol[2]), "bfs_solvable": float(sol[0]), "bfs_mean_solution_len": float(sol[1] / max(float(sol[0]), 1.0))})
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            dist.all_reduce(stats)
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        rec.update({"rooms": int(stats[2]), "ms": float(ms), "rooms_per_s": float(stats[2]) / (float(ms) * 1e-3),
                    "valid_fraction": float(stats[0] / stats[2]), "mean_explored": float(stats[1] / stats[2]), "max_explored": int(mx)})
        if "bfs_solvable" in rec:
            rec["bfs_solvable_fraction_of_valid"] = rec.pop("bfs_solvable") / max(float(stats[0]), 1.0)
        out[name] = rec
        del fixed, state, status, explored
    return out


def run_native(args):
    out = _claim_stdout()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cpu_base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_base = cpu_baseline_subprocess(args)

    import numpy as np
    import torch
    import torch.distributed as dist

    from ragen_b200.env.sokoban import SokobanBatch
    from ragen_b200.synth import synthetic_rooms

    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl native needs a B200 (there is no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    from ragen_b200.sharding import bind_host_to_gpu

    numa = bind_host_to_gpu(local_rank) if world > 1 else None  # pinned staging buffers next to the rank's GPU
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local_rank])
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n = args.envs
    fixed, state = synthetic_rooms(n, H, W, NUM_BOXES, seed=4 + rank)
    glyph = args.rooms == "glyph3"
    compact = args.rooms == "nibbles" or glyph  # both step packed rooms and report three-bit rows / deltas
    if compact and args.obs != "packed3":
        raise SystemExit("--rooms nibbles / glyph3 work with three-bit rows: use --obs packed3")
    # Several batches of n envs each, stepped in rotation: ONE batch's lines (58 MB with glyph rooms, 108 MB with nibble rooms) would
    # fit the 126 MB L2, so between two touches of a line the other batches' lines (>= 1.6 x L2 in all) go through it.
    NBATCH = resident_batches(n, args.obs, args.rooms)
    if os.environ.get("RAGEN_BENCH_BATCHES"):  # probing: more batches in rotation (an even divisor of --launches-per-step and of 8)
        NBATCH = int(os.environ["RAGEN_BENCH_BATCHES"])
    batches = []
    for j in range(NBATCH):
        if j:
            fixed, state = synthetic_rooms(n, H, W, NUM_BOXES, seed=1004 + 17 * j + rank)
        b_ = SokobanBatch(n, (H, W), NUM_BOXES, MAX_STEPS, device=local_rank, rooms=args.rooms)
        b_.load_rooms(fixed, state)
        batches.append(b_)
    batch = batches[0]
    RA = 8  # rotating action slabs, all resident in HBM
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)
    actions = torch.randint(1, 5, (RA, n), dtype=torch.uint8, device=dev, generator=g)
    host_actions = actions.cpu().pin_memory()  # the step's inputs live in pinned host memory (contract); step_host copies them up as they are

    sampler = ClockSampler(physical_gpu_index(local_rank))
    if rank == 0:
        sampler.start()

    # ---- device-resident timing -------------------------------------------------------------------
    # One bench "step" = R back-to-back launches (alternating batches), so that K steps are a timed region of tens of
    # milliseconds that host jitter cannot move; the R launches are captured once into a CUDA graph and replayed (no
    # Python / ctypes call per 30 us kernel inside the timed region).  R = 1 reproduces the round-1 line.
    packed = {"packed3": 3, "packed": True, "bytes": False}[args.obs]
    G = H * (W + 1)
    row_bytes = {"packed3": ((3 * H * W + 15) // 16) * 2, "packed": G // 2, "bytes": G}[args.obs]
    R = max(1, args.launches_per_step)

    def one_step(first):
        for i in range(R):
            batches[(first + i) % NBATCH].step(actions[(first + i) % RA], packed=packed)

    one_step(0)  # module load, cudaFuncSetAttribute: before any capture
    torch.cuda.synchronize()
    launch_mode, graph, chains = "loop", None, None
    if not args.no_graph and R % NBATCH == 0 and R % RA == 0:
        try:
            side = torch.cuda.Stream(dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                one_step(0)
            torch.cuda.current_stream(dev).wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                one_step(0)
            launch_mode = "cuda_graph"
            if not args.no_chains:
                # the resident batches are independent jobs: each batch's launches form one chain (its own stream inside the
                # capture), the chains meet again at the end of the step.  Launches of different batches overlap where one kernel
                # drains and the next ramps up; launches of the same batch stay in order.
                chains = torch.cuda.CUDAGraph()
                lanes = [torch.cuda.Stream(dev) for _ in range(NBATCH)]
                with torch.cuda.graph(chains, stream=side):
                    for b_i, lane_s in enumerate(lanes):
                        lane_s.wait_stream(side)
                        with torch.cuda.stream(lane_s):
                            for i in range(b_i, R, NBATCH):
                                batches[i % NBATCH].step(actions[i % RA], packed=packed)
                    for lane_s in lanes:
                        side.wait_stream(lane_s)
        except Exception as e:  # the plain loop is the same GPU work; say which one ran
            graph, chains, launch_mode = None, None, f"loop (graph capture failed: {type(e).__name__})"
            torch.cuda.synchronize()

    def run_step(i):
        if graph is not None:
            graph.replay()
        else:
            one_step(i * R)

    for i in range(args.warmup):
        run_step(i)
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    launches0 = batch.ctx.launches
    t_begin.record()
    for i in range(args.steps):
        run_step(i)
    t_end.record()
    barrier()
    launches = (R * args.steps) if graph is not None else (batch.ctx.launches - launches0)
    total_ms = t_begin.elapsed_time(t_end)
    # the step kernel is the ONLY kernel in the timed region (profiles/*_launches.csv), so its average launch duration is the
    # region divided by the launches; an event pair around every launch would itself cost ~2 us per launch
    kern_ms = total_ms / max(1, launches)
    total_ms = max_over_ranks(total_ms)
    serial = None
    if chains is not None:
        # `value`: the same launches with the NBATCH batches as independent chains.  The region above (one launch after the other)
        # stays the source of roofline.kernel_ms / achieved / frac - a launch's own duration, what ncu's launch list can be held
        # against - and is reported as `serial_chain`
        serial = {"value": n * world * launches / (total_ms * 1e-3), "unit": UNIT, "ms_per_step": total_ms / args.steps}
        for i in range(max(args.warmup, 3)):
            chains.replay()
        barrier()
        t_begin.record()
        for i in range(args.steps):
            chains.replay()
        t_end.record()
        barrier()
        total_ms = max_over_ranks(t_begin.elapsed_time(t_end))
        launch_mode = f"cuda_graph, {NBATCH} per-batch chains (serial_chain / roofline: one chain)"

    # ---- end to end with host buffers ----------------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        e2e_steps = max(args.steps, 20)
        api_tail = "_glyph3" if glyph else "_compact" if compact else {"packed3": "_packed3", "packed": "_packed", "bytes": ""}[args.obs]

        def timed(fn, steps):
            for i in range(max(args.warmup, 3)):
                fn(i)
            barrier()
            t0 = time.perf_counter()
            for i in range(steps):
                fn(i)
            barrier()
            return max_over_ranks(time.perf_counter() - t0)

        # (1) full results per step: glyph rows + reward / done / flags / n_executed, one device->host copy, synchronous
        rows_s = timed(lambda i: batches[i & 1].step_host(host_actions[i % RA], packed=packed), e2e_steps)
        e2e_rows = {"value": n * world * e2e_steps / rows_s, "unit": UNIT, "h2d_bytes_per_step": n * 1,
                    "d2h_bytes_per_step": n * (row_bytes + 8 + 1 + 1 + 1), "ms_per_step": 1e3 * rows_s / e2e_steps,
                    "api": "SokobanBatch.step_host -> ragen_sokoban_step%s (C ABI), pinned host buffers, one device->host copy" % api_tail}
        e2e = dict(e2e_rows, host_numa=numa)
        if compact:
            # (2) the same results as 4-byte step deltas (ragen_sokoban_step_delta): lossless for a consumer that mirrors the rooms
            # (HostMirror / decode_delta rebuild rooms, rows, rewards and flags bit for bit - tests/test_gpu_sokoban.py), two steps in
            # flight (launch i+1 while the results of step i drain on a copy stream)
            tickets = {}

            def run_pipelined(steps, fmt):
                def pipelined(i):
                    tickets[i + 1] = batches[(i + 1) & 1].step_host_begin(host_actions[(i + 1) % RA], fmt)
                    batches[i & 1].step_host_end(tickets.pop(i))

                for w in range(4):  # both batches, both pipeline slots: every pinned / device buffer exists before the clock starts
                    batches[w & 1].step_host_end(batches[w & 1].step_host_begin(host_actions[w % RA], fmt))
                    batches[w & 1].step_host_end(batches[w & 1].step_host_begin(host_actions[w % RA], fmt))
                barrier()
                t0 = time.perf_counter()
                tickets[0] = batches[0].step_host_begin(host_actions[0], fmt)
                for i in range(steps):
                    pipelined(i)
                batches[steps & 1].step_host_end(tickets.pop(steps))  # drain the look-ahead step (not counted)
                barrier()
                return max_over_ranks(time.perf_counter() - t0)

            # wall-clock over 20 short steps: one stray host hiccup (a pinned allocation, a collector pass) is 10x the region, so a
            # leg runs three times and the MEDIAN is the line's number; all three are listed
            def delta_leg(fmt, record_bytes, what):
                runs = sorted(run_pipelined(e2e_steps, fmt) for _ in range(3))
                return {"value": n * world * e2e_steps / runs[1], "unit": UNIT, "h2d_bytes_per_step": n * 1,
                        "d2h_bytes_per_step": n * record_bytes, "ms_per_step": 1e3 * runs[1] / e2e_steps,
                        "ms_per_step_runs": [round(1e3 * t / e2e_steps, 4) for t in runs], "api": what}

            rooms_name = "glyph rooms" if glyph else "nibble rooms"
            delta4 = delta_leg("delta", 4, "SokobanBatch.step_host_begin/_end -> ragen_sokoban_step_%s (C ABI): 4-byte step deltas down "
                               "(player cell before / after, box destination, info byte), self-contained records" % ("glyph3" if glyph else "delta"))
            e2e = delta_leg("delta1", 1, "SokobanBatch.step_host_begin/_end -> ragen_sokoban_step_delta1 (C ABI, %s): pinned host actions up, "
                            "ONE byte per action down (reward code + done / effective / success / executed / box-moved bits), two steps in "
                            "flight; decode_delta rebuilds rewards and flags, HostMirror (which knows the actions it sent and where every "
                            "player stands) the rooms and glyph rows, on the host on demand" % rooms_name)
            # the same one-byte records through SokobanBatch.host_stepper: [actions up | step | records down] captured ONCE per
            # (batch, pinned action buffer) as a CUDA graph, replayed per step - the per-step host work drops from a dozen framework
            # calls to a replay and a stream wait.  One stepper per rotating action slab and batch; two steps in flight.
            eager1 = e2e
            try:
                NBE = len(batches) if RA % len(batches) == 0 else 2  # batches stepped in turn; NBE - 1 steps are in flight behind the one awaited
                steppers = {}
                for i_ in range(RA):
                    steppers[(i_ % NBE, i_ % RA)] = batches[i_ % NBE].host_stepper(host_actions[i_ % RA], record_bytes=1)

                def run_graph(steps):
                    def st(i):
                        return steppers[(i % NBE, i % RA)]
                    for w in range(RA):
                        st(w).launch()
                        st(w).wait()
                    barrier()
                    t0 = time.perf_counter()
                    for i in range(NBE - 1):
                        st(i).launch()
                    for i in range(steps):
                        st(i + NBE - 1).launch()
                        st(i).wait()
                    for i in range(steps, steps + NBE - 1):
                        st(i).wait()  # drain the look-ahead steps (not counted)
                    barrier()
                    return max_over_ranks(time.perf_counter() - t0)

                runs = sorted(run_graph(e2e_steps) for _ in range(3))
                e2e = {"value": n * world * e2e_steps / runs[1], "unit": UNIT, "h2d_bytes_per_step": n * 1, "d2h_bytes_per_step": n * 1,
                       "ms_per_step": 1e3 * runs[1] / e2e_steps, "ms_per_step_runs": [round(1e3 * t / e2e_steps, 4) for t in runs],
                       "api": "SokobanBatch.host_stepper(pinned actions, record_bytes=1).launch() / .wait() -> one CUDA graph per step: "
                              "actions host->device, ragen_sokoban_step_delta1 (C ABI, %s), one byte per action device->host into pinned "
                              "memory; %d batches stepped in turn, %d steps in flight; decode_delta / HostMirror rebuild rewards, flags, rooms "
                              "and glyph rows on the host on demand" % (rooms_name, NBE, NBE),
                       "eager_delta1": eager1}
                if eager1["value"] > e2e["value"]:
                    # many ranks on one host: the link, not the per-step host work, is the limit, and the eager pipeline's separate copy
                    # stream uses it a little better - the line carries the faster of the two public paths and lists the other
                    graph_leg = {k: v for k, v in e2e.items() if k != "eager_delta1"}
                    e2e = dict(eager1, graph_replayed=graph_leg)
            except Exception as ex:  # capture unavailable: the eager leg stays the line's number, and says why
                e2e = dict(eager1, graph_leg_failed=f"{type(ex).__name__}: {ex}"[:200])
            e2e["host_numa"] = numa
            e2e["delta4"] = delta4
            e2e["full_rows"] = e2e_rows

        if compact:
            # (3) the same step with the observations as TEXT in host memory: the three-bit rows the step wrote are rendered to
            # UTF-8 strings on the device (ragen_glyph_utf8; fixed-width NUL-padded rows, one numpy 'S' view away from Python
            # bytes objects) and those travel instead of the codes
            from ragen_b200.env.sokoban import glyph_table  # noqa: F401

            lookup = {0: "#", 1: "_", 2: "O", 3: "√", 4: "X", 5: "P", 6: "S"}  # ragen/env/sokoban/config.py:10
            text_bufs = [b_.text_rows(lookup, fmt=3) for b_ in batches]
            stride = int(text_bufs[0][0].shape[1])
            pinned_text = [torch.empty((n, stride), dtype=torch.uint8, pin_memory=True) for _ in batches]
            pinned_len = [torch.empty(n, dtype=torch.int32, pin_memory=True) for _ in batches]

            def text_step(i):
                b_ = batches[i & 1]
                d_act = host_actions[i % RA].to(dev, non_blocking=True).unsqueeze(1)
                b_.step(d_act, packed=3)
                rows_d, lens_d = b_.text_rows(lookup, out=text_bufs[i & 1], fmt=3)
                pinned_text[i & 1].copy_(rows_d, non_blocking=True)
                pinned_len[i & 1].copy_(lens_d, non_blocking=True)
                h = b_._pin("slab_sc", (11 * n,), torch.uint8)
                h.copy_(b_._slab[b_._off_sc: b_._off_sc + 11 * n], non_blocking=True)
                torch.cuda.current_stream(dev).synchronize()

            text_s = timed(text_step, e2e_steps)
            e2e["text"] = {"value": n * world * e2e_steps / text_s, "unit": UNIT, "h2d_bytes_per_step": n * 1,
                           "d2h_bytes_per_step": n * (stride + 4 + 11), "ms_per_step": 1e3 * text_s / e2e_steps,
                           "api": "SokobanBatch.step (compact) + text_rows -> ragen_glyph_utf8 (C ABI): observations as UTF-8 text rows "
                                  "(%d B per env, the reference's glyphs) + lengths + reward / done / flags in pinned host memory" % stride}

    reset = None
    if not args.no_reset:
        reset = _reset_leg(args, rank, world, local_rank, barrier, dist, torch, np)

    clocks = sampler.stop() if rank == 0 else None
    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        kname = "sokoban_tile_kernel<10,10,false,%d>" % (4 if glyph else 3 if compact else {"packed3": 2, "packed": 1, "bytes": 0}[args.obs])
        moved = MOVED_PER_STEP - G + row_bytes - (H * W if compact else 0)
        if glyph:
            moved = 2 * row_bytes + 2 * 8 + 1 + 11  # glyph room in and out (it is the observation), record r+w, action, results
        achieved = moved * n / (kern_ms * 1e-3) / 1e9
        traffic, traffic_src = None, None
        tpath = os.path.join(ROOT, "profiles", TRAFFIC_FILE)
        if os.path.exists(tpath):
            rec = json.load(open(tpath)).get(kname)
            if rec:
                traffic, traffic_src = rec["dram_bytes_per_launch"], "profiles/%s (%s)" % (TRAFFIC_FILE, rec.get("how", ""))
        line = {"metric": METRIC, "value": n * world * launches / (total_ms * 1e-3), "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": workload_config(world, n, args.obs, args.rooms, R), "launch_mode": launch_mode,
                "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "serial_chain": serial,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": traffic, "traffic_source": traffic_src, "kernel": kname, "kernel_ms": kern_ms,
                             "bytes_per_env_step": moved, "peak_source": peak_src,
                             "frac_survey_431B": BYTES_PER_STEP * n / (kern_ms * 1e-3) / 1e9 / peak,
                             "achieved_with_overlap": moved * n * launches / (total_ms * 1e-3) / 1e9,  # per GPU: n and launches are a rank's
                             "frac_with_overlap": moved * n * launches / (total_ms * 1e-3) / 1e9 / peak,
                             "note": "kernel_ms / achieved / frac: launches one after the other (serial_chain) - a launch's own duration; "
                                     "achieved_with_overlap / frac_with_overlap: the bytes of `value`'s region, where launches of independent "
                                     "batches overlap at their ends (above 1.0 of the torch copy_ figure the peak was measured with; the same "
                                     "with 8 batches = 464 MB in rotation, so not L2).  "
                                     "achieved / frac count the bytes this kernel has to move per env-step (glyph rooms: room in and out 76 - the room is the "
                                     "observation - record r+w 16, action 1, results 11 = 104; nibble rooms: 100 + 16 + 1 + 11 + three-bit rows 38 = 166; 266 / 283 / "
                                     "338 B with byte rooms and three-bit / nibble / byte rows).  frac_survey_431B applies SURVEY.md 8(d)'s figure, "
                                     "which charges a 100 B room_fixed read and 110 B of glyph bytes this kernel does not move: it is "
                                     "not a physical fraction"}}
        if reset is not None:
            line["reset"] = reset
        if cpu_base is not None:
            line["cpu_baseline"] = cpu_base
        print(json.dumps(line), file=out, flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    _args = parse_args()
    if _args.rows:
        import bench_rows

        bench_rows.main(_args.quick, _args.only)
        sys.exit(0)
    sys.exit(run_reference(_args) if _args.impl == "reference" else run_native(_args))
