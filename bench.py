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
    ap.add_argument("--obs", default="packed3", choices=["packed3", "packed", "bytes"],
                    help="glyph-row format of the render: two 4-bit codes per byte (what the managers consume) or one code per byte")
    ap.add_argument("--rooms", default="nibbles", choices=["nibbles", "bytes"],
                    help="how the rooms rest in HBM: one cell per byte, or two (ragen_sokoban_step_compact; needs --obs packed3)")
    ap.add_argument("--rows", action="store_true", help="per-kernel lines for every SURVEY.md 8(d) config instead of the contract line")
    ap.add_argument("--quick", action="store_true", help="with --rows: smaller batches")
    ap.add_argument("--only", default="", help="with --rows: comma list of sokoban,generator,frozenlake,parse,c5,manager,rollout")
    return ap.parse_args()


def workload_config(n_gpus, envs_per_gpu, obs="packed", rooms="bytes"):
    rows = {"packed3": "three bits per cell, no separators (38 B per env; ragen_sokoban_step_packed3)",
            "packed": "two 4-bit glyph codes per byte (55 B per env)", "bytes": "one glyph code per byte (110 B per env)"}[obs]
    return {"glyph_rows": rows, "rooms_in_hbm": "two cells per byte (ragen_sokoban_step_compact)" if rooms == "nibbles" else "one cell per byte",
            "workload": "sokoban_10x10_2box_step_render_reward", "dim_room": [H, W], "num_boxes": NUM_BOXES,
            "max_steps": MAX_STEPS, "envs_per_gpu": envs_per_gpu, "envs_total": envs_per_gpu * n_gpus,
            "actions_per_launch": 1, "rooms": "synthetic random-walk rooms (ragen_b200/synth.py)",
            "actions": "uniform 1..4", "l2": "inputs larger than L2: two batches of envs_per_gpu envs stepped alternately, %d MB of distinct lines between two touches of a line (L2 is 126 MB)" % ((envs_per_gpu * 2 * ((50 if rooms == "nibbles" else 100) + 8 + 1 + 11 + {"packed3": 38, "packed": 55, "bytes": 110}[obs])) >> 20),
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


def _ref_step(_):
    envs = _W["envs"]
    acts = _W["rng"].integers(1, 5, len(envs)).tolist()
    t0 = time.perf_counter()
    tot = 0.0
    for e, a in zip(envs, acts):  # es_manager.py:149-169: serial loop, one env.step (with text render) each
        obs, r, d, info = e.step(a)
        tot += r
    return time.perf_counter() - t0, tot


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

        def one_step():
            t0 = time.perf_counter()
            for r in [p.apply_async(_ref_step, (0,)) for p in pools]:
                r.get()
            return time.perf_counter() - t0

        for _ in range(args.warmup):
            one_step()
        times = [one_step() for _ in range(args.steps)]
    finally:
        for p in pools:
            p.terminate()
    total = sum(times)
    n = per_core * cores
    value = n * args.steps / total
    sample = f"{n} envs ({per_core}/core x {cores} procs) x {args.steps} steps, 1 action each, text render per step"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": workload_config(args.gpus, args.envs, args.obs, args.rooms),
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

    n = args.envs
    fixed, state = synthetic_rooms(n, H, W, NUM_BOXES, seed=4 + rank)
    compact = args.rooms == "nibbles"
    if compact and args.obs != "packed3":
        raise SystemExit("--rooms nibbles renders three-bit rows: use --obs packed3")
    # Two batches of n envs each, stepped alternately: with the rooms as nibbles ONE batch's lines (rooms 52 MB, records 8 MB, rows
    # 40 MB, results 12 MB) would fit the 126 MB L2, so between two touches of a line the other batch's ~113 MB go through it.
    batches = []
    for j in range(2):
        if j:
            fixed, state = synthetic_rooms(n, H, W, NUM_BOXES, seed=1004 + rank)
        b_ = SokobanBatch(n, (H, W), NUM_BOXES, MAX_STEPS, device=local_rank, compact=compact)
        b_.load_rooms(fixed, state)
        batches.append(b_)
    batch = batches[0]
    R = 8  # rotating action slabs, all resident in HBM
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)
    actions = torch.randint(1, 5, (R, n), dtype=torch.uint8, device=dev, generator=g)
    host_actions = actions.cpu().pin_memory()  # the step's inputs live in pinned host memory (contract); step_host copies them up as they are

    sampler = ClockSampler(physical_gpu_index(local_rank))
    if rank == 0:
        sampler.start()

    # ---- device-resident timing -------------------------------------------------------------------
    packed = {"packed3": 3, "packed": True, "bytes": False}[args.obs]
    G = H * (W + 1)
    row_bytes = {"packed3": ((3 * H * W + 15) // 16) * 2, "packed": G // 2, "bytes": G}[args.obs]
    for i in range(args.warmup):
        batches[i & 1].step(actions[i % R], packed=packed)
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    launches0 = batch.ctx.launches
    t_begin.record()
    for i in range(args.steps):
        batches[i & 1].step(actions[i % R], packed=packed)
    t_end.record()
    barrier()
    launches = batch.ctx.launches - launches0
    total_ms = t_begin.elapsed_time(t_end)
    # the step kernel is the ONLY kernel in the timed region (gpu_launches == steps; profiles/*_launches.csv), so its average
    # launch duration is the region divided by the launches; an event pair around every launch would itself cost ~2 us per step
    kern_ms = total_ms / max(1, launches)
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())

    # ---- end to end with host buffers ----------------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        for i in range(max(args.warmup, 3)):
            batches[i & 1].step_host(host_actions[i % R], packed=packed)
        barrier()
        t0 = time.perf_counter()
        for i in range(args.steps):
            batches[i & 1].step_host(host_actions[i % R], packed=packed)
        barrier()
        e2e_s = time.perf_counter() - t0
        t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
        e2e = {"value": n * world * args.steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": n * 1,
               "d2h_bytes_per_step": n * (row_bytes + 8 + 1 + 1 + 1), "ms_per_step": 1e3 * e2e_s / args.steps,
               "api": "SokobanBatch.step_host -> ragen_sokoban_step%s (C ABI), pinned host buffers, one device->host copy" % ("_compact" if compact else {"packed3": "_packed3", "packed": "_packed", "bytes": ""}[args.obs]), "host_numa": numa}

    clocks = sampler.stop() if rank == 0 else None
    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = float(json.load(open(peaks_path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        achieved = BYTES_PER_STEP * n / (kern_ms * 1e-3) / 1e9
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            traffic = json.load(open(tpath)).get("sokoban_step_compact_bytes_per_launch" if compact else
                                                 {"packed3": "sokoban_step_packed3_bytes_per_launch", "packed": "sokoban_step_packed_bytes_per_launch",
                                                  "bytes": "sokoban_step_bytes_per_launch"}[args.obs])
        moved = MOVED_PER_STEP - G + row_bytes - (H * W if compact else 0)
        line = {"metric": METRIC, "value": n * world * args.steps / (total_ms * 1e-3), "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": workload_config(world, n, args.obs, args.rooms), "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": traffic, "kernel": "sokoban_tile_kernel<10,10,false,%d>" % (3 if compact else {"packed3": 2, "packed": 1, "bytes": 0}[args.obs]), "kernel_ms": kern_ms,
                             "bytes_per_env_step": BYTES_PER_STEP, "peak_source": peak_src,
                             "bytes_moved_per_env_step": moved, "frac_moved": moved * n / (kern_ms * 1e-3) / 1e9 / peak,
                             "note": "frac uses SURVEY.md 8(d)'s 431 B per env-step, which charges a 100 B room_fixed read the kernel "
                                     "avoids and 110 B of glyph rows; frac_moved uses the bytes this kernel has to move "
                                     "(338 B with byte glyph rows, 283 B with nibble rows, 266 B with three-bit rows, 166 B with the rooms kept as nibbles too)"}}
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
