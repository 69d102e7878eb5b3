"""CPU: the reference arm of bench.py runs without a GPU and prints ONE JSON line with the contract's keys; the native arm refuses to
run without a GPU (no CPU fallback)."""
import json
import os
import subprocess
import sys

import torch

from conftest import ROOT

BENCH = os.path.join(ROOT, "bench.py")


def test_reference_arm_line():
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE")}
    res = subprocess.run([sys.executable, BENCH, "--impl", "reference", "--steps", "2", "--warmup", "1", "--ref-envs", "256"],
                         capture_output=True, text=True, timeout=300, env=env)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [l for l in res.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["metric"] == "sokoban_env_steps_per_sec" and d["unit"] == "env-steps/s"
    assert d["steps"] == 2 and d["warmup"] == 1 and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["config"]["workload"].startswith("sokoban_10x10_2box") and d["cpu_baseline"]["kind"] == "port"
    assert d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"] > 0
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0


def test_reference_arm_other_ranks_are_silent():
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    res = subprocess.run([sys.executable, BENCH, "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=120, env=env)
    assert res.returncode == 0 and res.stdout.strip() == ""


def test_native_arm_needs_a_gpu():
    if torch.cuda.is_available():
        return
    res = subprocess.run([sys.executable, BENCH, "--steps", "1", "--warmup", "0", "--no-cpu-baseline", "--envs", "256"],
                         capture_output=True, text=True, timeout=300)
    assert res.returncode != 0 and res.stdout.strip() == ""
    assert "no CPU fallback" in res.stderr or "CUDA" in res.stderr


def test_committed_native_line_is_self_consistent():
    """The last native bench line committed under profiles/ carries every key of the contract and its numbers agree with each other
    (value = env-steps of the region / its time, roofline.frac = achieved / peak from the serial chain, e2e bytes per step)."""
    path = os.path.join(ROOT, "profiles", "r30_bench.json")
    d = json.loads(open(path).read().strip().splitlines()[-1])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype",
              "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"):
        assert k in d, k
    n, launches = d["config"]["envs_per_gpu"], d["gpu_launches"]
    assert d["metric"] == "sokoban_env_steps_per_sec" and d["scaling"] == "weak" and d["dtype"] == "u8" and d["vs_baseline"] is None
    assert launches == d["config"]["launches_per_step"] * d["steps"] > 0
    assert abs(d["value"] - n * d["n_gpus"] * launches / (d["ms_per_step"] * d["steps"] * 1e-3)) <= 1e-6 * d["value"]
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert abs(r["achieved"] - r["bytes_per_env_step"] * n / (r["kernel_ms"] * 1e-3) / 1e9) <= 1e-6 * r["achieved"]
    assert 0.5 < r["frac"] <= 1.0 and r["traffic"] is not None and 0.9 < r["traffic"] / (r["bytes_per_env_step"] * n) < 1.1
    s = d["serial_chain"]
    assert abs(s["value"] - n * d["n_gpus"] / (r["kernel_ms"] * 1e-3)) <= 2e-3 * s["value"]  # the same region, per launch
    e = d["e2e"]
    assert e["unit"] == d["unit"] and e["h2d_bytes_per_step"] == n and e["d2h_bytes_per_step"] == n and 0 < e["value"] < d["value"]
    assert d["clocks"]["reasons"] == [] and d["clocks"]["sm_mhz"] <= d["clocks"]["sm_max_mhz"]
    c = d["cpu_baseline"]
    assert c["kind"] == "port" and c["cores"] >= 1 and c["value"] > 0 and c["unit"] == d["unit"]
