"""DRAM bytes per launch of the bench's step kernel from an ncu launch list.

    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --cache-control none --clock-control none \
        --csv --log-file launches.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-reset --no-cpu-baseline --no-graph
    python scripts/traffic_from_ncu.py launches.csv 'sokoban_tile_kernel<10,10,false,4>' 40 [n_envs] >> entry for profiles/traffic_r2.json

The first `skip` launches are left out (cold L2); the rest is the steady state of the rotating batches, where the write-backs a
launch leaves dirty in L2 are counted in the launches that follow."""
import csv
import json
import re
import sys


def main(path, kernel, skip, n_envs):
    want = re.sub(r"[\s()]|int|bool", "", kernel)
    per = {}
    with open(path) as f:
        rows = [r for r in csv.reader(l for l in f if l.startswith('"'))]
    hdr = rows[0]
    ix = {h: i for i, h in enumerate(hdr)}
    for r in rows[1:]:
        name = re.sub(r"[\s()]|int|bool|void|<unnamed>::|\(anonymous namespace\)::", "", r[ix["Kernel Name"]])
        if want.split("<")[0] not in name:
            continue
        tmpl = re.sub(r"false", "0", re.sub(r"true", "1", want))
        flat = re.sub(r"false", "0", re.sub(r"true", "1", name))
        if tmpl.split("<")[1].rstrip(">") not in flat:
            continue
        d = per.setdefault(r[ix["ID"]], {})
        v, unit = float(r[ix["Metric Value"]].replace(",", "")), r[ix["Metric Unit"]]
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3}.get(unit, 1)
        d[r[ix["Metric Name"]]] = v * scale
    launches = [per[k] for k in sorted(per, key=int)][skip:]
    n = len(launches)
    rd = sum(x["dram__bytes_read.sum"] for x in launches) / n
    wr = sum(x["dram__bytes_write.sum"] for x in launches) / n
    us = sum(x["gpu__time_duration.sum"] for x in launches) / n
    print(json.dumps({kernel: {"dram_bytes_per_launch": rd + wr, "dram_read_bytes_per_launch": rd, "dram_write_bytes_per_launch": wr,
                               "launches_averaged": n, "mean_kernel_us_under_ncu": us, "bytes_per_env_step": (rd + wr) / n_envs}}, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 40, int(sys.argv[4]) if len(sys.argv) > 4 else 1 << 20)
