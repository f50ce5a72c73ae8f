"""Turn one gpurun scratch directory (bench.json, launches.csv, *.ncu-rep) into the tracked summaries under profiles/.

    python tools/summarise_profiles.py gpurun_out/h1 r01

Writes profiles/<tag>_bench_n1.json, <tag>_launches.csv, <tag>_launch_list_summary.md and, for every *.ncu-rep in the
directory, profiles/<tag>_<name>_ncu.json (a fixed set of metrics read with `ncu -i ... --page raw --csv`).
"""
from __future__ import annotations

import csv
import io
import json
import os
import shutil
import subprocess
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

STAGE_OF = {
    "tim_graph": "graph", "graph_prep": "graph", "graph_tc": "graph", "graph_fix": "graph", "degree": "degree", "heuristic": "heuristic", "prune": "prune_compact",
    "compact": "prune_compact", "bnb": "exact_search", "canon": "canonical", "rotation": "rotation",
    "translation": "translation", "gather": "rotation", "tail_kernel": "tail_fused",
}

NCU_KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
]


def short(name: str) -> str:
    n = name.split("(")[0]
    n = n.replace("void ", "")
    return n.split("<")[0].strip()


def read_launches(path: str):
    rows = []
    with open(path, newline="") as f:
        text = f.read()
    start = text.find('"ID"')
    for r in csv.DictReader(io.StringIO(text[start:])):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        ms = v / 1e6 if unit in ("ns", "nsecond") else v / 1e3 if unit in ("us", "usecond") else v
        rows.append((short(r["Kernel Name"]), r["Grid Size"], ms))
    return rows


def launch_summary(run_dir: str, tag: str, bench: dict) -> str:
    rows = read_launches(os.path.join(run_dir, "launches.csv"))
    batch = bench["config"].get("batch_per_gpu", 64)
    per = defaultdict(list)
    for name, grid, ms in rows:
        if not name.startswith("b2r::"):
            continue
        g = [int(x) for x in grid.strip("()").split(",")]
        if batch in g[:2] or name.endswith("prune_kernel") or name.endswith("bnb_kernel"):
            per[name].append(ms)
    stage = bench.get("stage_ms_per_step", {})
    total = sum(sum(v) / len(v) for v in per.values())
    lines = [
        f"# {tag} launch list (ncu --metrics gpu__time_duration.sum --clock-control none), `python bench.py --steps 2 --warmup 3 --no-cpu-baseline`",
        "",
        f"Raw list: `profiles/{tag}_launches.csv` ({len(rows)} launches of the command, torch fill kernels of the L2 flush included).",
        f"Per kernel: mean over its batch-{batch} launches.  Per-launch times under ncu are cold-cache and serialised, so compare SHARES with the",
        f"CUDA-event stage times of the plain run (`profiles/{tag}_bench_n1.json`, `stage_ms_per_step`, total {stage.get('total')} ms).",
        "",
        "| kernel | launches | ms per launch (ncu) | share (ncu) | plain-run stage (CUDA events) | share (plain) |",
        "|---|---|---|---|---|---|",
    ]
    st_total = stage.get("total") or 1.0
    for name, v in sorted(per.items(), key=lambda kv: -sum(kv[1]) / len(kv[1])):
        m = sum(v) / len(v)
        key = next((s for k, s in STAGE_OF.items() if k in name), None)
        sv = stage.get(key) if key else None
        lines.append(f"| `{name}` | {len(v)} | {m:.4f} | {100 * m / total:.1f}% | {key}: {sv} ms | "
                     + (f"{100 * sv / st_total:.1f}%" if sv is not None else "-") + " |")
    lines += ["", f"Sum over kernels (ncu): {total:.3f} ms per step; plain run: {bench['ms_per_step']:.4f} ms per step (device events), "
              f"value {bench['value']:.0f} {bench['unit']}, e2e {bench['e2e']['value']:.0f}."]
    return "\n".join(lines) + "\n"


def ncu_raw(rep: str) -> dict:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rd = list(csv.reader(io.StringIO(out)))
    hdr, units, vals = rd[0], rd[1], rd[2]
    d = {}
    for h, u, v in zip(hdr, units, vals):
        d[h] = (v, u)
    return d


def ncu_summary(rep: str) -> dict:
    raw = ncu_raw(rep)
    res = {"kernel": raw.get("Kernel Name", ("?", ""))[0], "report": os.path.basename(rep),
           "how": "ncu --set full --clock-control none --import-source on; read with ncu -i ... --page raw --csv"}
    for k in NCU_KEYS:
        if k in raw:
            v, u = raw[k]
            try:
                res[k] = float(v.replace(",", ""))
            except ValueError:
                res[k] = v
            if u:
                res[k + ".unit"] = u
    if "dram__bytes_read.sum" in res and "dram__bytes_write.sum" in res:
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        r = res["dram__bytes_read.sum"] * scale.get(res.get("dram__bytes_read.sum.unit", "byte"), 1)
        w = res["dram__bytes_write.sum"] * scale.get(res.get("dram__bytes_write.sum.unit", "byte"), 1)
        res["traffic_bytes_per_launch"] = int(r + w)
    return res


def main() -> None:
    run_dir, tag = sys.argv[1], sys.argv[2]
    prof = os.path.join(ROOT, "profiles")
    bench = json.loads(open(os.path.join(run_dir, "bench.json")).read().strip().splitlines()[-1])
    with open(os.path.join(prof, f"{tag}_bench_n1.json"), "w") as f:
        f.write(json.dumps(bench) + "\n")
    ref = os.path.join(run_dir, "bench_ref.json")
    if os.path.exists(ref) and os.path.getsize(ref):
        shutil.copy(ref, os.path.join(prof, f"{tag}_bench_reference_arm.json"))
    if os.path.exists(os.path.join(run_dir, "launches.csv")):
        shutil.copy(os.path.join(run_dir, "launches.csv"), os.path.join(prof, f"{tag}_launches.csv"))
        with open(os.path.join(prof, f"{tag}_launch_list_summary.md"), "w") as f:
            f.write(launch_summary(run_dir, tag, bench))
    for fn in sorted(os.listdir(run_dir)):
        if fn.endswith(".ncu-rep"):
            s = ncu_summary(os.path.join(run_dir, fn))
            with open(os.path.join(prof, f"{tag}_{fn[:-8]}_ncu.json"), "w") as f:
                json.dump(s, f, indent=1)
                f.write("\n")
            print(fn, "->", s.get("kernel"), s.get("gpu__time_duration.sum"), s.get("traffic_bytes_per_launch"))


if __name__ == "__main__":
    main()
