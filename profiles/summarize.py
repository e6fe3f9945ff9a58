#!/usr/bin/env python
"""profiles/summarize.py <tag> — reads gpurun_out/full_*_<tag>.ncu-rep (ncu --set full) and gpurun_out/launches_<tag>.csv here
(no GPU needed) and writes profiles/<tag>_kernels.md + profiles/<tag>_launches.csv (per-kernel totals of the launch list)."""
import csv
import glob
import io
import os
import subprocess
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TAG = sys.argv[1] if len(sys.argv) > 1 else "r01"
METRICS = [
    ("gpu__time_duration.sum", "time"),
    ("launch__grid_size", "grid"), ("launch__block_size", "block"), ("launch__registers_per_thread", "regs"),
    ("launch__shared_mem_per_block_static", "smem_static"), ("launch__shared_mem_per_block_dynamic", "smem_dyn"),
    ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_pct"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"),
    ("l1tex__t_sector_hit_rate.pct", "l1_hit_pct"), ("lts__t_sector_hit_rate.pct", "l2_hit_pct"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "threads_per_inst"),
    ("smsp__cycles_active.avg", "smsp_cycles_active"), ("sm__cycles_elapsed.max", "sm_cycles_elapsed"),
    ("sass__inst_executed_local_loads", "local_loads"), ("sass__inst_executed_local_stores", "local_stores"),
]


def raw(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    return hdr, units, rows[2:]


def main():
    lines = [f"# ncu --set full summaries ({TAG})", "",
             "Commands: `python bench.py --steps 2 --warmup 3 --no-cpu-baseline --workload <w>` and `python profiles/sort_probe.py 24` (profiles/capture.sh), `--clock-control none`; section names carry the workload.",
             "Numbers under a profiler are cold-cache and serialised; they explain, the bench line measures.", ""]
    for path in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", f"full_*_{TAG}.ncu-rep"))):
        hdr, units, rows = raw(path)
        idx = {h: i for i, h in enumerate(hdr)}
        name = rows[0][idx["Kernel Name"]].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
        tag = os.path.basename(path)[len("full_"):].split("_")[0]
        lines.append(f"## [{tag}] {name}  ({len(rows)} launches captured)")
        lines.append("")
        lines.append("| metric | " + " | ".join(f"launch {k}" for k in range(len(rows))) + " | unit |")
        lines.append("|---|" + "---|" * (len(rows) + 1))
        for m, label in METRICS:
            if m in idx:
                lines.append(f"| {label} (`{m}`) | " + " | ".join(r[idx[m]] for r in rows) + f" | {units[idx[m]]} |")
        # top stall reasons
        stalls = [(h, [float(r[i] or 0) for r in rows]) for h, i in idx.items()
                  if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
        stalls.sort(key=lambda x: -x[1][0])
        lines.append("")
        lines.append("Top stall reasons (warps stalled per issue-active cycle): " +
                     ", ".join(f"{h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]}={v[0]:.2f}" for h, v in stalls[:5]))
        lines.append("")
    open(os.path.join(ROOT, "profiles", f"{TAG}_kernels.md"), "w").write("\n".join(lines) + "\n")
    # launch lists (one per workload) -> per-kernel totals
    for lp in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", f"launches_*_{TAG}.csv"))):
        wl = os.path.basename(lp)[len("launches_"):-len(f"_{TAG}.csv")]
        text = open(lp).read()
        start = text.index('"ID"')
        rows = list(csv.DictReader(io.StringIO(text[start:])))
        tot = defaultdict(lambda: [0, 0.0])
        for r in rows:
            if r.get("Metric Name") != "gpu__time_duration.sum":
                continue
            k = r["Kernel Name"].split("(")[0]
            v = float(r["Metric Value"].replace(",", ""))
            if r["Metric Unit"] in ("ns", "nsecond"):
                v /= 1e3
            elif r["Metric Unit"] in ("ms", "msecond"):
                v *= 1e3
            tot[k][0] += 1; tot[k][1] += v
        total = sum(v[1] for v in tot.values())
        with open(os.path.join(ROOT, "profiles", f"{TAG}_launches_{wl}.csv"), "w") as f:
            f.write("kernel,launches,total_us,share_pct,avg_us\n")
            for k, (c, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
                f.write(f"\"{k}\",{c},{t:.1f},{100 * t / total:.2f},{t / c:.2f}\n")
        import shutil
        shutil.copy(lp, os.path.join(ROOT, "profiles", f"{TAG}_launch_list_raw_{wl}.csv"))
    # DRAM traffic per launch of the captured kernels -> profiles/<tag>_traffic.json (bench.py's roofline.traffic)
    import json
    traffic = {}
    for path in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", f"full_*_{TAG}.ncu-rep"))):
        hdr, units, rows = raw(path)
        idx = {h: i for i, h in enumerate(hdr)}
        def val(m, r):
            v = float(r[idx[m]].replace(",", "")); u = units[idx[m]]
            return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        name = os.path.basename(path)[len("full_"):-len(f"_{TAG}.ncu-rep")]
        traffic[name] = val("dram__bytes_read.sum", rows[0]) + val("dram__bytes_write.sum", rows[0])
    json.dump(traffic, open(os.path.join(ROOT, "profiles", f"{TAG}_traffic.json"), "w"), indent=1, sort_keys=True)
    print(open(os.path.join(ROOT, "profiles", f"{TAG}_kernels.md")).read()[:200])


if __name__ == "__main__":
    main()
