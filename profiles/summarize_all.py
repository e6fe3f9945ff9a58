#!/usr/bin/env python
"""profiles/summarize_all.py <tag> — reads gpurun_out/all_<tag>.csv (the raw page of the `ncu --set full` pass of profiles/capture_all.sh:
one row per launch) and writes profiles/<tag>_all_kernels.md (one row per KERNEL: its longest launch) and profiles/<tag>_traffic.json
(per kernel: DRAM bytes per launch, lanes per instruction, issue-active %, L1 wavefront pipe % — bench.py's roofline.traffic)."""
import csv, json, os, re, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TAG = sys.argv[1] if len(sys.argv) > 1 else "r02"
import glob
# all_<tag>.csv, plus all_<tag>_*.csv from filtered passes over the same probe (profiles/capture_all.sh <tag> <kernel regex>): rows are
# re-ordered to the first file's columns
hdr = units = None
rows = []
for path in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", f"all_{TAG}.csv")) + glob.glob(os.path.join(ROOT, "gpurun_out", f"all_{TAG}_*.csv"))):
    part = list(csv.reader(open(path)))
    if len(part) < 3:
        continue
    if hdr is None:
        hdr, units = part[0], part[1]
        rows += part[2:]
    else:
        pos = {h: i for i, h in enumerate(part[0])}
        rows += [[r[pos[h]] if h in pos and pos[h] < len(r) else "" for h in hdr] for r in part[2:]]
idx = {h: i for i, h in enumerate(hdr)}
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "nsecond": 1e-3, "ns": 1e-3, "usecond": 1, "us": 1, "msecond": 1e3, "ms": 1e3, "second": 1e6}


def val(r, m, default=0.0):
    if m not in idx or r[idx[m]] in ("", "n/a"):
        return default
    return float(r[idx[m]].replace(",", "")) * SCALE.get(units[idx[m]], 1)


def short(name):
    name = name.replace("void ", "").replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
    return re.sub(r"\(.*$", "", name)


best = {}
count = {}
for r in rows:
    k = short(r[idx["Kernel Name"]])
    if k.startswith("at::") or "at::" in k or k.startswith("elementwise_kernel") or "vectorized_elementwise" in k:
        continue                                   # torch's own kernels of the probe (random inputs, copies): not this library's
    count[k] = count.get(k, 0) + 1
    if k not in best or val(r, "gpu__time_duration.sum") > val(best[k], "gpu__time_duration.sum"):
        best[k] = r
out = [f"# ncu --set full: every kernel of libclrt_cuda.so ({TAG})", "",
       "`bash profiles/capture_all.sh %s` = `ncu --set full --clock-control none python profiles/all_kernels_probe.py` (Soup-1M, 1920x1080, "
       "2^24-element blocks; one GPU).  One row per kernel: its longest launch of the pass.  Times under a profiler are cold-cache and serialised — they "
       "explain, bench.py measures.  lanes = thread instructions per warp instruction; L1 pipe = l1tex__data_pipe_lsu_wavefronts (%% of peak); "
       "stalls = warps stalled per issue-active cycle, top three." % TAG, "",
       "| kernel | launches | time us | grid x block | regs | smem B | DRAM rd MB | DRAM wr MB | DRAM % | SM % | occ % | issue % | L1 hit % | L2 hit % | lanes | L1 pipe % | local ld/st | top stalls |",
       "|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
facts = {}
for k, r in sorted(best.items(), key=lambda kv: -val(kv[1], "gpu__time_duration.sum")):
    stalls = sorted(((h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")], val(r, h)) for h in idx
                     if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")), key=lambda x: -x[1])[:3]
    smem = val(r, "launch__shared_mem_per_block_static") + val(r, "launch__shared_mem_per_block_dynamic")
    rd, wr = val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum")
    out.append("| `%s` | %d | %.1f | %d x %d | %d | %d | %.2f | %.2f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %d / %d | %s |" % (
        k, count[k], val(r, "gpu__time_duration.sum"), val(r, "launch__grid_size"), val(r, "launch__block_size"), val(r, "launch__registers_per_thread"), smem,
        rd / 1e6, wr / 1e6, val(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"), val(r, "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
        val(r, "sm__warps_active.avg.pct_of_peak_sustained_active"), val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        val(r, "l1tex__t_sector_hit_rate.pct"), val(r, "lts__t_sector_hit_rate.pct"), val(r, "smsp__thread_inst_executed_per_inst_executed.ratio"),
        val(r, "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"), val(r, "sass__inst_executed_local_loads"), val(r, "sass__inst_executed_local_stores"),
        ", ".join("%s %.2f" % s for s in stalls)))
    base = re.sub(r"<.*$", "", k)
    f = {"dram_bytes": rd + wr, "threads_per_inst": val(r, "smsp__thread_inst_executed_per_inst_executed.ratio"),
         "issue_active_pct": val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
         "l1_pipe_pct": val(r, "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"), "time_us": val(r, "gpu__time_duration.sum"), "kernel": k}
    if "soup1m_" + base not in facts or f["time_us"] > facts["soup1m_" + base]["time_us"]:
        facts["soup1m_" + base] = f
open(os.path.join(ROOT, "profiles", f"{TAG}_all_kernels.md"), "w").write("\n".join(out) + "\n")
json.dump(facts, open(os.path.join(ROOT, "profiles", f"{TAG}_traffic.json"), "w"), indent=1, sort_keys=True)
print("\n".join(out[:6]) + "\n... %d kernels" % len(best))
