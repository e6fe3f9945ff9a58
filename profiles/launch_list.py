"""profiles/launch_list.py <tag> — per-kernel totals of the ncu launch lists gpurun_out/launches_<workload>_<tag>.csv
(`ncu --metrics gpu__time_duration.sum --clock-control none --csv` over a bench.py run) -> profiles/<tag>_launches_<workload>.csv
(kernel, launches, total us, share of the run's kernel time, average us).  Runs here, no GPU needed."""
import csv, glob, io, os, sys
from collections import defaultdict
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TAG = sys.argv[1] if len(sys.argv) > 1 else "r02"
for lp in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", f"launches_*_{TAG}.csv"))):
    wl = os.path.basename(lp)[len("launches_"):-len(f"_{TAG}.csv")]
    text = open(lp).read()
    rows = list(csv.DictReader(io.StringIO(text[text.index('"ID"'):])))
    tot = defaultdict(lambda: [0, 0.0])
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        k = r["Kernel Name"].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
        v = float(r["Metric Value"].replace(",", ""))
        v = v / 1e3 if r["Metric Unit"] in ("ns", "nsecond") else v * 1e3 if r["Metric Unit"] in ("ms", "msecond") else v
        tot[k][0] += 1; tot[k][1] += v
    total = sum(v[1] for v in tot.values())
    out = os.path.join(ROOT, "profiles", f"{TAG}_launches_{wl}.csv")
    with open(out, "w") as f:
        f.write("kernel,launches,total_us,share_pct,avg_us\n")
        for k, (c, t) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
            f.write(f"\"{k}\",{c},{t:.1f},{100 * t / total:.2f},{t / c:.2f}\n")
    print(out, len(rows), "rows;", "total %.1f us" % total)
    print(open(out).read()[:900])
