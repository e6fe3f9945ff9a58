"""profiles/hotspots.py <tag>: per-SOURCE-LINE view of the ncu full captures of the traversal kernels (no GPU needed).
ncu's source page (`ncu -i <rep> --page source --csv`) gives warp-stall samples and executed instructions per SASS instruction;
`nvdisasm -g` on the library's cubin gives the source line of every instruction (-lineinfo build).  Joined on the instruction
offset (opcodes are cross-checked) and summed per source line -> profiles/<tag>_source_hotspots.md."""
import csv, io, os, re, subprocess, sys, tempfile, collections

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TAG = sys.argv[1] if len(sys.argv) > 1 else "r01"
LIB = os.path.join(ROOT, "openclraytracer_b200", "libclrt_cuda.so")
# (capture label, kernel, substring of the mangled name, object file of the library)
KERNELS = [("soup1m", "shadow_shade_kernel", "shadow_shade_kernelILb0", "trace"), ("soup1m", "trace_primary_kernel", "trace_primary_kernelILi1ELb0", "trace"),
           ("clumsydragon", "shadow_shade_kernel", "shadow_shade_kernelILb0", "trace"),
           ("clumsydragon", "trace_primary_kernel", "trace_primary_kernelILi1ELb0", "trace"),
           ("blocks", "onesweep_pass_kernel", "onesweep_pass_kernelILi16ELb1", "sort"), ("soup1m", "refit_atomic_kernel", "refit_atomic_kernel", "bvh"),
           ("soup1m", "karras_emit_kernel", "karras_emit_kernel", "bvh"), ("soup1m", "widen_kernel", "widen_kernel", "trace_tree")]


def disassemble(obj):
    d = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", obj + ".", LIB], cwd=d, check=True, stdout=subprocess.DEVNULL)
    cubin = [f for f in os.listdir(d) if f.startswith(obj + ".") and f.endswith(".cubin")][0]
    text = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cubin)], check=True, capture_output=True, text=True).stdout
    funcs, cur, line, inl = {}, None, None, None
    for l in text.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", l)
        if m:
            cur = funcs.setdefault(m.group(1), []); line = None
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', l)
        if m:
            line = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m and cur is not None:
            cur.append((int(m.group(1), 16), m.group(2).strip(), line))
    return funcs


def ncu_sass(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    col = {n: i for i, n in enumerate(rows[h])}
    res = []
    for r in rows[h + 1:]:
        if len(r) <= col["Avg. Threads Executed"] or not r[0].startswith("0x"):
            continue
        f = lambda k: float(r[col[k]] or 0)
        res.append((int(r[0], 16), r[col["Source"]].strip(), f("Warp Stall Sampling (All Samples)"), f("Warp Stall Sampling (Not-issued Samples)"),
                    f("Instructions Executed"), f("Thread Instructions Executed")))
    base = res[0][0]
    return [(a - base, s, x, y, n, t) for a, s, x, y, n, t in res]


def main():
    dis_cache, src_cache = {}, {}

    def source_line(name, no):
        if name not in src_cache:
            p = os.path.join(ROOT, "openclraytracer_b200", "csrc", name)
            src_cache[name] = open(p).read().splitlines() if os.path.exists(p) else []
        lines = src_cache[name]
        return lines[no - 1].strip()[:110].replace("|", "\\|") if 0 < no <= len(lines) else ""
    out = ["# Source-line hot spots of the hot kernels (%s)" % TAG, "",
           "From the `ncu --set full --import-source on` captures of profiles/capture.sh, joined with `nvdisasm -g` of the shipped cubin "
           "(profiles/hotspots.py).  samples = warp-stall samples attributed to the line's instructions (all / not issued); "
           "inst = warp-level instructions executed; lanes = thread instructions / warp instructions.", ""]
    for wl, kernel, mangled, obj in KERNELS:
        rep = os.path.join(ROOT, "gpurun_out", "full_%s_%s_%s.ncu-rep" % (wl, kernel, TAG))
        if not os.path.exists(rep):
            continue
        if obj not in dis_cache:
            dis_cache[obj] = disassemble(obj)
        funcs = dis_cache[obj]
        name = next(k for k in funcs if mangled in k)
        dis, sass = funcs[name], ncu_sass(rep)
        by_off = {o: (op, ln) for o, op, ln in dis}
        agree = sum(1 for o, s, *_ in sass if o in by_off and by_off[o][0].split()[0] == s.split()[0])
        per = collections.defaultdict(lambda: [0.0, 0.0, 0.0, 0.0])
        for o, s, x, y, n, t in sass:
            ln = by_off.get(o, (None, None))[1] or ("?", 0)
            p = per[ln]; p[0] += x; p[1] += y; p[2] += n; p[3] += t
        tot = [sum(p[i] for p in per.values()) for i in range(4)]
        out += ["## [%s] %s — %d SASS instructions, %d opcodes agree with the cubin, %.0f samples, %.3g warp instructions, %.1f lanes on average"
                % (wl, kernel, len(sass), agree, tot[0], tot[2], tot[3] / max(tot[2], 1)), "",
                "| file:line | samples (all) | % | not issued | inst % | lanes | source |", "|---|---|---|---|---|---|---|"]
        for ln, p in sorted(per.items(), key=lambda kv: -kv[1][0])[:(22 if obj == "trace" else 12)]:
            code = source_line(ln[0], ln[1])
            out.append("| %s:%d | %.0f | %.1f | %.0f | %.1f | %.1f | `%s` |" % (ln[0], ln[1], p[0], 100 * p[0] / max(tot[0], 1), p[1], 100 * p[2] / max(tot[2], 1),
                                                                              p[3] / max(p[2], 1), code))
        out.append("")
    open(os.path.join(ROOT, "profiles", "%s_source_hotspots.md" % TAG), "w").write("\n".join(out) + "\n")
    print("\n".join(out[:60]))


if __name__ == "__main__":
    main()
