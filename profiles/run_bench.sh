timeout 700 python -m pytest tests -m gpu -q -x 2>&1 | tail -6
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-kernels --workload clumsydragon 2>&1 | tail -1 > gpurun_out/b_cd.json
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-kernels --workload soup1m 2>&1 | tail -1 > gpurun_out/b_soup.json
python - <<PY
import json
for f in ["b_cd","b_soup"]:
    try:
        d=json.loads(open("gpurun_out/%s.json"%f).read())
        print(f, "Mrays/s %.0f ms/step %.3f e2e %.0f build %.3f"%(d["value"], d["ms_per_step"], d["e2e"]["value"], d["build_ms"]), d["build_stage_ms"])
        for k,v in d["trace_kernels"].items(): print("   ",k, {a:round(b,3) for a,b in v.items()})
        print("   ", d.get("trace_structure"))
    except Exception as e: print(f, "ERR", e, open("gpurun_out/%s.json"%f).read()[-800:])
PY
