# usage: bash scripts/gpu_bench2.sh <tag> -- smoke, default bench (10+3), driver-like bench (20+5)
tag=${1:-b2}
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"; tail -3 gpurun_out/${tag}_bench.err
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${tag}_bench_20.json 2> gpurun_out/${tag}_bench_20.err; echo "bench20 exit $?"; tail -3 gpurun_out/${tag}_bench_20.err
python - <<PY
import json
for f in ("gpurun_out/${tag}_bench.json","gpurun_out/${tag}_bench_20.json"):
    d=json.load(open(f))
    print(f, round(d["value"]), round(d["ms_per_step"],2), {k:round(v,2) for k,v in d["kernel_ms_per_step"].items()}, d["parity"], d["faults"], d["dropped"], d["gpu_launches"])
    print("  roofline", round(d["roofline"]["frac"],3), d["roofline"]["traffic"], "e2e", d.get("e2e",{}).get("value"), d.get("cpu_baseline",{}).get("value"))
PY
