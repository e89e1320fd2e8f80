# usage: bash scripts/gpu_final.sh <tag> -- GPU tests, smoke, the bench lines quoted in DESIGN, matched-filter ncu capture
tag=${1:-r4}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/${tag}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
b() { name=$1; shift; timeout 400 python bench.py "$@" > gpurun_out/${tag}_bench_${name}.json 2> gpurun_out/${tag}_bench_${name}.err; echo "bench ${name} exit $?"; }
b default
b d20 --steps 20 --warmup 5 --no-cpu-baseline
b c4 --workload c4 --steps 20 --warmup 5 --no-cpu-baseline
b c2 --workload c2 --steps 20 --warmup 5 --no-cpu-baseline
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/${tag}_bench_*.json")):
    try: d=json.load(open(f))
    except Exception as e: print(f, "unreadable", e); continue
    print(f.split("_bench_")[1], round(d["value"]), round(d["ms_per_step"],2), d.get("parity",{}).get("mismatching_records"), d.get("parity",{}).get("records_compared"), d.get("faults"), "e2e", round(d.get("e2e",{}).get("value",0)), "frac", d.get("roofline",{}).get("frac"), "traffic", d.get("roofline",{}).get("traffic"), d["clocks"])
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fir_kernel -s 8 -c 1 -o gpurun_out/${tag}_fir python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/${tag}_ncu_fir.log 2>&1; echo "ncu fir exit $?"
