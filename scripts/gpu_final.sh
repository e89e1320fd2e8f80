# usage: bash scripts/gpu_final.sh <tag>  -- tests, default bench line, c4/c2 lines, launch list
tag=${1:-r01v}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 gpurun_out/${tag}_pytest.log
timeout 900 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"; tail -2 gpurun_out/${tag}_bench.err
timeout 600 python bench.py --workload c4 --no-cpu-baseline > gpurun_out/${tag}_bench_c4.json 2> gpurun_out/${tag}_bench_c4.err; echo "c4 exit $?"
timeout 600 python bench.py --workload c2 --no-cpu-baseline > gpurun_out/${tag}_bench_c2.json 2> gpurun_out/${tag}_bench_c2.err; echo "c2 exit $?"
timeout 600 python bench.py --fma --no-cpu-baseline --no-e2e > gpurun_out/${tag}_bench_fma.json 2> gpurun_out/${tag}_bench_fma.err; echo "fma exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "ncu launches exit $?"
python - <<PY
import json
for f in ["bench","bench_c4","bench_c2","bench_fma"]:
    d=json.load(open("gpurun_out/${tag}_%s.json" % f))
    print(f, round(d["value"]), round(d["ms_per_step"],2), d.get("parity"), d["faults"], d["dropped"], (d.get("e2e") or {}).get("value"), d["roofline"]["frac"])
PY
