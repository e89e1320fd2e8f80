# usage: bash scripts/gpu_r2a.sh <tag>  -- GPU parity tests, bench line, single-metric launch list
tag=${1:-r2a}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/${tag}_pytest.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"; tail -3 gpurun_out/${tag}_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/${tag}_bench.json"))
print({k:d[k] for k in ["value","ms_per_step","kernel_ms_per_step","parity","frames_by_type","faults","dropped"] if k in d})
print(d["roofline"]["frac"], d["roofline"]["avg_launch_ms"], d["roofline"]["in_step"])
print(d.get("e2e"))
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "ncu launches exit $?"
