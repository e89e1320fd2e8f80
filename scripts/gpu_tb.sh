# usage: bash scripts/gpu_tb.sh <tag> [bench args]  -- GPU parity tests + one bench line (no CPU baseline)
tag=${1:-tb}; shift
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/${tag}_pytest.log
timeout 900 python bench.py --no-cpu-baseline "$@" > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"; tail -3 gpurun_out/${tag}_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/${tag}_bench.json"))
print({k:d[k] for k in ["value","ms_per_step","kernel_ms_per_step","parity","frames_by_type","faults","dropped","search"] if k in d})
print(d["roofline"]["frac"], d["roofline"]["avg_launch_ms"], d["roofline"]["in_step"]["avg_launch_ms"])
print(d.get("e2e",{}).get("value"))
PY
