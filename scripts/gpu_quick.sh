# usage: bash scripts/gpu_quick.sh <tag> [ncu]   -- GPU parity tests + bench (+ optional steady-state ncu capture)
tag=${1:-quick}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${tag}_pytest.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench exit $?"; tail -3 gpurun_out/${tag}_bench.err
python - <<PY
import json
d=json.load(open("gpurun_out/${tag}_bench.json"))
print({k:d[k] for k in ["value","ms_per_step","kernel_ms_per_step","parity","roofline"] if k in d})
print(d.get("e2e"))
PY
if [ "$2" = "ncu" ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'scan_kernel|fir_kernel|decode_kernel|dcd_kernel|fixup_kernel' -s 270 -c 27 -o gpurun_out/${tag}_steady python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/${tag}_ncu.log 2>&1; echo "ncu exit $?"
fi
