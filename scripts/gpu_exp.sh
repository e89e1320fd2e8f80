# experiment: bench lines under different environment settings
mkdir -p gpurun_out
for pad in 0 6 20; do
  M17B_FIR_PAD_KB=$pad timeout 600 python bench.py --no-cpu-baseline --no-e2e --no-parity > gpurun_out/exp_pad$pad.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/exp_pad$pad.json"))
print("pad $pad", round(d["ms_per_step"],2), {k:round(v,2) for k,v in d["kernel_ms_per_step"].items()}, round(d["roofline"]["avg_launch_ms"],2))
PY
done
