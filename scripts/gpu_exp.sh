# experiment: bench lines under different environment settings
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 600 python bench.py --no-cpu-baseline --no-e2e --steps 20 --warmup 5 $EXTRA > gpurun_out/exp_$tag.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/exp_$tag.json"))
print("$tag", round(d["value"]), round(d["ms_per_step"],2), {k:round(v,2) for k,v in d["kernel_ms_per_step"].items()}, d.get("parity",{}).get("mismatching"), d["faults"])
PY
}
for sp in 1 4 8; do
run d20_spread$sp M17B_SPREAD=$sp
EXTRA="--workload c4" run c4_spread$sp M17B_SPREAD=$sp
done
