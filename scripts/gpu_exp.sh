# experiment: bench lines under different environment settings
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" timeout 600 python bench.py --no-cpu-baseline --no-e2e --steps 20 --warmup 5 $EXTRA > gpurun_out/exp_$tag.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/exp_$tag.json"))
print("$tag", round(d["value"]), round(d["ms_per_step"],2), {k:round(v,2) for k,v in d["kernel_ms_per_step"].items()}, d.get("parity",{}).get("mismatching"), d["faults"])
PY
}
run pad6 M17B_FIR_PAD_KB=6
run pad30 M17B_FIR_PAD_KB=30
run aux6 M17B_AUX_ROUNDS=6
run main6 M17B_MAIN_ROUNDS=6
