# experiment: bench lines of the configs the acquiring streams' spread factor matters for
mkdir -p gpurun_out
run() { tag=$1; shift; timeout 600 python bench.py --no-cpu-baseline --no-e2e "$@" > gpurun_out/exp_$tag.json 2>/dev/null
  python - <<PY
import json
d=json.load(open("gpurun_out/exp_$tag.json"))
print("$tag", round(d["value"]), round(d["ms_per_step"],2), {k:round(v,2) for k,v in d["kernel_ms_per_step"].items()}, d.get("parity",{}).get("mismatching"), d["faults"])
PY
}
run c4_128k --workload c4 --channels 131072 --steps 10 --warmup 3
run c4 --workload c4 --steps 20 --warmup 5
run d20 --steps 20 --warmup 5
