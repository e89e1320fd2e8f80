# usage: bash scripts/gpu_campaign1.sh <tag>  -- single-GPU bench lines of every workload, both arms (JSON lines only)
tag=${1:-r3}
mkdir -p gpurun_out
b() { name=$1; shift; timeout 400 python bench.py "$@" > gpurun_out/${tag}_bench_${name}.json 2> gpurun_out/${tag}_bench_${name}.err; echo "bench ${name} exit $?"; }
b default
b d20 --steps 20 --warmup 5 --no-cpu-baseline
b reference --impl reference
b c2 --workload c2 --steps 20 --warmup 5 --no-cpu-baseline
b c4 --workload c4 --steps 20 --warmup 5 --no-cpu-baseline
b fma --fma --steps 30 --warmup 5 --no-cpu-baseline --no-e2e
b 128k --channels 131072 --steps 20 --warmup 5 --no-cpu-baseline --no-e2e
b unique65536 --gen device --unique 65536 --parity-waveforms 2048 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/${tag}_bench_*.json")):
    try: d=json.load(open(f))
    except Exception as e: print(f, "unreadable", e); continue
    print(f.split("_bench_")[1], round(d["value"]), round(d["ms_per_step"],2), d.get("parity",{}).get("mismatching"), d.get("parity",{}).get("records_compared"), d.get("faults"), "e2e", round(d.get("e2e",{}).get("value",0)), "frac", d.get("roofline",{}).get("frac"))
PY
