#!/bin/bash
# diagnostic bench variants: no step overlap, twice the channels per GPU
tag=$1
mkdir -p gpurun_out
python bench.py --no-overlap --no-e2e --no-cpu-baseline > gpurun_out/${tag}_bench_noverlap.json 2> gpurun_out/${tag}_noverlap.err; echo "no-overlap exit $?"
python bench.py --channels 131072 --no-e2e --no-cpu-baseline > gpurun_out/${tag}_bench_128k.json 2> gpurun_out/${tag}_128k.err; echo "128k exit $?"
python bench.py --channels 32768 --no-e2e --no-cpu-baseline > gpurun_out/${tag}_bench_32k.json 2> gpurun_out/${tag}_32k.err; echo "32k exit $?"
python - <<P
import json
for t in ["noverlap","128k","32k"]:
    try:
        j=json.loads(open("gpurun_out/${tag}_bench_%s.json"%t).read().strip().splitlines()[-1])
        print(t, round(j["value"]), round(j["ms_per_step"],2), j["parity"], j["faults"], {k:round(v,2) for k,v in j["kernel_ms_per_step"].items()}, round(j["roofline"]["frac"],4))
    except Exception as e: print(t, "ERR", e)
P
