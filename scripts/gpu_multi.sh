# usage: bash scripts/gpu_multi.sh <tag> <n_gpus> <name> [bench args]  -- one bench line on N GPUs of this box (torchrun, one rank per GPU)
tag=$1; n=$2; name=$3; shift 3
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n "$@" \
  > gpurun_out/${tag}_bench_${name}_n${n}.json 2> gpurun_out/${tag}_bench_${name}_n${n}.err; echo "bench ${name} n=$n exit $?"
tail -c 600 gpurun_out/${tag}_bench_${name}_n${n}.err
python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/${tag}_bench_${name}_n${n}.json") if l.startswith("{")][-1])
    print(round(d["value"]), round(d["ms_per_step"],2), d.get("parity"), d.get("faults"), d.get("dropped"), "e2e", d.get("e2e",{}).get("value"))
except Exception as e: print("no line", e)
PY
