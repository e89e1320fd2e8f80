# usage: bash scripts/gpu_multi.sh <tag> <N> [second-workload]  -- the driver's N-GPU launch of both arms
# (configs[2] workload), then optionally our arm once more on another workload (e.g. c4 = configs[3])
tag=${1:-r01m}; N=${2:-2}; W2=${3:-}
mkdir -p gpurun_out
nvidia-smi -L | head -8; free -g | head -2; nproc
run() { timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:3}" > gpurun_out/${tag}_$2_n$N.json 2> gpurun_out/${tag}_$2_n$N.err; echo "$2 exit $?"; }
run 29511 ref --impl reference --steps 2 --warmup 1
run 29512 bench --steps 10 --warmup 3
[ -n "$W2" ] && run 29513 bench_$W2 --steps 10 --warmup 3 --workload $W2 --no-cpu-baseline
tail -3 gpurun_out/${tag}_bench_n$N.err
for f in gpurun_out/${tag}_bench*_n$N.json gpurun_out/${tag}_ref_n$N.json; do echo "== $f"; cut -c1-2500 $f; done
