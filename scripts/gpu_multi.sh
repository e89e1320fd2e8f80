# usage: bash scripts/gpu_multi.sh <tag> <N>  -- the driver's N-GPU launch of both arms
tag=${1:-r01m}; N=${2:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8; free -g | head -2; nproc
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/${tag}_ref_n$N.json 2> gpurun_out/${tag}_ref_n$N.err; echo "ref exit $?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/${tag}_bench_n$N.json 2> gpurun_out/${tag}_bench_n$N.err; echo "bench exit $?"
tail -3 gpurun_out/${tag}_bench_n$N.err
cut -c1-3000 gpurun_out/${tag}_bench_n$N.json; cut -c1-600 gpurun_out/${tag}_ref_n$N.json
