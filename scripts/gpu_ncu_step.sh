# usage: bash scripts/gpu_ncu_step.sh <tag> [skip] [count] -- ncu --set full of the scan / fixup / decode launches of steady steps
tag=${1:-r2}; skip=${2:-200}; count=${3:-30}
mkdir -p gpurun_out
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"${KREGEX:-scan_kernel}" -s $skip -c $count -o gpurun_out/${tag}_steady python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/${tag}_ncu.log 2>&1; echo "ncu exit $?"
ls -la gpurun_out/${tag}_steady.ncu-rep
