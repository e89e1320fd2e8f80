set -x
nvidia-smi -L
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01b_pytest_gpu.log 2>&1; echo "pytest exit $?"
tail -5 gpurun_out/r01b_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/r01b_bench.json 2> gpurun_out/r01b_bench.err; echo "bench exit $?"
cat gpurun_out/r01b_bench.json; tail -3 gpurun_out/r01b_bench.err
timeout 600 python bench.py --fma --no-cpu-baseline > gpurun_out/r01b_bench_fma.json 2> gpurun_out/r01b_bench_fma.err; echo "bench fma exit $?"
cat gpurun_out/r01b_bench_fma.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01b_launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/r01b_ncu_launch.log 2>&1; echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'scan_kernel|fir_kernel|decode_kernel' -s 51 -c 17 -o gpurun_out/r01b_full python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/r01b_ncu_full.log 2>&1; echo "ncu full exit $?"
ls -la gpurun_out
