set -x
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'scan_kernel|fir_kernel|decode_kernel' -s 170 -c 17 -o gpurun_out/r01b_full_steady python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/r01b_ncu_full_steady.log 2>&1; echo "ncu full exit $?"
