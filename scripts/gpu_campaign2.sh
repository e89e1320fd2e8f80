# usage: bash scripts/gpu_campaign2.sh <tag>  -- ncu evidence of the default bench command: single-metric launch list, and
# --set full captures of the matched filter (with source), the scan (with source) and the other kernels.  Keep the sum of
# the reports below gpurun's 64 MiB.
tag=${1:-r3}
mkdir -p gpurun_out
B="python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-parity"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2400 --csv --log-file gpurun_out/${tag}_launches.csv $B > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "launch list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fir_kernel -s 8 -c 1 -o gpurun_out/${tag}_fir $B > gpurun_out/${tag}_ncu_fir.log 2>&1; echo "ncu fir exit $?"
timeout 1200 ncu --set full --clock-control none -k regex:'dcd_kernel|classify_kernel|decode_kernel|search_|fixup|tail_kernel' -s 500 -c 16 -o gpurun_out/${tag}_others $B > gpurun_out/${tag}_ncu_others.log 2>&1; echo "ncu others exit $?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:scan_kernel -s 112 -c 6 -o gpurun_out/${tag}_scan $B > gpurun_out/${tag}_ncu_scan.log 2>&1; echo "ncu scan exit $?"
ls -la gpurun_out/; du -sm gpurun_out
