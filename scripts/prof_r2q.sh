set -x
B="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-extra"
$B > gpurun_out/r2q_plain_C.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'kp_weak|kp_mark|sweep_kernel|agg_dense|epi_emit|epi_count|internal_kernel' -c 14 -o gpurun_out/r2q_full_tail $B > gpurun_out/r2q_ncu_full_tail.log 2>&1
BE="python bench.py --workload E --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-extra --no-partition"
$BE > gpurun_out/r2q_plain_E.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'sw_band|sw_prep' -s 5 -c 5 -o gpurun_out/r2q_full_E $BE > gpurun_out/r2q_ncu_full_E.log 2>&1
ls -la gpurun_out/ | tail
