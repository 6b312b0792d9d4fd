set -x
B="python bench.py --steps 3 --warmup 1 --no-cpu-baseline --no-e2e --no-extra"
$B > gpurun_out/r2h_plain_C.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2h_launches_benchC.csv $B > gpurun_out/r2h_ncu_list.log 2>&1
BC="$B --no-partition"
$BC > gpurun_out/r2h_plain_C2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'porec_tile|incidence_tile|expand_diag|row_sums' -s 4 -c 4 -o gpurun_out/r2h_full_C $BC > gpurun_out/r2h_ncu_full_C.log 2>&1
BD="python bench.py --workload D --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-extra --no-partition"
$BD > gpurun_out/r2h_plain_D.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'sw_band|sw_prep' -s 13 -c 13 -o gpurun_out/r2h_full_D $BD > gpurun_out/r2h_ncu_full_D.log 2>&1
ls -la gpurun_out/
