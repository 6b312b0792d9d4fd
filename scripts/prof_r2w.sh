set -x
python bench.py > gpurun_out/r2w_bench_C.json 2> gpurun_out/r2w_bench_C.err; echo "bench rc=$?"
B="python bench.py --steps 3 --warmup 1 --no-cpu-baseline --no-e2e --no-extra"
$B > gpurun_out/r2w_plain_C.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2w_launches_benchC.csv $B > gpurun_out/r2w_ncu_list.log 2>&1
BC="$B --no-partition"
$BC > gpurun_out/r2w_plain_C2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'porec_tile|expand_diag|row_sums' -s 3 -c 3 -o gpurun_out/r2w_full_C $BC > gpurun_out/r2w_ncu_full_C.log 2>&1
ls -la gpurun_out/ | tail -8
