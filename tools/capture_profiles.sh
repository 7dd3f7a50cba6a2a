# Launch lists and `ncu --set full` reports of the bench workloads (run on a B200 under gpurun; TAG names the files).
# Each ncu pass follows a plain run of the same command that exited 0.
TAG=${TAG:-r02z}
B="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e"
for w in c2 c1 c3 n1; do
  $B --workload $w > gpurun_out/${TAG}_bench_$w.log 2>&1 || { tail -5 gpurun_out/${TAG}_bench_$w.log; continue; }
  ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${TAG}_launches_$w.csv $B --workload $w > /dev/null 2>&1
done
ncu --set full --clock-control none --import-source on -k regex:'pb_decode_dense_kernel|zone_count_kernel|zone_scan_mc_kernel|zone_emit_kernel' -s 15 -c 4 -o gpurun_out/${TAG}_c2 $B --workload c2 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:'json_template_kernel|json_decode_kernel' -s 9 -c 3 -o gpurun_out/${TAG}_c1 $B --workload c1 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:'pb_decode_dense_kernel|merge_fixup_kernel' -s 12 -c 2 -o gpurun_out/${TAG}_c3 $B --workload c3 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:'enriched_fast_kernel|enriched_decode_kernel' -s 9 -c 3 -o gpurun_out/${TAG}_n1 $B --workload n1 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:'encode_size_kernel|encode_emit_kernel' -s 8 -c 2 -o gpurun_out/${TAG}_enc $B --workload c2 > /dev/null 2>&1
ls -la gpurun_out | tail -20
