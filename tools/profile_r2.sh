#!/bin/bash
# Round-2 ncu evidence (run on the GPU box): launch list of the default bench command, --set full captures of the
# kernels that are new or changed this round.  Each command runs plain first (recipe: B200_PROFILING.md).
set -u
mkdir -p gpurun_out
BENCH="python bench.py --steps 5 --warmup 3 --no-cpu --e2e-steps 0 --no-extras"
$BENCH > gpurun_out/r2_plain_c2.json 2> gpurun_out/r2_plain_c2.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_c2_launches.csv $BENCH > /dev/null 2>&1
for spec in "c4:uniform:scatter_partition_kernel" "c4:zipf1.1:scatter_cache_kernel" "c4:all_equal:scatter_cache_kernel"; do
  IFS=: read w dist kern <<< "$spec"
  tag=${w}_${dist//./_}
  $BENCH --workload $w --dist $dist > gpurun_out/r2_plain_$tag.json 2> gpurun_out/r2_plain_$tag.err && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 100 --csv --log-file gpurun_out/r2_${tag}_launches.csv \
      $BENCH --workload $w --dist $dist > /dev/null 2>&1
  ncu --set full --clock-control none --import-source on -k regex:$kern -s 4 -c 1 -f -o gpurun_out/r2_prof_$tag \
      $BENCH --workload $w --dist $dist > gpurun_out/r2_ncu_$tag.log 2>&1
  ncu -i gpurun_out/r2_prof_$tag.ncu-rep --page details > gpurun_out/r2_${tag}_ncu_full_details.txt 2>&1
  ncu -i gpurun_out/r2_prof_$tag.ncu-rep --page raw --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed > gpurun_out/r2_${tag}_dram_bytes.csv 2>&1
  echo "== $tag"; head -c 400 gpurun_out/r2_plain_$tag.json; echo; tail -2 gpurun_out/r2_${tag}_dram_bytes.csv
done
python tools/inner_first_probe.py > gpurun_out/r2_inner_first_plain.txt 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 12 -c 6 --csv --log-file gpurun_out/r2_inner_first_launches.csv python tools/inner_first_probe.py > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:cols_small_kernel -s 4 -c 1 -f -o gpurun_out/r2_prof_cols_small python tools/inner_first_probe.py > gpurun_out/r2_ncu_cols_small.log 2>&1
ncu -i gpurun_out/r2_prof_cols_small.ncu-rep --page details > gpurun_out/r2_cols_small_ncu_full_details.txt 2>&1
cat gpurun_out/r2_inner_first_plain.txt; tail -7 gpurun_out/r2_inner_first_launches.csv | cut -c1-300
rm -f gpurun_out/*.ncu-rep
