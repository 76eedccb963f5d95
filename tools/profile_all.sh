#!/bin/bash
# One ncu --set full capture of the dominant kernel of every benchmark config + launch lists.
# Run on the GPU box AFTER the plain bench of the same workload exited 0; writes gpurun_out/prof_<w>.*
set -u
mkdir -p gpurun_out
declare -A K=( [c2]=split_reduce_kernel [c3]=rows_max_sum_kernel [c3max]=rows_group_kernel [c4]=scatter_partition_kernel
               [c5]=cols_wide_kernel [c1]=rows_group_kernel )
for w in ${WORKLOADS:-c3 c3max c4 c5 c1}; do
  python bench.py --workload $w --steps 5 --warmup 3 --no-cpu > gpurun_out/plain_$w.json 2> gpurun_out/plain_$w.err || { echo "plain $w failed"; continue; }
  ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_$w.csv \
      python bench.py --workload $w --steps 5 --warmup 3 --no-cpu > /dev/null 2>&1
  ncu --set full --clock-control none --import-source on -k regex:${K[$w]} -s 4 -c 1 -f -o gpurun_out/prof_$w \
      python bench.py --workload $w --steps 5 --warmup 3 --no-cpu > gpurun_out/ncu_$w.log 2>&1
  ncu -i gpurun_out/prof_$w.ncu-rep --page details > gpurun_out/prof_${w}_details.txt 2>&1
  ncu -i gpurun_out/prof_$w.ncu-rep --page raw --csv --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,dram__throughput.avg.pct_of_peak_sustained_elapsed > gpurun_out/prof_${w}_raw.csv 2>&1
  echo "== $w"; cat gpurun_out/plain_$w.json | head -c 600; echo; tail -2 gpurun_out/prof_${w}_raw.csv
done
