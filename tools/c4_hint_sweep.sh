#!/bin/bash
# L2-residency hints x chunk size for the partitioned histogram (uniform C4 keys); see wcr_partition.cu.
for lg in 24 23 22; do
  for hint in 0 1 3 7 5; do
    echo "chunk_log2=$lg hint=$hint: $(DACE_B200_PART_CHUNK_LOG2=$lg DACE_B200_PART_HINT=$hint python tools/c4_probe.py uniform 2>&1 | grep selected)"
  done
done
