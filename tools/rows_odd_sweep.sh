for spec in "float32 sum 1 2097152 129" "float32 sum 1 1048576 250" "float32 sum 1 786432 333" "float32 sum 1 524288 511" "float32 sum 1 524288 513" "float32 sum 1 262144 1001" "bfloat16 sum 1 2097152 250" "float64 sum 1 1048576 129"; do
  python tools/one_reduce.py $spec
  DACE_B200_NO_TINY_ROWS=1 python tools/one_reduce.py $spec | sed 's/$/   (lanes-per-row kernel)/'
done
