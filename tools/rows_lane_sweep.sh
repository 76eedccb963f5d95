for lb in 64 128 256; do
  echo "== DACE_B200_ROWS_LANE_BYTES=$lb"
  for spec in "float32 sum 1 1048576 256" "float32 sum 1 524288 512" "float32 sum 1 2097152 129" "float32 sum 1 262144 1000" "float64 max 1 524288 256" "bfloat16 sum 1 2097152 256" "int64 max 1 524288 256" "float32 sum 1 131072 2048"; do
    DACE_B200_ROWS_LANE_BYTES=$lb python tools/one_reduce.py $spec
  done
done
