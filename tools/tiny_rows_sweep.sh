# tiny dense rows: plain reduce and the fused max + sum, several lengths and types
for n in 8 16 32 33 48 64 96 100 128; do python tools/one_reduce.py float32 sum 1 $((268435456/n)) $n; done
for n in 32 64 100 128; do python tools/one_reduce.py bfloat16 sum 1 $((536870912/n)) $n; done
for n in 32 64; do python tools/one_reduce.py float64 sum 1 $((134217728/n)) $n; done
