#!/bin/bash
# Sweep the hybrid knob of the partitioned histogram (1 tile in N through RED.global) on C4.
for d in 0 2 3 4 5 6 8; do
  DACE_B200_PART_DIVERT=$d python bench.py --workload c4 --steps 20 --no-cpu --e2e-steps 0 2>/dev/null | python -c "
import json,sys; l=json.loads(sys.stdin.readline()); print('divert=$d', 'kernel_ms=%.4f' % l['roofline']['kernel_ms'], 'GB/s=%.0f' % l['roofline']['achieved'])"
done
