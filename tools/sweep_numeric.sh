#!/usr/bin/env bash
# Sweep the numeric-pass tuning knobs (env hooks read by csrc/bsp_numeric.cu) on the bench workload.
# usage: tools/sweep_numeric.sh "Q list" "panel MB list" "stages list" "warps list" [workload]
wl=${5:-cfg2}
for w in $4; do for st in $3; do for pm in $2; do for q in $1; do
  BSP_NUM_Q=$q BSP_NUM_PANEL_MB=$pm BSP_NUM_STAGES=$st BSP_NUM_WARPS=$w python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e --workload $wl 2>/dev/null |
    python -c "import json,sys; d=json.loads(sys.stdin.read()); print('Q=$q panel_mb=$pm stages=$st warps=$w', 'sym %.3f num %.3f ms  %.0f GFLOP/s  refolds %d' % (d['phases_ms']['symbolic'], d['phases_ms']['numeric'], d['value'], d['config']['exact_refolds']))"
done; done; done; done
