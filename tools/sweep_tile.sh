#!/bin/bash
# experiment helper: CSC tiling geometry sweep on the bench workload
for cfg in "16000 64" "16000 128" "16000 256" "20000 128" "25000 128" "12000 128" "8000 128"; do
  set -- $cfg
  echo "max_rows=$1 partial_mb=$2"
  PML_TILE_MAX_ROWS=$1 PML_TILE_PARTIAL_MB=$2 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-convergence 2>/dev/null | python -c "import json,sys; j=json.loads(sys.stdin.read()); k=j['kernels']; print(round(j['value'],1), j['config']['csc_tiling'], 'tiled', round(k['col_pass_tiled']['ms_per_launch'],3), 'cold', round(k['col_pass']['ms_per_launch'],3), 'fixup', round(k['fixup']['ms_per_launch'],3), 'fin', round(k['finalize']['ms_per_launch'],3), 'row', round(k['row_pass']['ms_per_launch'],3))"
done
