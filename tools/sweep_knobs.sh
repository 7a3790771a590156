# bench.py under the packer's tuning knobs (one line per setting)
run() {
  env "$@" python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-convergence 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=j['kernels']
print('$*', round(j['value'],1), 'row', round(k['row_pass']['ms_per_launch'],3), 'tiled', round(k['col_pass_tiled']['ms_per_launch'],3), 'cold', round(k['col_pass']['ms_per_launch'],3), 'fin', round(k['finalize']['ms_per_launch'],3), j['config']['csc_tiling'])"
}
run PML_TILE_MIN_RUN=2
run PML_TILE_MIN_RUN=1
run PML_TILE_MIN_RUN=1 PML_TILE_PARTIAL_MB=256
run PML_TILE_MIN_RUN=3
run PML_TILE_MAX_ROWS=16000
run PML_TILE_MAX_ROWS=28000
