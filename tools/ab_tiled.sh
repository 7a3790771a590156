# bench.py under the switches of the tiled CSC pass (one line per setting)
run() {
  env "$@" python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-convergence --no-parity 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=j['kernels']
print('$*', round(j['value'],1), 'row', round(k['row_pass']['ms_per_launch'],3), 'epi', round(k['row_epilogue']['ms_per_launch'],3), 'tiled', round(k['col_pass_tiled']['ms_per_launch'],3), 'cold', round(k['col_pass']['ms_per_launch'],3), 'fix', round(k['fixup']['ms_per_launch'],3), k['fixup']['launches'], 'fin', round(k['finalize']['ms_per_launch'],3), k['finalize']['launches'], 'pack_s', j['config']['pack_s'], 'hv', round(j['hessian_vector']['cold_ms'],3), round(j['hessian_vector']['cached_ms'],3))"
}
run PML_NOP=1
run PML_NO_UNIFORM_FAST=1
run PML_NO_BANK_ORDER=1
run PML_TILE_MAX_ROWS=28000
run PML_TILE_MIN_RUN=1 PML_TILE_PARTIAL_MB=256
