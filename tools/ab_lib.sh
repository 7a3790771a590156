# bench.py with alternative builds of the library (PML_LIB=path), one line per build
run() {
  env "$@" python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-convergence --no-parity 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=j['kernels']
print('$*', round(j['value'],1), 'row', round(k['row_pass']['ms_per_launch'],3), 'epi', round(k['row_epilogue']['ms_per_launch'],3), 'tiled', round(k['col_pass_tiled']['ms_per_launch'],3), 'cold', round(k['col_pass']['ms_per_launch'],3), 'hv', round(j['hessian_vector']['cold_ms'],3), round(j['hessian_vector']['cached_ms'],3))"
}
run PML_NOP=1
for f in build/ab/*/lib.so; do run PML_LIB=$PWD/$f; done
