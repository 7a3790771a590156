# bench.py with and without the sliced-ELL margin pass / its bank-aware entry order, one line per setting
run() {
  env "$@" python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-convergence 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=j['kernels']
print('$*', round(j['value'],1), 'e2e', round(j['e2e']['value'],1), 'row', round(k['row_pass']['ms_per_launch'],3), 'epi', round(k['row_epilogue']['ms_per_launch'],3), 'tiled', round(k['col_pass_tiled']['ms_per_launch'],3), 'cold', round(k['col_pass']['ms_per_launch'],3), 'hv', round(j['hessian_vector']['cold_ms'],3), round(j['hessian_vector']['cached_ms'],3), 'parity', j.get('parity',{}).get('ok'), 'obj', j['e2e'].get('objective'), 'pack_s', j.get('pack_s'))"
}
for v in "$@"; do run $v; done
