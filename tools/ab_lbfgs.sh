# the lock-step solves with alternative builds of the library (tools/build_variant.sh), one line per build
for lib in "" build/ab/*/lib.so; do
  echo "== $lib"
  PML_LIB=${lib:+$PWD/$lib} PML_FORMS=fused,lbfgs python tools/profile_re_solve.py 2>/dev/null | python -c "
import json,sys
j=json.load(sys.stdin)
for k in ('fused','lbfgs'):
    v=j[k]
    print(k, round(v['wall_s_unprofiled'],4), round(v['device_ms_total'],1), [(round(r['ms'],2),r['n']) for r in v['top'] if 'lbfgs_direction' in r['kernel'] or 'cg_step' in r['kernel']])"
done
