run() { env "$@" timeout 300 python bench.py --workload cfg4 --freqs 148 --no-north-star --no-cpu 2>/dev/null | grep '^{' | python -c "
import json,sys; d=json.loads(sys.stdin.read()); p=d['phases_ms']; print('$*', round(d['value'],1), 'factor %.1f fwd %.1f adj %.1f'%(p['factor'],p['fwd_solve'],p['adj_solve']))"; }
run JPB_X=0
run JPB_PANEL_MAX=72
run JPB_PANEL_MAX=110
run JPB_FRONT_BIG=96
run JPB_FRONT_BIG=192
run JPB_NT_MID=256
