B="timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --kernel wave"
run() { env $ENVV $B "$@" 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV $*', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'])"; }
ENVV="DLBC_WAVE_FREE=0" run
for la in 5 6 7 8 10; do ENVV="DLBC_WAVE_FREE=1 DLBC_WAVE_LA=$la" run; done
ENVV="DLBC_WAVE_FREE=1 DLBC_WAVE_LA=8 DLBC_WAVE_TYB=16" run
ENVV="DLBC_WAVE_FREE=1 DLBC_WAVE_LA=12 DLBC_WAVE_TYB=16" run
