B="timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --kernel wave"
run() { env $ENVV $B "$@" 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV $*', round(d['value']), round(d['roofline']['kernel_ms'],2), d['clocks']['sm_mhz'])"; }
ENVV="" run
ENVV="DLBC_WAVE_BATCH=0" run
ENVV="DLBC_WAVE_BATCH=0 DLBC_WAVE_LA=3" run
ENVV="DLBC_WAVE_TYB=24 DLBC_WAVE_LA=5" run
ENVV="DLBC_WAVE_TYB=40 DLBC_WAVE_LA=3" run
ENVV="DLBC_WAVE_TYB=40 DLBC_WAVE_LA=4" run
ENVV="" run --workload porous
ENVV="DLBC_WAVE_LA=5" run --workload porous
ENVV="" run
bash scripts/gpu_ncu_env.sh rowsync
