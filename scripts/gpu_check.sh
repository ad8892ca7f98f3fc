# parity (all GPU tests) + quick bench of the main variants
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | grep -E "^E  |^FAILED|passed|failed|Error" | head -20
B="timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e"
run() { env $ENVV $B "$@" 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV $*', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
ENVV="" run --kernel wave
ENVV="" run --kernel two_pass
for la in 4 5 6; do for b in 0 1; do ENVV="DLBC_WAVE_BATCH=$b DLBC_WAVE_LA=$la" run --kernel wave; done; done
ENVV="DLBC_WAVE_BATCH=1 DLBC_WAVE_LA=5 DLBC_WAVE_TYB=40" run --kernel wave
ENVV="DLBC_WAVE_BATCH=1 DLBC_WAVE_LA=6 DLBC_WAVE_TYB=24" run --kernel wave
ENVV="DLBC_WAVE_BATCH=1 DLBC_WAVE_LA=5" run --kernel wave
