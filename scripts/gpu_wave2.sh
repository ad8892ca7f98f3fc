mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "wave" 2>&1 | grep -E "^E  |^FAILED|passed|failed|Error" | head -20
B="timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --kernel wave"
run() { env "$@" $B 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$*', round(d['value']), round(d['roofline']['kernel_ms'],3), round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
run DLBC_WAVE_SPLIT=0
run DLBC_WAVE_SPLIT=1
run DLBC_WAVE_SPLIT=1 DLBC_WAVE_LA=3
run DLBC_WAVE_SPLIT=1 DLBC_WAVE_LA=5
run DLBC_WAVE_SPLIT=1 DLBC_WAVE_LA=6
run DLBC_WAVE_SPLIT=1 DLBC_WAVE_TYB=16 DLBC_WAVE_LA=5
run DLBC_WAVE_SPLIT=1 DLBC_WAVE_TYB=64 DLBC_WAVE_LA=3
run DLBC_WAVE_SPLIT=1 DLBC_WAVE_TYB=64 DLBC_WAVE_LA=2
