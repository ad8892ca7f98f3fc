timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "wave" 2>&1 | grep -E "^E  |^FAILED|passed|failed|Error" | head -20
B="timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e"
for tyb in 16 32 64 128; do for la in 2 4 8; do DLBC_WAVE_TYB=$tyb DLBC_WAVE_LA=$la $B --kernel wave 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('wave tyb=$tyb la=$la', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; done; done
$B --kernel wave 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('wave default', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3))"
