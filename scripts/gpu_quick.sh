# usage: bash scripts/gpu_quick.sh [pytest -k expr]
python -m pytest tests/test_gpu_parity.py -m gpu -q -x --deselect tests/test_gpu_parity.py::test_golden_merging_droplets_3d_t1000 2>&1 | grep -E "^E  |^FAILED|passed|failed|Error" | head -20
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e"
$B --kernel two_pass 2>&1 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('two_pass', round(d['value']), d['roofline']['kernels_ms'])"
for ch in 64 128; do DLBC_ZM_CHUNK=$ch $B --kernel zmarch 2>&1 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('zmarch $ch', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3))"; done
