python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "zpipe" 2>&1 | grep -E "^E  |^FAILED|passed|failed|Error" | head -20
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e"
for tile in 32x8 64x4; do for ch in 64 128 512; do DLBC_ZP_TILE=$tile DLBC_ZP_CHUNK=$ch $B --kernel zpipe 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('zpipe $tile $ch', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks'])"; done; done
