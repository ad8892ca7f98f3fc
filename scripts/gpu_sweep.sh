python -m pytest tests/test_gpu_parity.py -m gpu -q --deselect tests/test_gpu_parity.py::test_golden_merging_droplets_3d_t1000 2>&1 | grep -E "^E  |^FAILED|passed|failed" > gpurun_out/t5.log; cat gpurun_out/t5.log
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e"
for k in two_pass; do $B --kernel $k 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$k', round(d['value']), d['roofline']['kernels_ms'])"; done
for tile in 64x8 64x4 32x8; do for ch in 32 128 512; do DLBC_ZM_TILE=$tile DLBC_ZM_CHUNK=$ch $B --kernel zmarch 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('zmarch $tile $ch', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3))"; done; done
