timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
B="timeout 600 python bench.py --workload porous --steps 20 --no-cpu-baseline"
run() { env $ENVV $B "$@" 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV $*', round(d['value']), d['roofline']['kernels_ms'], d['clocks']['sm_mhz'])"; }
for k in wave two_pass; do ENVV="" run --kernel $k; done
python bench.py --workload porous --no-cpu-baseline > gpurun_out/bench_porous.json 2>/dev/null
