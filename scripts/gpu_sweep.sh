mkdir -p gpurun_out
timeout 100 python __graft_entry__.py smoke 2>&1 | tail -3
B="timeout 120 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --no-checks --kernel wave"
run() { $B "$@" 2>gpurun_out/last.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$*', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'])" || tail -3 gpurun_out/last.err; }
run
run --wave-pf 148
run --wave-pf 296
run --wave-pf 592
run --wave-pf 1184
run --wave-pf 592 --wave-la 3
run --wave-pf 296 --wave-la 3
run
