mkdir -p gpurun_out
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log | head -1
B="timeout 120 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --no-checks"
run() { env $ENVV $B "$@" 2>gpurun_out/last.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV $*', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])" || tail -3 gpurun_out/last.err; }
ENVV="" run --kernel tma || exit 1
ENVV="" run --kernel tma --wave-la 5
ENVV="" run --kernel tma --wave-la 6
ENVV="" run --kernel tma --wave-la 8
