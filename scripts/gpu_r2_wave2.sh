mkdir -p gpurun_out
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 500 python -m pytest tests -m gpu -x -q --timeout 60 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" gpurun_out/pytest_gpu.log | head -20
grep -q "rc=0" gpurun_out/pytest_gpu.log || { tail -30 gpurun_out/pytest_gpu.log; exit 1; }
B="timeout 120 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --no-checks"
run() { env $ENVV $B "$@" 2>gpurun_out/last.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV $*', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])" || tail -3 gpurun_out/last.err; }
for i in 1 2; do
ENVV="" run --kernel wave
ENVV="DLBC_CUDA_LIB=$PWD/ab/wave_v1.so" run --kernel wave
done
ENVV="" run --kernel wave --wave-la 3
ENVV="" run --kernel wave --wave-la 5
ENVV="" run --kernel wave --workload porous
ENVV="DLBC_CUDA_LIB=$PWD/ab/wave_v1.so" run --kernel wave --workload porous
