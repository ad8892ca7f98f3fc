# new TMA kernel: smoke, TMA-related parity tests (fail fast, bounded), A/B against the wave kernel (12 vs 8 consumer warps)
mkdir -p gpurun_out
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/smoke.log
timeout 400 python -m pytest tests -m gpu -x -q --timeout 60 -k "tma or multi_band or forced_bands or spinodal or golden or roundtrip or tables" > gpurun_out/pytest_tma.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_tma.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" gpurun_out/pytest_tma.log | head -20
grep -q "rc=0" gpurun_out/pytest_tma.log || { tail -30 gpurun_out/pytest_tma.log; exit 1; }
B="timeout 120 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --no-checks"
run() { env $ENVV $B "$@" 2>gpurun_out/last.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV $*', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])" || tail -3 gpurun_out/last.err; }
ENVV="" run --kernel wave
ENVV="" run --kernel tma || exit 1
ENVV="DLBC_CUDA_LIB=$PWD/ab/cw8.so" run --kernel tma
ENVV="" run --kernel wave
ENVV="" run --kernel tma
ENVV="DLBC_CUDA_LIB=$PWD/ab/cw8.so" run --kernel tma
ENVV="" run --kernel tma --wave-la 3
ENVV="" run --kernel tma --wave-la 5
ENVV="DLBC_CUDA_LIB=$PWD/ab/cw8.so" run --kernel tma --wave-la 3
ENVV="DLBC_CUDA_LIB=$PWD/ab/cw8.so" run --kernel tma --wave-la 5
ENVV="" run --kernel wave --workload porous
ENVV="" run --kernel two_pass --workload porous
