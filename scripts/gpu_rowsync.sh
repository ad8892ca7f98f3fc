timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
DLBC_WAVE_FREE=1 timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "wave or ragged" 2>&1 | tail -1
B="timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --kernel wave"
run() { env $ENVV $B "$@" 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV $*', round(d['value']), round(d['roofline']['kernel_ms'],2), d['clocks']['sm_mhz'])"; }
ENVV="DLBC_CUDA_LIB=$PWD/ab/libdlbc_prev.so" run
for la in 3 4 5; do ENVV="DLBC_WAVE_LA=$la" run; done
ENVV="DLBC_WAVE_LA=3 DLBC_WAVE_TYB=40" run
ENVV="DLBC_WAVE_LA=3 DLBC_WAVE_TYB=48" run
ENVV="DLBC_CUDA_LIB=$PWD/ab/libdlbc_prev.so" run
ENVV="" run
