# A/B of two library builds on the same box: ab/libdlbc_head.so (previous commit) vs the in-tree build
B="timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e"
run() { env $ENVV $B "$@" 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV $*', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for i in 1 2; do
ENVV="DLBC_CUDA_LIB=$PWD/ab/libdlbc_head.so" run --kernel two_pass
ENVV="" run --kernel two_pass
ENVV="DLBC_CUDA_LIB=$PWD/ab/libdlbc_head.so" run --kernel wave
ENVV="" run --kernel wave
done
