# A/B of library builds on the same box: every ab/*.so vs the in-tree build
B="timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e"
run() { env $ENVV $B "$@" 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV $*', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for i in 1 2; do
ENVV="" run --kernel ${K:-wave}
for so in ab/*.so; do ENVV="DLBC_CUDA_LIB=$PWD/$so $EXTRA" run --kernel ${K:-wave}; done
done
