# A/B of library builds on the same box: every ab/*.so vs the in-tree build
mkdir -p gpurun_out
B="timeout 120 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --no-checks"
run() { env $ENVV $B "$@" 2>gpurun_out/last.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV $*', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])" || tail -3 gpurun_out/last.err; }
for i in 1 2; do
ENVV="" run --kernel ${K:-wave} $EXTRA
for so in ab/*.so; do ENVV="DLBC_CUDA_LIB=$PWD/$so" run --kernel ${K:-wave} $EXTRA; done
done
