# one round of scripts/gpu_ab.sh (every ab/*.so vs the in-tree build), spinodal wave
mkdir -p gpurun_out
B="timeout 120 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --no-checks --kernel wave"
run() { env $ENVV $B 2>gpurun_out/last.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'])" || tail -3 gpurun_out/last.err; }
{ ENVV="" run; for so in ab/*.so; do ENVV="DLBC_CUDA_LIB=$PWD/$so" run; done; ENVV="" run; for so in ab/*.so; do ENVV="DLBC_CUDA_LIB=$PWD/$so" run; done; } 2>&1 | tee gpurun_out/ab_once.txt
