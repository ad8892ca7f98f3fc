# experiment: L2 persisting set-aside size vs the wave kernel (evict_last on the psi items' loads)
mkdir -p gpurun_out
B="timeout 120 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --no-checks --kernel wave"
run() { env $ENVV $B "$@" 2>gpurun_out/last.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV $*', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'])" || tail -3 gpurun_out/last.err; grep -h "persisting" gpurun_out/last.err | head -1; }
{ for mb in "" 0 16 32 48 64 96 9999; do
  if [ -z "$mb" ]; then ENVV="" run; else ENVV="DLBC_L2_PERSIST_MB=$mb" run; fi
done
ENVV="" run
ENVV="DLBC_L2_PERSIST_MB=32" run --workload porous
ENVV="" run --workload porous; } 2>&1 | tee gpurun_out/ab_l2_persist.txt
