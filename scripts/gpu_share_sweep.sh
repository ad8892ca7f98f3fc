# band height / lag sweep with the shared psi halo rows (no redundant psi work any more: small bands cost nothing extra)
mkdir -p gpurun_out
B="timeout 120 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --no-checks --kernel wave"
run() { env DLBC_CUDA_LIB=$PWD/ab/share.so $B "$@" 2>gpurun_out/last.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$*', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks']['sm_mhz'])" || tail -3 gpurun_out/last.err; }
{ run
for c in "8 16" "16 8" "16 6" "16 5" "24 6" "24 5" "24 4" "32 5" "32 3" "48 3"; do set -- $c; run --wave-tyb $1 --wave-la $2; done
run; } 2>&1 | tee gpurun_out/ab_share_sweep.txt
