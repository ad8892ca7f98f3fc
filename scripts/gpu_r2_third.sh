# TMA feed microbenchmark + band/lag sweep of the wave kernel on the porous workload
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench_tma scripts/ubench_tma.cu && timeout 300 /tmp/ubench_tma > gpurun_out/ubench_tma.txt 2>&1
tail -100 gpurun_out/ubench_tma.txt
B="timeout 300 python bench.py --workload porous --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-checks --kernel wave --shape 512,512,256"
for opt in "" "--wave-tyb 16" "--wave-tyb 16 --wave-la 5" "--wave-tyb 16 --wave-la 3" "--wave-la 3" "--wave-tyb 24 --wave-la 3" "--wave-tyb 8 --wave-la 6"; do
  $B $opt 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('porous $opt', round(d['value']), d['roofline']['kernels_ms'], d['clocks']['sm_mhz'])"
done
