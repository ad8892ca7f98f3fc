# solid carry, second call: (1) two-pass kernels with / without the carry on the porous workload, (2) porous bench line
# and ncu capture of the MASK wave kernel of the in-tree build (wave carry = default), (3) GPU suite with ab/carry2p.so
mkdir -p gpurun_out
K=two_pass EXTRA="--workload porous" bash scripts/gpu_ab.sh 2>&1 | tee gpurun_out/ab_carry2p.txt
timeout 300 python bench.py --workload porous --no-cpu-baseline --steps 20 > gpurun_out/bench_porous.json 2> gpurun_out/bench.err; cat gpurun_out/bench_porous.json
DLBC_CUDA_LIB=$PWD/ab/carry2p.so timeout 600 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/pytest_gpu_carry2p.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_carry2p.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" gpurun_out/pytest_gpu_carry2p.log | head -20
timeout 600 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" gpurun_out/pytest_gpu.log | head -20
OUT=porous
B2="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-checks --kernel wave --shape 512,512,128 --workload porous"
timeout 300 $B2 > gpurun_out/plain_$OUT.log 2>&1 || { echo "plain $OUT failed"; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wave_kernel -s 4 -c 1 -f -o gpurun_out/prof_$OUT $B2 > gpurun_out/ncu_$OUT.log 2>&1; tail -1 gpurun_out/ncu_$OUT.log
r=gpurun_out/prof_$OUT.ncu-rep
[ -f $r ] || exit 0
ncu -i $r --page raw --csv > gpurun_out/prof_${OUT}_raw.csv 2>/dev/null
ncu -i $r --page source --csv > gpurun_out/prof_${OUT}_source.csv 2>/dev/null
ncu -i $r --page details > gpurun_out/prof_${OUT}_details.txt 2>/dev/null
gzip -f gpurun_out/prof_${OUT}_source.csv
rm -f $r
