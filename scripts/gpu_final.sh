# end-of-round refresh on one GPU: suite, smoke, bench (both arms), porous line, launch list, one full capture of the wave kernel
mkdir -p gpurun_out
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python -m pytest tests -m gpu -x -q --timeout 180 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -2 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench.json
timeout 400 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; cut -c1-300 gpurun_out/bench_ref.json
timeout 300 python bench.py --workload porous --no-cpu-baseline --steps 20 > gpurun_out/bench_porous.json 2>> gpurun_out/bench.err; cut -c1-200 gpurun_out/bench_porous.json
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-checks"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launches.log 2>&1
B2="$B --kernel wave --shape 512,512,128"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wave_kernel -s 4 -c 1 -f -o gpurun_out/prof_wave $B2 > gpurun_out/ncu_wave.log 2>&1; tail -1 gpurun_out/ncu_wave.log
r=gpurun_out/prof_wave.ncu-rep
if [ -f $r ]; then
ncu -i $r --page raw --csv > gpurun_out/prof_wave_raw.csv 2>/dev/null
ncu -i $r --page source --csv > gpurun_out/prof_wave_source.csv 2>/dev/null
ncu -i $r --page details > gpurun_out/prof_wave_details.txt 2>/dev/null
gzip -f gpurun_out/prof_wave_source.csv; rm -f $r
fi
