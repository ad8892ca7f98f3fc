# one gpurun call: parity tests, smoke, bench (both arms), ncu launch list, full captures of the top kernels
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader > gpurun_out/box.txt; nproc >> gpurun_out/box.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
for k in two_pass zmarch; do
timeout 300 python bench.py --kernel $k --no-cpu-baseline --no-e2e --steps 20 > gpurun_out/bench_$k.json 2>> gpurun_out/bench.err
done
timeout 300 python bench.py --components 1 --size 256 --no-cpu-baseline --no-e2e --steps 200 > gpurun_out/bench_c1_256.json 2>> gpurun_out/bench.err; cat gpurun_out/bench_c1_256.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; cat gpurun_out/bench_ref.json
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launches.log 2>&1
B2="python bench.py --shape 512,512,128 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
for k in wave two_pass; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wave_kernel|fused_kernel|psi_kernel" -s 6 -c 2 -f -o gpurun_out/prof_$k $B2 --kernel $k > gpurun_out/ncu_$k.log 2>&1; tail -1 gpurun_out/ncu_$k.log
done
# gpurun brings back <= 64 MiB: export the pages read offline, keep the .ncu-rep only when small
for k in wave two_pass; do
  r=gpurun_out/prof_$k.ncu-rep
  [ -f $r ] || continue
  ncu -i $r --page raw --csv > gpurun_out/prof_${k}_raw.csv 2>/dev/null
  ncu -i $r --page source --csv > gpurun_out/prof_${k}_source.csv 2>/dev/null
  ncu -i $r --page details > gpurun_out/prof_${k}_details.txt 2>/dev/null
  gzip -f gpurun_out/prof_${k}_source.csv
  ls -la $r
  [ $(stat -c %s $r) -gt 20000000 ] && rm -f $r
done
du -sh gpurun_out
