# one gpurun call: parity tests, smoke, bench (both arms), other configs, ncu launch list, full captures of the top kernels
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader > gpurun_out/box.txt; nproc >> gpurun_out/box.txt
timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -x -q --timeout 180 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
grep -q "rc=0" gpurun_out/pytest_gpu.log || { tail -40 gpurun_out/pytest_gpu.log; exit 1; }
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
timeout 300 python bench.py --kernel two_pass --no-cpu-baseline --no-e2e --no-checks --steps 20 > gpurun_out/bench_two_pass.json 2>> gpurun_out/bench.err
timeout 300 python bench.py --kernel tma --wave-la 6 --no-cpu-baseline --no-e2e --no-checks --steps 20 > gpurun_out/bench_tma.json 2>> gpurun_out/bench.err
timeout 300 python bench.py --components 1 --size 256 --no-cpu-baseline --no-e2e --steps 200 > gpurun_out/bench_c1_256.json 2>> gpurun_out/bench.err; cat gpurun_out/bench_c1_256.json
timeout 300 python bench.py --workload porous --no-cpu-baseline --steps 20 > gpurun_out/bench_porous.json 2>> gpurun_out/bench.err; cat gpurun_out/bench_porous.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; cat gpurun_out/bench_ref.json
[ -n "$NO_NCU" ] && exit 0
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-checks"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launches.log 2>&1
cap() { # name regex bench-args...
  local OUT=$1 RX=$2; shift 2
  local B2="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-checks $*"
  timeout 300 $B2 > gpurun_out/plain_$OUT.log 2>&1 || { echo "plain $OUT failed"; return; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$RX" -s 4 -c 1 -f -o gpurun_out/prof_$OUT $B2 > gpurun_out/ncu_$OUT.log 2>&1; tail -1 gpurun_out/ncu_$OUT.log
  local r=gpurun_out/prof_$OUT.ncu-rep
  [ -f $r ] || return
  ncu -i $r --page raw --csv > gpurun_out/prof_${OUT}_raw.csv 2>/dev/null
  ncu -i $r --page source --csv > gpurun_out/prof_${OUT}_source.csv 2>/dev/null
  ncu -i $r --page details > gpurun_out/prof_${OUT}_details.txt 2>/dev/null
  gzip -f gpurun_out/prof_${OUT}_source.csv
  rm -f $r
}
cap wave wave_kernel --kernel wave --shape 512,512,128
cap porous wave_kernel --kernel wave --shape 512,512,128 --workload porous
cap c1 fused_kernel --components 1 --size 256
cap two_pass_psi psi_kernel --kernel two_pass --shape 512,512,128
cap two_pass_fused fused_kernel --kernel two_pass --shape 512,512,128
du -sh gpurun_out
