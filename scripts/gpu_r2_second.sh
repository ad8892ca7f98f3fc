# round 2, second GPU call: full parity suite on the fixed build + ncu captures (tma, wave, porous wave)
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" gpurun_out/pytest_gpu.log | head -40
cap() { # name kernel regex extra-args
  local OUT=$1 K=$2 RX=$3; shift 3
  local B2="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-checks --kernel $K $*"
  timeout 300 $B2 > gpurun_out/plain_$OUT.log 2>&1 || { echo "plain $OUT failed"; return; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$RX" -s 4 -c 1 -f -o gpurun_out/prof_$OUT $B2 > gpurun_out/ncu_$OUT.log 2>&1; tail -1 gpurun_out/ncu_$OUT.log
  local r=gpurun_out/prof_$OUT.ncu-rep
  [ -f $r ] || return
  ncu -i $r --page raw --csv > gpurun_out/prof_${OUT}_raw.csv 2>/dev/null
  ncu -i $r --page source --csv > gpurun_out/prof_${OUT}_source.csv 2>/dev/null
  ncu -i $r --page details > gpurun_out/prof_${OUT}_details.txt 2>/dev/null
  gzip -f gpurun_out/prof_${OUT}_source.csv
  ls -la $r
  [ $(stat -c %s $r) -gt 15000000 ] && rm -f $r
}
cap tma tma tma_kernel --shape 512,512,64
cap wave wave wave_kernel --shape 512,512,128
cap porous wave wave_kernel --shape 512,512,128 --workload porous
du -sh gpurun_out
