B2="python bench.py --shape 512,512,64 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --kernel zmarch"
$B2 > gpurun_out/plainz.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"zmarch_kernel" -s 4 -c 1 -o gpurun_out/prof_r1_zmarch_b $B2 > gpurun_out/ncu3.log 2>&1; tail -2 gpurun_out/ncu3.log; tail -c 600 gpurun_out/plainz.log
