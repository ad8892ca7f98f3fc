# usage: [ENV=..] bash scripts/gpu_ncu_one.sh <kernel variant> <regex> <outname>
K=$1; RX=$2; OUT=$3
B2="python bench.py --shape 512,512,128 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --kernel $K"
$B2 > gpurun_out/plain_$OUT.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"$RX" -s 4 -c 1 -f -o gpurun_out/prof_$OUT $B2 > gpurun_out/ncu_$OUT.log 2>&1; tail -1 gpurun_out/ncu_$OUT.log
