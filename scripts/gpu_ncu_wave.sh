# usage: bash scripts/gpu_ncu_wave.sh <outname> [env...] -- ncu capture of wave_kernel at 512x512x128
mkdir -p gpurun_out
OUT=$1; shift
B2="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-checks --kernel wave --shape 512,512,128 $EXTRA"
env "$@" timeout 120 $B2 > gpurun_out/plain_$OUT.log 2>&1 || { echo "plain failed"; tail -3 gpurun_out/plain_$OUT.log; exit 1; }
env "$@" timeout 600 ncu --set full --clock-control none --import-source on -k regex:wave_kernel -s 4 -c 1 -f -o gpurun_out/prof_$OUT $B2 > gpurun_out/ncu_$OUT.log 2>&1; tail -1 gpurun_out/ncu_$OUT.log
r=gpurun_out/prof_$OUT.ncu-rep
ncu -i $r --page raw --csv > gpurun_out/prof_${OUT}_raw.csv 2>/dev/null
ls -la $r
