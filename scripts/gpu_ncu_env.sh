# usage: bash scripts/gpu_ncu_env.sh <tag> [ENV=val ...]   -- full ncu capture of one wave launch under the given env
TAG=$1; shift
B2="python bench.py --shape 512,512,128 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --kernel wave"
env "$@" timeout 600 ncu --set full --clock-control none --import-source on -k regex:"wave_kernel" -s 6 -c 1 -f -o /tmp/prof_$TAG $B2 > gpurun_out/ncu_$TAG.log 2>&1
ncu -i /tmp/prof_$TAG.ncu-rep --page raw --csv > gpurun_out/prof_${TAG}_raw.csv 2>/dev/null
ncu -i /tmp/prof_$TAG.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/prof_${TAG}_source.csv.gz
