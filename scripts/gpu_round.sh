# one gpurun call: parity tests, bench (both arms), ncu launch list, one full capture of the top kernel
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
python bench.py --kernel two_pass --no-cpu-baseline --no-e2e --steps 20 > gpurun_out/bench_two_pass.json 2>> gpurun_out/bench.err
python bench.py --kernel zmarch --no-cpu-baseline --no-e2e --steps 20 > gpurun_out/bench_zmarch.json 2>> gpurun_out/bench.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; cat gpurun_out/bench_ref.json
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launches.log 2>&1
B2="python bench.py --shape 512,512,128 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
for k in zmarch two_pass; do
  ncu --set full --clock-control none --import-source on -k regex:"zmarch_kernel|fused_kernel|psi_kernel" -s 6 -c 2 -f -o gpurun_out/prof_$k $B2 --kernel $k > gpurun_out/ncu_$k.log 2>&1; tail -1 gpurun_out/ncu_$k.log
done
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader; nproc
