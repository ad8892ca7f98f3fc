# round 2, first GPU call: does the TMA kernel run at all, parity tests, quick A/B of the three schedules
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader > gpurun_out/box.txt; nproc >> gpurun_out/box.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -5 gpurun_out/smoke.log
timeout 1700 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
grep -E "^FAILED|^ERROR|passed|failed|rc=" gpurun_out/pytest_gpu.log | head -40
B="timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --no-checks"
for k in tma wave two_pass; do
  $B --kernel $k > gpurun_out/bench_q_$k.json 2> gpurun_out/bench_q_$k.err; echo "bench $k rc=$?"
  python -c "import json,sys; d=json.loads(open('gpurun_out/bench_q_$k.json').read()); print('$k', round(d['value']), d['roofline']['kernels_ms'], round(d['roofline']['frac'],3), d['clocks'])" || tail -5 gpurun_out/bench_q_$k.err
done
