# usage: N=<gpus> bash scripts/gpu_multi_tests.sh   (under gpurun --gpus N): the decomposition tests only, log kept
mkdir -p gpurun_out
N=${N:-2}
timeout 600 python -m pytest tests/test_multi_gpu.py -q --timeout 200 -rs > gpurun_out/pytest_multi_n$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi_n$N.log
tail -6 gpurun_out/pytest_multi_n$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531"
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --workload porous > gpurun_out/bench_n${N}_porous.json 2> gpurun_out/bench_n${N}_porous.err; echo "porous rc=$?"; cut -c1-700 gpurun_out/bench_n${N}_porous.json
