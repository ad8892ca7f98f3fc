# usage: N=<gpus> bash scripts/gpu_multi.sh   (run under gpurun --gpus N)
# decomposition-invariance tests on N GPUs (log kept), weak and strong scaling bench lines with the driver-visible checks
mkdir -p gpurun_out
N=${N:-2}
nvidia-smi --query-gpu=name --format=csv,noheader | sort | uniq -c > gpurun_out/box_n$N.txt
timeout 600 python -m pytest tests/test_multi_gpu.py -q --timeout 200 -rs > gpurun_out/pytest_multi_n$N.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi_n$N.log
tail -4 gpurun_out/pytest_multi_n$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531"
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n${N}_weak.json 2> gpurun_out/bench_n${N}_weak.err; echo "weak rc=$?"; cat gpurun_out/bench_n${N}_weak.json | cut -c1-1500
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --scaling strong > gpurun_out/bench_n${N}_strong512.json 2> gpurun_out/bench_n${N}_strong512.err; echo "strong rc=$?"; cat gpurun_out/bench_n${N}_strong512.json | cut -c1-600
if [ -n "$BIG" ]; then
timeout 900 $TR bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-checks --scaling strong --size $BIG > gpurun_out/bench_n${N}_strong$BIG.json 2> gpurun_out/bench_n${N}_strong$BIG.err; echo "strong $BIG rc=$?"; cat gpurun_out/bench_n${N}_strong$BIG.json | cut -c1-600
fi
