N=2
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -2
T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N"
for k in wave two_pass wave two_pass; do
$T --steps 20 --warmup 3 --no-e2e --kernel $k 2>gpurun_out/multi2.err | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$k x2', round(d['value']), d['ms_per_step'], d['clocks']['sm_mhz'])"
done
