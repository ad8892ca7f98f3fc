N=2
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -2
T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N"
$T --impl reference --steps 3 --warmup 1 2>/dev/null | tail -1 | cut -c1-200
$T --steps 20 --warmup 3 2>gpurun_out/multi2.err | tail -1 > gpurun_out/bench_n2.json; python -c "import json; d=json.load(open('gpurun_out/bench_n2.json')); print('weak x2', round(d['value']), d['ms_per_step'], d['config']['lattice'], d['e2e'], d['clocks'])"
