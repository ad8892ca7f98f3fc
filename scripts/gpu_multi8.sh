mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_multi8.log 2>&1; tail -1 gpurun_out/pytest_multi8.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 50 --warmup 5 2>gpurun_out/multi8.err | tail -1 > gpurun_out/bench_n8.json
python -c "import json; d=json.load(open('gpurun_out/bench_n8.json')); print('auto weak x8', round(d['value']), d['ms_per_step'], d['config']['lattice'], d['e2e']['value'], d['clocks'])"
