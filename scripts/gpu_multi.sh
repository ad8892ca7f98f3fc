# usage: bash scripts/gpu_multi.sh N   (run under gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | grep -E "^E  |^FAILED|passed|failed|Error|skipped" | head -20
T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N"
$T --steps 50 --warmup 5 2>gpurun_out/multi_auto.err | tail -1 > gpurun_out/bench_n$N.json; python -c "import json; d=json.load(open('gpurun_out/bench_n$N.json')); print('auto weak x$N', round(d['value']), d['ms_per_step'], d['config']['lattice'], d['e2e']['value'] if d['e2e'] else None)"
for k in wave; do
$T --steps 20 --warmup 3 --no-e2e --kernel $k 2>gpurun_out/multi_$k.err | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$k weak x$N', round(d['value']), d['ms_per_step'], d['config']['lattice'])"
done
$T --steps 50 --warmup 5 --no-e2e --scaling strong 2>gpurun_out/multi_strong.err | tail -1 > gpurun_out/bench_strong_n$N.json; python -c "import json; d=json.load(open('gpurun_out/bench_strong_n$N.json')); print('auto strong x$N', round(d['value']), d['ms_per_step'], d['config']['lattice'])"
