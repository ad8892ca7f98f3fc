# usage: bash scripts/gpu_multi.sh N   (run under gpurun --gpus N)
N=${1:-2}
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | grep -E "^E  |^FAILED|passed|failed|Error|skipped" | head -20
for k in two_pass wave zpipe; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 --no-e2e --kernel $k 2>gpurun_out/multi_$k.err | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$k x$N', round(d['value']), d['ms_per_step'], d['config']['lattice'])"
done
