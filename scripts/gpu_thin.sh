N=2
T="timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N"
for nz in 64 128 256; do for k in wave two_pass; do
$T --steps 40 --warmup 5 --no-e2e --kernel $k --shape 512,512,$nz 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$k nz=$nz x2', round(d['value']), round(d['ms_per_step'],3))"
done; done
