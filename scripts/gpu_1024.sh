# BASELINE configs[4]: 1024^3 two-component Shan-Chen on 8 GPUs (128-plane slabs of 1024 x 1024)
mkdir -p gpurun_out
N=8
T="timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N"
$T --shape 1024,1024,128 --steps 30 --warmup 3 2>gpurun_out/b1024.err | tail -1 > gpurun_out/bench_1024_n8.json
python -c "import json; d=json.load(open('gpurun_out/bench_1024_n8.json')); print('1024^3 x8', round(d['value']), d['ms_per_step'], d['config']['lattice'], d['e2e']['value'], d['clocks'], d['total_mass'])"
tail -3 gpurun_out/b1024.err
timeout 300 python bench.py --shape 1024,1024,64 --steps 20 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('1 GPU 1024x1024x64 wave', round(d['value']), d['roofline']['kernels_ms'])"
