# wave-kernel parameter sweep: L2 prefetch distance / policy, batched psi gathers
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "wave" 2>&1 | grep -E "^E  |^FAILED|passed|failed|Error" | head -20
B="timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --kernel wave"
run() { env "$@" $B 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$*', round(d['value']), round(d['roofline']['kernel_ms'],3), round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
for i in 1 2; do
run DLBC_WAVE_PF=0 DLBC_WAVE_BATCH=0
run DLBC_WAVE_PF=0 DLBC_WAVE_BATCH=1
run DLBC_WAVE_PF=592 DLBC_WAVE_BATCH=0
run DLBC_WAVE_PF=592 DLBC_WAVE_BATCH=1
done
nvidia-smi -q -d POWER | head -30
