B="timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --kernel wave"
run() { env $ENVV $B "$@" 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV $*', round(d['value']), round(d['roofline']['kernel_ms'],2), d['clocks']['sm_mhz'])"; }
for tyb in 24 32 48 64; do for la in 3 4 5 6 7; do ENVV="DLBC_WAVE_TYB=$tyb DLBC_WAVE_LA=$la" run; done; done
