B="timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --kernel wave"
run() { env $ENVV $B "$@" 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$ENVV $*', round(d['value']), round(d['roofline']['kernel_ms'],3), round(d['roofline']['frac'],3), d['clocks']['sm_mhz'], d['clocks']['reasons'])"; }
export DLBC_WAVE_SPLIT=0
ENVV="" run --shape 512,512,512
ENVV="" run --shape 384,384,512
ENVV="" run --shape 640,640,256
ENVV="" run --shape 640,384,512
ENVV="" run --shape 512,384,512
ENVV="" run --shape 384,512,512
ENVV="DLBC_WAVE_BATCH=1 DLBC_WAVE_LA=3" run
ENVV="DLBC_WAVE_BATCH=1 DLBC_WAVE_LA=5" run
ENVV="DLBC_WAVE_BATCH=1 DLBC_WAVE_LA=6" run
ENVV="DLBC_WAVE_BATCH=0 DLBC_WAVE_LA=5" run
ENVV="DLBC_WAVE_BATCH=0 DLBC_WAVE_LA=6" run
ENVV="DLBC_WAVE_BATCH=1 DLBC_WAVE_TYB=16 DLBC_WAVE_LA=6" run
ENVV="DLBC_WAVE_BATCH=1 DLBC_WAVE_TYB=24 DLBC_WAVE_LA=4" run
