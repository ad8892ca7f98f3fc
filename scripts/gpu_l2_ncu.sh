# diagnostic: DRAM bytes / L2 hit rate of the wave kernel vs the persisting-L2 set-aside (few metrics, one launch each)
mkdir -p gpurun_out
for mb in 0 16 32 48; do
  DLBC_L2_PERSIST_MB=$mb timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_sectors_srcunit_tex_op_read.sum --clock-control none -k regex:wave_kernel -s 4 -c 1 --csv --log-file gpurun_out/l2_persist_$mb.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-checks --kernel wave --shape 512,512,128 > gpurun_out/l2_persist_$mb.log 2>&1
  echo "== set-aside request $mb MB"; grep -h "persisting" gpurun_out/l2_persist_$mb.log | head -1
  python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/l2_persist_$mb.csv")) if len(r)>5]
for r in rows[1:]: print(r[-3], r[-2], r[-1])
PY
done 2>&1 | tee gpurun_out/l2_persist_ncu.txt
