run() {
  echo "== lanes $1 streams '$2' lane_min_group $3"
  RPVG_B200_EM_LANES=$1 RPVG_B200_EM_STREAMS="$2" RPVG_B200_EM_LANE_MIN_GROUP=$3 timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('ms_per_step',round(d['ms_per_step'],2),'kernel_ms',round(d['roofline']['kernel_ms'],2),'checksum',d['work_per_step']['checksum'])
    elif 'rror' in l or 'capacity' in l: print(l.strip()[:300])
" | head -20
}
RPVG_B200_DEBUG=1 python bench.py --steps 1 --warmup 1 --no-cpu-baseline 2>&1 | grep "em bin" | head -8
run 7 "" 16
run 5 "" 16
run 4 "" 16
run 7 "0 0 2 4 6 8 9 10" 16
run 7 "" 8
run 8 "" 8
