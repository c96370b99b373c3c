run() {
  echo "== big_group $1 big_slice $2 slice $3"
  RPVG_B200_EM_BIG_GROUP=$1 RPVG_B200_EM_BIG_SLICE_WEIGHT=$2 RPVG_B200_EM_SLICE_WEIGHT=$3 timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('ms_per_step',round(d['ms_per_step'],2),'kernel_ms',round(d['roofline']['kernel_ms'],2),'checksum',d['work_per_step']['checksum'])
    elif 'rror' in l or 'capacity' in l: print(l.strip()[:300])
" | head -20
}
run 32 12000 12000
run 8 6000 12000
run 8 7000 12000
run 4 6000 12000
run 8 6000 14000
run 8 5000 12000
