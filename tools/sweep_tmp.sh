run() {
  echo "== tiers '$1' slice $2"
  RPVG_B200_EM_TIERS="$1" RPVG_B200_EM_SLICE_WEIGHT=$2 timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('ms_per_step',round(d['ms_per_step'],2),'kernel_ms',round(d['roofline']['kernel_ms'],2),'checksum',d['work_per_step']['checksum'])
    elif 'rror' in l or 'capacity' in l: print(l.strip()[:300])
" | head -20
}
run "" 12000
run "8:64,24:128,52:128,104:256,216:512" 12000
run "8:64,24:128,52:128,104:256,216:512" 4500
run "8:64,24:128,52:128,104:256,216:512" 5500
run "8:64,24:64,52:128,104:256,216:512" 12000
run "8:64,24:128,56:256,108:256,216:512" 12000
