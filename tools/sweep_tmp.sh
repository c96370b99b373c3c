run() {
  echo "== lib $1"
  RPVG_B200_LIB=$PWD/$1 timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('ms_per_step',round(d['ms_per_step'],2),'kernel_ms',round(d['roofline']['kernel_ms'],2),'checksum',d['work_per_step']['checksum'])
    elif 'rror' in l or 'capacity' in l: print(l.strip()[:300])
" | head -20
}
timeout 600 python -m pytest tests/test_em_gpu.py -x -q 2>&1 | tail -5
run tools/bin/librpvg_b200_prev.so
run rpvg_b200/librpvg_b200.so
run tools/bin/librpvg_b200_prev.so
run rpvg_b200/librpvg_b200.so
