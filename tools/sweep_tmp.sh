run() {
  echo "== lib $1 lanes $2 conns $3"
  CUDA_DEVICE_MAX_CONNECTIONS=$3 RPVG_B200_LIB=$PWD/$1 RPVG_B200_EM_LANES=$2 timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('ms_per_step',round(d['ms_per_step'],2),'kernel_ms',round(d['roofline']['kernel_ms'],2),'checksum',d['work_per_step']['checksum'])
    elif 'rror' in l or 'priorit' in l: print(l.strip()[:300])
"
}
python -m pytest tests/test_em_gpu.py -x -q 2>&1 | tail -2
run tools/bin/librpvg_b200_head.so 0 8
run rpvg_b200/librpvg_b200.so 0 8
run rpvg_b200/librpvg_b200.so 7 8
run rpvg_b200/librpvg_b200.so 8 8
RPVG_B200_EM_LANES=7 timeout 300 python tools/em_timeline.py > gpurun_out/timeline_l7.txt 2>&1
