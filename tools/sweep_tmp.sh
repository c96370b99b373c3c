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
timeout 300 python -m pytest tests/test_em_gpu.py -x -q 2>&1 | tail -15
run tools/bin/librpvg_b200_head.so
run rpvg_b200/librpvg_b200.so
for lib in rpvg_b200/librpvg_b200.so; do
echo "== single $lib"
RPVG_B200_LIB=$PWD/$lib timeout 300 python - <<'PY'
import os, sys, time
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from rpvg_b200 import capi, synth
arena = synth.transcripts_arena()
shapes = np.array([arena.cluster_shape(c) for c in range(arena.n_clusters)])
weight = shapes[:, 2] + shapes[:, 0]
for target in (190000, 60000, 30000, 12000, 8000, 5000):
    c = int(np.argmin(np.abs(weight - target)))
    sub = arena.select([c])
    ctx = capi.Context(0)
    b = ctx.upload(sub)
    b.em(); torch.cuda.synchronize()
    t0 = time.perf_counter(); b.em(); torch.cuda.synchronize(); t1 = time.perf_counter()
    ab = np.zeros(sub.n_paths); nz = np.zeros(1); tc = np.zeros(1); its = np.zeros(1, dtype=np.uint32)
    b.download(ab, nz, tc, its)
    print(f"  weight {weight[c]} its {its[0]}  {1e3*(t1-t0):.2f} ms  {1e6*(t1-t0)/its[0]:.2f} us/it")
PY
done
for cfg in "12000 2" "30000 4" "190000 16"; do
  set -- $cfg
  echo "== trace weight $1 group $2"
  RPVG_B200_LIB=$PWD/tools/bin/librpvg_b200_trace.so timeout 120 python tools/em_single.py $1 $2 2>&1 | grep -E "(layout|trace)\] rank 0/|^cluster" | head -4
done
timeout 300 python tools/em_timeline.py > gpurun_out/timeline_pull.txt 2>&1
