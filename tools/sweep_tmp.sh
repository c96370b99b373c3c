timeout 600 python -m pytest tests/test_batch_gpu.py -x -q 2>&1 | tail -2
for m in haplotype-transcripts haplotypes strains transcripts; do
  RPVG_B200_BATCH_TRACE=1 RPVG_BENCH_MODEL=$m timeout 600 python tools/nested_bench.py 4000 16 2>&1 | grep -v "^\[rpvg_b200 batch\]" | tail -5
done
RPVG_B200_BATCH_TRACE=1 timeout 600 python tools/nested_bench.py 4000 16 batch-only 2>&1 | grep "batch\]" | tail -8
