for m in haplotype-transcripts haplotypes strains transcripts; do
  RPVG_BENCH_MODEL=$m timeout 600 python tools/nested_bench.py 4000 16 2>&1 | tail -5
done
