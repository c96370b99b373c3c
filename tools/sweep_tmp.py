import sys; sys.path.insert(0,'.')
import numpy as np
from rpvg_b200 import capi, synth
from oracle import loader
for n,r in ((50_000,1_000_000),(200_000,4_000_000)):
    a=synth.transcripts_arena(seed=synth.BASE_SEED+7, n_reads=r, sizes=[n], reads_cap=1e12)
    ctx=capi.Context(0)
    ab,noise,total,its=ctx.em_abundances(a, capi.default_em_params(max_em_its=200))
    o=loader.port_em_abundances(a, max_em_its=200, num_threads=0)
    print(n, "its",its,o[3],"sum",ab.sum(),o[0].sum(),"noise",noise,o[1],"total",total,o[2], "nonzero", (ab>0).sum(), (o[0]>0).sum())
    nz=o[0]!=0
    if nz.any(): print(" max rel", np.max(np.abs(ab[nz]-o[0][nz])/o[0][nz]))
    ctx.close()
