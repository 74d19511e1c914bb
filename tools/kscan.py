import sys, numpy as np
sys.path.insert(0, ".")
import libff_b200 as m
from libff_b200 import synthetic as syn
gid=int(sys.argv[1]); lg=int(sys.argv[2]); n=1<<lg
m.init(0)
bases,_=syn.random_bases(gid,n,seed=1); sc=syn.random_scalars(gid,n,seed=2)
rb=m.ResidentBases(gid,bases); rs=m.ResidentScalars(gid,sc)
for k in (0,4,5,6,7,8,9):
    m.set_option("reduce_chunk",k)
    best=None
    for it in range(3):
        m.multi_exp_resident(rb,rs); t=m.last_timings()
        if best is None or t["total"]<best["total"]: best=t
    print("gid",gid,"lg",lg,"K=2^%d"%k, "bucket_reduce=%.2f final=%.2f total=%.2f"%(best["bucket_reduce"],best["final"],best["total"]),flush=True)
