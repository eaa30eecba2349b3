import sys, os, time
sys.path[:0]=["/root/repo","/root/repo/mqt-qmap_b200","/root/repo/tests"]
import numpy as np
from concurrent.futures import ProcessPoolExecutor
import multiprocessing as mp
def work(seeds):
    sys.path[:0]=["/root/repo","/root/repo/mqt-qmap_b200"]
    from oracle import refapi
    from qmap_b200 import workloads
    R=refapi.Ref(); a,c=workloads.c4_files(); ra=R.arch_load(a,c)
    D=ra.distance_table(False)
    cfg=refapi.make_config(debug=True, **workloads.C4_CONFIG)
    out=[]
    for s in seeds:
        circ=workloads.c4_circuit(s,30); cc,tt=circ.gate_arrays()
        rr=R.map_gates(ra,127,cc,tt,cfg)
        _,_,ident=circ.mapper_view(127)
        c2,t2,_=circ.mapper_view(127)
        est=0.
        for x,y in zip(c2,t2):
            if x>=0:
                est+=2.0**min(D[ident[x],ident[y]]/30.,40.)
        out.append((s,rr.expanded_nodes,est,int(rr.calls["generated_nodes"].max())))
    return out
if __name__=="__main__":
    seeds=list(range(96)); chunks=[seeds[i::16] for i in range(16)]
    with ProcessPoolExecutor(16, mp_context=mp.get_context("spawn")) as ex:
        res=[r for ch in ex.map(work,chunks) for r in ch]
    res.sort()
    e=np.array([r[1] for r in res],float); s=np.array([r[2] for r in res]); g=np.array([r[3] for r in res])
    from scipy.stats import spearmanr
    print("spearman est vs expanded", spearmanr(e,s)); print("expanded: mean",e.mean(),"max",e.max(),"p90",np.percentile(e,90)); print("max per-layer gen: max",g.max(),"p90",np.percentile(g,90))
    print("D sample", sorted(set((s*1).round(1)))[:5])
